"""Small hand-written models shared by the parity tests and the golden-vector generators: they reach the row types
and geometry pairs the humanoid does not (joint equality, dof friction, slide limits, ball joints, plane-box,
sphere-capsule, sphere-box, capsule-box, box-box)."""

SMALL_MODELS = {
  "equality_friction_limits": """
    <mujoco><option timestep="0.002"/>
      <worldbody>
        <body pos="0 0 1"><joint name="a" type="hinge" axis="0 1 0" frictionloss="0.3" damping="0.1"/><geom type="capsule" size=".05 .2" fromto="0 0 0 .4 0 0"/>
          <body pos=".4 0 0"><joint name="b" type="hinge" axis="0 1 0" limited="true" range="-20 20" stiffness="2"/><geom type="capsule" size=".04 .2" fromto="0 0 0 .3 0 0"/></body>
        </body>
        <body pos="0 1 1"><joint name="c" type="slide" axis="0 0 1" limited="true" range="-0.05 0.05"/><geom type="sphere" size=".1"/></body>
      </worldbody>
      <equality><joint joint1="a" joint2="b" polycoef="0.1 0.5 0 0 0"/></equality>
      <actuator><motor joint="a" gear="2" ctrllimited="true" ctrlrange="-1 1"/><position joint="c" kp="20" kv="2"/><velocity joint="b" kv="1.5" forcelimited="true" forcerange="-0.4 0.4"/></actuator>
    </mujoco>""",
  "box_and_spheres_on_plane": """
    <mujoco>
      <worldbody>
        <geom type="plane" size="5 5 .01" contype="3"/>
        <body pos="0 0 0.09"><freejoint/><geom type="box" size=".1 .15 .1" contype="0" conaffinity="2"/></body>
        <body pos="0.5 0 0.08"><freejoint/><geom type="sphere" size=".1"/></body>
        <body pos="0.62 0 0.12"><freejoint/><geom type="capsule" size=".05 .1"/></body>
        <body pos="0.5 0.15 0.1"><freejoint/><geom type="sphere" size=".08" condim="1"/></body>
      </worldbody>
    </mujoco>""",
  "no_constraints_no_actuators": """
    <mujoco><worldbody>
      <body pos="0 0 1"><joint type="ball"/><geom type="capsule" size=".05 .2"/>
        <body pos="0 0 -.5"><joint type="hinge" axis="1 0 0"/><geom type="sphere" size=".1"/></body>
      </body>
    </worldbody></mujoco>""",
}


BOX_PAIRS_XML = """
<mujoco>
  <worldbody>
    <geom type="plane" size="5 5 .01" pos="0 0 -1"/>
    <body pos="0 0 0"><freejoint/><geom type="box" size=".2 .3 .15" margin=".4"/></body>
    <body pos="1 0 0"><freejoint/><geom type="sphere" size=".1" margin=".4"/></body>
    <body pos="0 1 0"><freejoint/><geom type="capsule" size=".07 .25" margin=".4"/></body>
    <body pos="1 1 0"><freejoint/><geom type="capsule" size=".05 .1" margin=".4"/></body>
    <body pos="2 1 0"><freejoint/><geom type="box" size=".25 .1 .2" margin=".02"/></body>
  </worldbody>
</mujoco>"""

# every pair type of the convex (GJK / EPA) narrowphase plus the ellipsoid / cylinder primitives (collision_driver.py:43)
CONVEX_PAIRS_XML = """
<mujoco>
  <worldbody>
    <geom type="plane" size="5 5 .01" pos="0 0 -0.45"/>
    <body pos="0 0 0"><freejoint/><geom type="ellipsoid" size=".2 .3 .15"/></body>
    <body pos="1 0 0"><freejoint/><geom type="cylinder" size=".15 .25"/></body>
    <body pos="0 1 0"><freejoint/><geom type="sphere" size=".18"/><geom type="ellipsoid" size=".12 .1 .3" pos=".35 0 0" margin=".01"/></body>
    <body pos="1 1 0"><freejoint/><geom type="capsule" size=".1 .25"/><geom type="cylinder" size=".2 .1" pos="0 .4 0" euler="30 20 0"/></body>
    <body pos="2 1 0"><freejoint/><geom type="box" size=".25 .15 .2"/></body>
  </worldbody>
</mujoco>"""

# convex-hull meshes given inline (no asset files): an irregular 14-vertex hull and a tetrahedron, against each other, the plane and
# every primitive type (plane_convex + GJK / EPA with the exhaustive mesh support function)
MESH_PAIRS_XML = """
<mujoco>
  <asset>
    <mesh name="rock" vertex="0.001 0.143 -0.197 -0.147 -0.060 -0.196 0.010 0.178 -0.098 -0.146 0.092 0.101 0.026 -0.180 -0.009 0.099 -0.153 -0.078
      -0.157 -0.085 -0.183 -0.039 -0.168 0.054 0.014 -0.014 -0.275 -0.198 -0.014 0.050 -0.179 -0.045 -0.137 -0.112 0.118 -0.134
      -0.006 0.138 -0.137 -0.159 0.126 0.109 0.2 0.05 0.1 0.05 0.02 0.28"/>
    <mesh name="tetra" vertex="0 0 0.3  0.28 0 -0.1  -0.14 0.24 -0.1  -0.14 -0.24 -0.1"/>
  </asset>
  <worldbody>
    <geom type="plane" size="5 5 .01" pos="0 0 -0.45"/>
    <body pos="0 0 0"><inertial pos="0 0 0" mass="1" diaginertia=".02 .02 .02"/><freejoint/><geom type="mesh" mesh="rock"/></body>
    <body pos="1 0 0"><inertial pos="0 0 0" mass=".5" diaginertia=".01 .01 .01"/><freejoint/><geom type="mesh" mesh="tetra" margin=".01"/></body>
    <body pos="0 1 0"><freejoint/><geom type="sphere" size=".18"/><geom type="ellipsoid" size=".12 .1 .3" pos=".35 0 0"/></body>
    <body pos="1 1 0"><freejoint/><geom type="capsule" size=".1 .25"/><geom type="cylinder" size=".2 .1" pos="0 .4 0" euler="30 20 0"/></body>
    <body pos="2 1 0"><freejoint/><geom type="box" size=".25 .15 .2"/></body>
  </worldbody>
</mujoco>"""

XML = dict(SMALL_MODELS, box_pairs=BOX_PAIRS_XML, convex_pairs=CONVEX_PAIRS_XML, mesh_pairs=MESH_PAIRS_XML)
