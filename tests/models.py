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
  # connect (loop closure between two chains, and a body pinned to the world), weld (two free bodies; one body welded to the world
  # with a torquescale and an anchor), an inactive weld, next to a joint equality
  "connect_weld": """
    <mujoco><option timestep="0.002"/>
      <worldbody>
        <body name="l1" pos="0 0 1"><joint name="a" type="hinge" axis="0 1 0" damping="0.05"/><geom type="capsule" size=".03 .2" fromto="0 0 0 .4 0 0"/>
          <body name="l2" pos=".4 0 0"><joint name="b" type="hinge" axis="0 1 0"/><geom type="capsule" size=".03 .2" fromto="0 0 0 .3 0 -.2"/></body>
        </body>
        <body name="r1" pos=".7 0 1"><joint name="c" type="ball"/><geom type="capsule" size=".03 .1" fromto="0 0 0 0 0 -.2"/></body>
        <body name="f1" pos="0 1 1"><freejoint/><geom type="box" size=".1 .05 .08"/></body>
        <body name="f2" pos=".3 1 1.1" euler="10 20 30"><freejoint/><geom type="sphere" size=".07"/></body>
        <body name="f3" pos="0 -1 1" euler="0 45 0"><freejoint/><geom type="ellipsoid" size=".1 .06 .04"/></body>
        <body name="p1" pos="1.5 0 1"><joint name="d" type="slide" axis="0 0 1"/><joint name="e" type="hinge" axis="1 0 0"/><geom type="sphere" size=".05"/></body>
      </worldbody>
      <equality>
        <connect body1="l2" body2="r1" anchor=".3 0 -.2"/>
        <weld body1="f1" body2="f2" solref="0.01 1"/>
        <connect body1="p1" anchor="0 .1 0" solimp="0.8 0.9 0.01 0.5 2"/>
        <weld body1="f3" torquescale="0.5" anchor=".05 0 .02"/>
        <weld body1="f2" body2="f3" active="false"/>
        <joint joint1="d" joint2="e" polycoef="0 0.2 0 0 0"/>
      </equality>
      <actuator><motor joint="a" gear="3"/><motor joint="e" gear="1"/></actuator>
    </mujoco>""",
  # joint springs of every joint type (potential energy, passive forces), springref away from qpos0
  "springs": """
    <mujoco><option timestep="0.002"/>
      <worldbody>
        <body pos="0 0 1"><joint type="free" stiffness="2"/><geom type="box" size=".1 .1 .1"/></body>
        <body pos="1 0 1"><joint type="ball" stiffness="3" damping="0.1"/><geom type="capsule" size=".05 .2" fromto="0 0 0 .3 0 0"/>
          <body pos=".3 0 0"><joint type="hinge" axis="0 1 0" stiffness="5" springref="20"/><geom type="capsule" size=".04 .2" fromto="0 0 0 .3 0 0"/>
            <body pos=".3 0 0"><joint type="slide" axis="1 0 0" stiffness="40" springref="0.05"/><geom type="sphere" size=".06"/></body></body>
        </body>
      </worldbody>
    </mujoco>""",
  # gravity compensation: fully and partially compensated bodies, a joint whose compensation is routed through the actuator
  # channel (actuatorgravcomp, forward.py:806-829), a free body that floats
  "gravcomp": """
    <mujoco><option timestep="0.002"/>
      <worldbody>
        <body pos="0 0 1" gravcomp="1"><joint name="a" type="hinge" axis="0 1 0" damping="0.2"/><geom type="capsule" size=".04 .2" fromto="0 0 0 .4 0 0"/>
          <body pos=".4 0 0" gravcomp="0.5"><joint name="b" type="hinge" axis="0 1 0" actuatorgravcomp="true" actuatorfrclimited="true" actuatorfrcrange="-2 2"/>
            <geom type="capsule" size=".03 .2" fromto="0 0 0 .3 0 .1"/>
            <body pos=".3 0 .1"><joint name="c" type="ball"/><geom type="box" size=".05 .04 .03" pos=".05 0 0"/></body></body>
        </body>
        <body pos="0 1 1" gravcomp="1"><freejoint/><geom type="sphere" size=".1"/></body>
      </worldbody>
      <actuator><motor joint="b" gear="1.5"/><motor joint="a"/></actuator>
    </mujoco>""",
  # connect / weld defined between sites (the site frames, not eq_data, give the anchors and the relative orientation)
  "site_equalities": """
    <mujoco><option timestep="0.002"/>
      <worldbody>
        <site name="w0" pos=".2 0 1.2" euler="0 0 20"/>
        <body pos="0 0 1"><joint type="hinge" axis="0 1 0"/><geom type="capsule" size=".03 .2" fromto="0 0 0 .4 0 0"/><site name="a1" pos=".4 0 .05"/>
          <body pos=".4 0 0"><joint type="ball"/><geom type="capsule" size=".03 .1" fromto="0 0 0 .2 0 0"/><site name="a2" pos=".2 0 0" euler="10 0 0"/></body></body>
        <body pos=".7 0 1"><freejoint/><geom type="box" size=".05 .05 .05"/><site name="b1" pos="-.1 0 0" euler="0 15 0"/><site name="b2" pos="0 .1 0"/></body>
        <body pos="0 1 1"><freejoint/><geom type="sphere" size=".06"/><site name="c1" pos="0 0 .1" euler="5 5 5"/></body>
      </worldbody>
      <equality>
        <connect site1="a2" site2="b1"/>
        <weld site1="c1" site2="w0" torquescale="0.8"/>
        <connect site1="b2" site2="a1" active="false"/>
        <weld site1="b2" site2="c1" solref="0.03 1"/>
      </equality>
    </mujoco>""",
  "no_constraints_no_actuators": """
    <mujoco><worldbody>
      <body pos="0 0 1"><joint type="ball"/><geom type="capsule" size=".05 .2"/>
        <body pos="0 0 -.5"><joint type="hinge" axis="1 0 0"/><geom type="sphere" size=".1"/></body>
      </body>
    </worldbody></mujoco>""",
}


# Runge-Kutta with stateful actuators: filterexact + activation clamp, integrator with a clamp that binds, plain filter, actearly;
# free / ball / hinge / slide joints for the position integration (forward.py:52-112, :130-167, support.py:38-64)
SMALL_MODELS["rk4_actuators"] = """
    <mujoco><option timestep="0.004"/>
      <worldbody>
        <geom type="plane" size="5 5 .01"/>
        <body pos="0 0 1"><joint name="a" type="hinge" axis="0 1 0" damping="0.1"/><geom type="capsule" size=".05 .2" fromto="0 0 0 .4 0 0"/>
          <body pos=".4 0 0"><joint name="b" type="hinge" axis="0 1 0" stiffness="2"/><geom type="capsule" size=".04 .2" fromto="0 0 0 .3 0 0"/>
            <body pos=".3 0 0"><joint name="c" type="ball" damping="0.05"/><geom type="capsule" size=".03 .1" fromto="0 0 0 .2 0 0"/></body></body>
        </body>
        <body pos="0 1 1"><joint name="s" type="slide" axis="0 0 1" damping="1"/><geom type="sphere" size=".1"/></body>
        <body pos="1 1 0.5" euler="10 20 30"><freejoint/><geom type="box" size=".1 .08 .05"/></body>
      </worldbody>
      <actuator>
        <general joint="a" dyntype="filterexact" dynprm="0.03" gainprm="4" actlimited="true" actrange="-0.25 0.25"/>
        <general joint="b" dyntype="integrator" gainprm="2" actlimited="true" actrange="-0.004 0.004"/>
        <general joint="s" dyntype="filter" dynprm="0.02" gainprm="30" biastype="affine" biasprm="0 -30 -1"/>
        <general joint="b" dyntype="filterexact" dynprm="0.01" gainprm="1.5" actearly="true"/>
        <motor joint="a" gear="0.5"/>
      </actuator>
    </mujoco>"""

# limited ball joints (constraint.py:1422 _limit_ball): one well past its range, one inside, one with a margin that activates it
SMALL_MODELS["ball_limits"] = """
    <mujoco><option timestep="0.002"/>
      <worldbody>
        <body pos="0 0 1"><joint name="a" type="ball" limited="true" range="0 20" damping="0.05"/><geom type="capsule" size=".04 .2" fromto="0 0 0 .4 0 0"/>
          <body pos=".4 0 0"><joint name="b" type="ball" limited="true" range="0 60"/><geom type="capsule" size=".03 .15" fromto="0 0 0 .3 0 0"/>
            <body pos=".3 0 0"><joint name="c" type="hinge" axis="0 1 0" limited="true" range="-10 10"/><geom type="sphere" size=".05" pos=".1 0 0"/></body></body>
        </body>
        <body pos="0 1 1"><joint name="d" type="ball" limited="true" range="0 25" margin="0.2" solreflimit="0.01 1" solimplimit="0.8 0.9 0.01"/><geom type="box" size=".1 .05 .03" pos=".1 0 0"/></body>
      </worldbody>
      <actuator><motor joint="c" gear="1"/></actuator>
    </mujoco>"""

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

# zero-margin boxes: box-box through GJK / EPA + multi-contact face clipping, the reference's default route (collision_driver.py:74,
# :736, collision_gjk.py:1930); five free boxes of different proportions
BOX_STACK_XML = """
<mujoco>
  <worldbody>
    <geom type="plane" size="5 5 .01" pos="0 0 -1"/>
    <body pos="0 0 0"><freejoint/><geom type="box" size=".3 .3 .1"/></body>
    <body pos="1 0 0"><freejoint/><geom type="box" size=".15 .2 .1"/></body>
    <body pos="0 1 0"><freejoint/><geom type="box" size=".1 .1 .25"/></body>
    <body pos="1 1 0"><freejoint/><geom type="box" size=".2 .05 .15"/></body>
    <body pos="2 1 0"><freejoint/><geom type="box" size=".12 .12 .12"/></body>
  </worldbody>
</mujoco>"""

# a height field (inline elevations, 7 x 9 grid) under one free body of every convex type: hfield-vs-{sphere, capsule, ellipsoid,
# cylinder, box, mesh} through the per-prism GJK / EPA of collision_convex.py:158
HFIELD_XML = """
<mujoco>
  <option jacobian="dense"/>
  <asset>
    <hfield name="terrain" nrow="7" ncol="9" size="1.2 0.9 0.3 0.1"
      elevation="0 1 2 3 2 1 0 1 2   1 2 3 4 3 2 1 2 3   2 3 5 6 5 3 2 3 4   1 2 4 7 4 2 1 2 3   0 1 3 5 3 1 0 1 2   1 1 2 3 2 1 1 2 2   2 1 1 2 1 1 2 3 1"/>
    <mesh name="tetra" vertex="0 0 0.2  0.2 0 -0.08  -0.1 0.17 -0.08  -0.1 -0.17 -0.08"/>
  </asset>
  <worldbody>
    <geom type="hfield" hfield="terrain" pos="0.1 -0.05 0.02" euler="0 0 15"/>
    <body pos="0 0 1"><freejoint/><geom type="sphere" size=".12"/></body>
    <body pos="0 0 1"><freejoint/><geom type="capsule" size=".06 .15"/></body>
    <body pos="0 0 1"><freejoint/><geom type="box" size=".12 .08 .1"/></body>
    <body pos="0 0 1"><freejoint/><geom type="ellipsoid" size=".14 .09 .07"/></body>
    <body pos="0 0 1"><freejoint/><geom type="cylinder" size=".08 .1"/></body>
    <body pos="0 0 1"><inertial pos="0 0 0" mass=".5" diaginertia=".01 .01 .01"/><freejoint/><geom type="mesh" mesh="tetra"/></body>
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

XML = dict(SMALL_MODELS, box_pairs=BOX_PAIRS_XML, box_stack=BOX_STACK_XML, hfield=HFIELD_XML, convex_pairs=CONVEX_PAIRS_XML, mesh_pairs=MESH_PAIRS_XML)
