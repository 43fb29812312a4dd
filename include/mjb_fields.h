/* mjb_fields.h -- single source of truth for the Model / Data field tables.
 *
 * The reference has no C interface: `Model` and `Data` are Python dataclasses of Warp arrays
 * (reference mujoco_warp/_src/types.py:834-1603 Model, :1703-1896 Data, :1618 Contact,
 * :1659 Constraint, :706 Option).  This header restates the subset of those fields that the
 * `step` path reads or writes as X-macro tables so that
 *   - the CUDA library (float, device pointers),
 *   - the CPU oracle (double, host pointers, test infrastructure only), and
 *   - the Python host layer (which queries the tables through mjb_*_field_name())
 * all agree on names and order without a hand-maintained copy.
 *
 * Table entry: X(name)  -- name equals the reference field name (contact / efc members are
 * prefixed `contact_` / `efc_`).  Logical shapes are documented in DESIGN.md and in
 * mujoco_warp_b200/types.py.
 */
#ifndef MJB_FIELDS_H_
#define MJB_FIELDS_H_

/* ---- model sizes (int) --------------------------------------------------------------------- */
#define MJB_MODEL_SIZES(X) \
  X(nq) X(nv) X(nu) X(na) X(nbody) X(njnt) X(ngeom) X(nsite) X(nmocap) X(neq) X(npair) \
  X(nxn)        /* statically filtered geom pairs (reference io.py:356 nxn_geom_pair_filtered) */ \
  X(nlevel)     /* kinematic tree depth levels (reference io.py:235-239 body_tree) */ \
  X(nv_pad)     /* reference io.py:650-661 */ \
  X(nlimit)     /* limited slide/hinge joints (reference io.py:264-265) */ \
  X(nlimitball) /* limited ball joints (reference io.py:266) */ \
  X(nfrictiondof) /* dofs with frictionloss > 0 */ \
  X(neqjnt)     /* joint equalities (reference io.py:447) */ \
  X(nJmom)      /* actuator moment non-zeros */ \
  X(nmaxcondim) X(nmaxpyramid) X(ngravcomp) X(is_sparse) X(nM) X(ntree) \
  X(ngeompair)  /* ngeom (ngeom - 1) / 2: entries of the upper-triangular pair table (SAP broadphase) */ \
  X(nmesh) X(nmeshvert) X(nmeshgraph) /* convex-hull vertex tables of mesh geoms (reference types.py Model.mesh_vert*) */ \
  X(neqconnect) X(neqweld) X(blob_hot_bytes) /* connect / weld equalities (reference io.py:445-448 eq_connect_adr, eq_wld_adr) */ \
  X(nhfield) X(nhfielddata) /* height fields (reference types.py:1046-1050) */ \
  X(ngroup) X(ngrouppair) /* two-level broadphase of large pair lists: geom groups (a body; each world geom alone) and group pairs */

/* ---- model int arrays ---------------------------------------------------------------------- */
#define MJB_MODEL_INT_FIELDS(X) \
  X(body_parentid) X(body_rootid) X(body_weldid) X(body_mocapid) X(eq_active0) X(body_jntnum) X(body_jntadr) \
  X(body_dofnum) X(body_dofadr) \
  X(body_subtreeend) /* one past the last body of the (pre-order contiguous) subtree */ \
  X(body_tree) X(body_tree_adr) \
  X(jnt_type) X(jnt_qposadr) X(jnt_dofadr) X(jnt_bodyid) X(jnt_limited_adr) X(jnt_limited_ball_adr) \
  X(jnt_actfrclimited) X(jnt_actgravcomp) \
  X(dof_bodyid) X(dof_jntid) X(dof_parentid) X(dof_friction_adr) \
  X(dof_tree0) /* first dof of the dof's kinematic tree: row i of qM is zero left of it (envelope of the factorisations) */ \
  X(geom_type) X(geom_bodyid) X(geom_condim) X(geom_priority) X(geom_static) \
  X(site_bodyid) \
  X(nxn_pair) X(nxn_pairid) \
  X(nxn_lookup) /* upper_tri_index(g1, g2) -> index into nxn_pair, or -1 (reference io.py:346 nxn_pairid) */ \
  X(pair_dim) \
  X(actuator_trntype) X(actuator_trnid) X(actuator_dyntype) X(actuator_gaintype) \
  X(actuator_biastype) X(actuator_actadr) X(actuator_actnum) X(actuator_ctrllimited) \
  X(actuator_forcelimited) X(actuator_actlimited) X(actuator_actearly) \
  X(eq_type) X(eq_obj1id) X(eq_obj2id) X(eq_objtype) X(eq_jnt_adr) X(eq_connect_adr) X(eq_wld_adr) \
  X(geom_dataid) X(mesh_vertadr) X(mesh_vertnum) X(mesh_graphadr) X(mesh_graph) \
  X(hfield_nrow) X(hfield_ncol) X(hfield_adr) \
  X(group_bodyid) /* body whose frame carries the group's bounding sphere */ \
  X(grouppair)    /* (group, group) of every group pair that owns filtered geom pairs; nxn_pair is sorted by group pair */ \
  X(grouppair_adr) /* ngrouppair + 1: range of nxn_pair rows of each group pair */ \
  X(grouppair_always) /* 1: never rejected by the group test (a plane, or a pair with a collision sensor, inside) */

/* ---- model real arrays --------------------------------------------------------------------- */
#define MJB_MODEL_REAL_FIELDS(X) \
  X(qpos0) X(qpos_spring) \
  X(body_pos) X(body_quat) X(body_ipos) X(body_iquat) X(body_mass) X(body_subtreemass) \
  X(body_inertia) X(body_invweight0) X(body_gravcomp) \
  X(jnt_pos) X(jnt_axis) X(jnt_stiffness) X(jnt_range) X(jnt_margin) X(jnt_solref) X(jnt_solimp) \
  X(jnt_actfrcrange) \
  X(dof_armature) X(dof_damping) X(dof_invweight0) X(dof_frictionloss) X(dof_solref) X(dof_solimp) \
  X(geom_size) X(geom_pos) X(geom_quat) X(geom_friction) X(geom_solmix) X(geom_solref) \
  X(geom_solimp) X(geom_margin) X(geom_gap) X(geom_rbound) X(geom_aabb) \
  X(site_pos) X(site_quat) \
  X(pair_solref) X(pair_solreffriction) X(pair_solimp) X(pair_margin) X(pair_gap) X(pair_friction) \
  X(actuator_dynprm) X(actuator_gainprm) X(actuator_biasprm) X(actuator_ctrlrange) \
  X(actuator_forcerange) X(actuator_actrange) X(actuator_gear) \
  X(eq_solref) X(eq_solimp) X(eq_data) \
  X(mesh_vert) \
  X(hfield_size) /* nhfield x 4: x, y, z_top, z_bottom */ \
  X(hfield_data) /* nhfielddata: elevations in [0, 1], row major, rows along +y */ \
  X(group_sphere) /* ngroup x 4: centre (body frame) and radius of a sphere around the group's geoms, margins included */

/* ---- options: re-read on every call (reference users mutate m.opt.* after put_model) -------- */
#define MJB_OPT_INT_FIELDS(X) \
  X(integrator) X(cone) X(solver) X(iterations) X(ls_iterations) X(disableflags) X(enableflags) \
  X(broadphase) X(broadphase_filter) X(run_collision_detection) X(ccd_iterations)
#define MJB_OPT_REAL_FIELDS(X) \
  X(timestep) X(tolerance) X(ls_tolerance) X(gravity_x) X(gravity_y) X(gravity_z) \
  X(impratio_invsqrt) X(meaninertia) X(ccd_tolerance)

/* ---- data sizes ---------------------------------------------------------------------------- */
#define MJB_DATA_SIZES(X) \
  X(nworld) X(naconmax) X(njmax) X(njmax_pad) X(nwconmax) /* per-world contact index list cap */

/* ---- data arrays: X(name, per_world_count); count 0 marks a global (pooled) array ---------------
 * Count expressions use the size identifiers of MJB_MODEL_SIZES / MJB_DATA_SIZES, which must be
 * in scope (as ints) where the table is expanded with a two-argument X. */
#define MJB_DATA_INT_FIELDS(X) \
  X(solver_niter, 1) X(ne, 1) X(nf, 1) X(nl, 1) X(nefc, 1) X(nacon, 0) X(ncollision, 0) \
  X(eq_active, neq) /* int32 mirror of the bool field */ \
  X(moment_rownnz, nu) X(moment_rowadr, nu) X(moment_colind, nJmom) \
  X(contact_dim, 0) X(contact_geom, 0) X(contact_efc_address, 0) X(contact_worldid, 0) \
  X(contact_type, 0) X(contact_geomcollisionid, 0) \
  X(efc_type, njmax) X(efc_id, njmax) X(efc_state, njmax_pad) \
  X(wcon_n, 1) X(wcon, nwconmax) /* per-world contact index lists (library workspace) */ \
  X(status, 1)                   /* per-world status bits (replaces device printf) */

#define MJB_DATA_REAL_FIELDS(X) \
  X(time, 1) X(energy, 2) X(qpos, nq) X(qvel, nv) X(act, na) X(qacc_warmstart, nv) X(ctrl, nu) \
  X(qfrc_applied, nv) X(xfrc_applied, 6 * nbody) X(mocap_pos, 3 * nmocap) \
  X(mocap_quat, 4 * nmocap) X(qacc, nv) X(act_dot, na) \
  X(xpos, 3 * nbody) X(xquat, 4 * nbody) X(xmat, 9 * nbody) X(xipos, 3 * nbody) \
  X(ximat, 9 * nbody) X(xanchor, 3 * njnt) X(xaxis, 3 * njnt) X(geom_xpos, 3 * ngeom) \
  X(geom_xmat, 9 * ngeom) X(site_xpos, 3 * nsite) X(site_xmat, 9 * nsite) \
  X(subtree_com, 3 * nbody) X(cdof, 6 * nv) X(cinert, 10 * nbody) \
  X(actuator_length, nu) X(actuator_moment, nJmom) X(crb, 10 * nbody) X(qM, nv_pad * nv_pad) \
  X(qLD, nv * nv) X(qLDiagInv, nv) \
  X(actuator_velocity, nu) X(cvel, 6 * nbody) X(cdof_dot, 6 * nv) X(qfrc_bias, nv) \
  X(qfrc_spring, nv) X(qfrc_damper, nv) X(qfrc_gravcomp, nv) X(qfrc_fluid, nv) \
  X(qfrc_passive, nv) X(actuator_force, nu) X(qfrc_actuator, nv) \
  X(qfrc_smooth, nv) X(qacc_smooth, nv) X(qfrc_constraint, nv) X(cacc, 6 * nbody) \
  X(cfrc_int, 6 * nbody) X(cfrc_ext, 6 * nbody) \
  X(contact_dist, 0) X(contact_pos, 0) X(contact_frame, 0) X(contact_includemargin, 0) \
  X(contact_friction, 0) X(contact_solref, 0) X(contact_solreffriction, 0) X(contact_solimp, 0) \
  X(efc_J, njmax_pad * nv_pad) X(efc_pos, njmax) X(efc_margin, njmax) X(efc_D, njmax_pad) \
  X(efc_vel, njmax) X(efc_aref, njmax) X(efc_frictionloss, njmax) X(efc_force, njmax) \
  X(efc_Ma, nv)

/* ---- pipeline stage bits (order = execution order inside one launch) ------------------------ */
#define MJB_STAGE_KINEMATICS       (1u << 0)  /* smooth.kinematics   smooth.py:358 */
#define MJB_STAGE_COM_POS          (1u << 1)  /* smooth.com_pos      smooth.py:602 */
#define MJB_STAGE_CRB              (1u << 2)  /* smooth.crb          smooth.py:889 */
#define MJB_STAGE_FACTOR_M         (1u << 3)  /* smooth.factor_m     smooth.py:1104 */
#define MJB_STAGE_COLLISION        (1u << 4)  /* collision_driver.collision :755 */
#define MJB_STAGE_MAKE_CONSTRAINT  (1u << 5)  /* constraint.make_constraint :2210 */
#define MJB_STAGE_TRANSMISSION     (1u << 6)  /* smooth.transmission smooth.py:2605 */
#define MJB_STAGE_FWD_VELOCITY     (1u << 7)  /* forward.fwd_velocity forward.py:593 */
#define MJB_STAGE_FWD_ACTUATION    (1u << 8)  /* forward.fwd_actuation forward.py:837 */
#define MJB_STAGE_FWD_ACCELERATION (1u << 9)  /* forward.fwd_acceleration(factorize=False) :950 */
#define MJB_STAGE_FWD_ACC_FACTOR   (1u << 10) /* forward.fwd_acceleration(factorize=True) */
#define MJB_STAGE_SOLVE            (1u << 11) /* solver.solve        solver.py:3297 */
#define MJB_STAGE_EULER            (1u << 12) /* forward.euler       forward.py:327 */
#define MJB_STAGE_IMPLICIT         (1u << 13) /* forward.implicit    forward.py:495 */
#define MJB_STAGE_ZERO_SENSOR      (1u << 14) /* d.sensordata.zero_(), d.energy.zero_() :978,:984 */
/* sub-functions of fwd_velocity / forward that the reference also exports on their own (mujoco_warp/__init__.py:25-112) */
#define MJB_STAGE_COM_VEL          (1u << 15) /* smooth.com_vel      smooth.py:2015 */
#define MJB_STAGE_PASSIVE          (1u << 16) /* passive.passive     passive.py:729 */
#define MJB_STAGE_RNE              (1u << 17) /* smooth.rne          smooth.py:1259 */
#define MJB_STAGE_ENERGY_POS       (1u << 18) /* sensor.energy_pos   sensor.py:2854 */
#define MJB_STAGE_ENERGY_VEL       (1u << 19) /* sensor.energy_vel   sensor.py:2922 */

#define MJB_STAGES_FWD_POSITION (MJB_STAGE_KINEMATICS | MJB_STAGE_COM_POS | MJB_STAGE_CRB | \
  MJB_STAGE_COLLISION | MJB_STAGE_MAKE_CONSTRAINT | MJB_STAGE_TRANSMISSION)
#define MJB_STAGES_FORWARD (MJB_STAGES_FWD_POSITION | MJB_STAGE_ZERO_SENSOR | MJB_STAGE_FWD_VELOCITY | \
  MJB_STAGE_FWD_ACTUATION | MJB_STAGE_FWD_ACC_FACTOR | MJB_STAGE_SOLVE)

/* ---- per-world status bits ----------------------------------------------------------------- */
#define MJB_STATUS_NEFC_OVERFLOW  (1 << 0)  /* reference forward.py:194 printf */
#define MJB_STATUS_NCON_OVERFLOW  (1 << 1)  /* reference forward.py:208 printf */
#define MJB_STATUS_WCON_OVERFLOW  (1 << 2)  /* per-world contact index list full */
#define MJB_STATUS_NAN_QACC       (1 << 3)

#endif /* MJB_FIELDS_H_ */
