#ifndef MJB_CON_QUAD
#define MJB_CON_QUAD 1
#endif
// mjb_constraint.cuh -- constraint.make_constraint (reference constraint.py:2210), one warp per world.
//
// Row classes are emitted in the reference's launch order (equality, friction, limit, contact;
// constraint.py:2220-2779) because the solver classifies rows by index (solver.py:584-598).
// Within a class the reference's order is an atomic race; here rows are in MuJoCo-C order
// (object id, then contact index, then pyramid edge), assigned with ballots / prefix sums.
// Dense Jacobian rows (nv_pad floats) are written cooperatively: lane = dof column.
#pragma once

#include "mjb_types.cuh"

struct EfcRow {
  float D, vel, aref, pos, margin, frictionloss;
};

// constraint.py:53-121, the part of _efc_row that does not depend on the row's velocity:
// D and the spring / damper coefficients of aref = -ki * pos_aref - b * vel.
struct EfcParams {
  float D, ki, b;
};
__device__ __forceinline__ EfcParams efc_params(const KParams& K, float pos_imp, float invweight, const float* solref, const float* solimp) {
  float timeconst = solref[0];
  const float dampratio = solref[1];
  float dmin = solimp[0], dmax = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  if (!(K.o.disableflags & DSBL_REFSAFE)) timeconst = fmaxf(timeconst, 2.0f * K.o.timestep);
  dmin = clampf(dmin, MJ_MINIMP, MJ_MAXIMP);
  dmax = clampf(dmax, MJ_MINIMP, MJ_MAXIMP);
  width = fmaxf(MJ_MINVAL, width);
  mid = clampf(mid, MJ_MINIMP, MJ_MAXIMP);
  power = fmaxf(1.0f, power);
  const float dmax_sq = dmax * dmax;
  float k = 1.0f / (dmax_sq * timeconst * timeconst * dampratio * dampratio);
  float b = 2.0f / (dmax * timeconst);
  if (solref[0] <= 0.f) k = -solref[0] / dmax_sq;
  if (solref[1] <= 0.f) b = -solref[1] / dmax;
  const float imp_x = fabsf(pos_imp) / width;
  float imp_y;
  if (power == 1.0f) {  // pow(x, 0) == 1: both branches reduce to imp_x
    imp_y = imp_x;
  } else if (power == 2.0f) {
    imp_y = imp_x < mid ? imp_x * imp_x / mid : 1.0f - (1.0f - imp_x) * (1.0f - imp_x) / (1.0f - mid);
  } else if (imp_x < mid) {
    imp_y = (1.0f / powf(mid, power - 1.0f)) * powf(imp_x, power);
  } else {
    imp_y = 1.0f - (1.0f / powf(1.0f - mid, power - 1.0f)) * powf(1.0f - imp_x, power);
  }
  float imp = clampf(dmin + imp_y * (dmax - dmin), dmin, dmax);
  if (imp_x > 1.0f) imp = dmax;
  EfcParams r;
  r.D = 1.0f / fmaxf(invweight * (1.0f - imp) / imp, MJ_MINVAL);
  r.ki = k * imp;
  r.b = b;
  return r;
}

__device__ __forceinline__ EfcRow efc_row(const KParams& K, float pos_aref, float pos_imp, float invweight, const float* solref,
                                          const float* solimp, float margin, float vel, float frictionloss) {
  const EfcParams p = efc_params(K, pos_imp, invweight, solref, solimp);
  EfcRow r;
  r.D = p.D;
  r.vel = vel;
  r.aref = -p.ki * pos_aref - p.b * vel;
  r.pos = pos_aref + margin;
  r.margin = margin;
  r.frictionloss = frictionloss;
  return r;
}

__device__ __forceinline__ void store_row(const KParams& K, int w, int r, const EfcRow& e, int type, int id) {
  DW(efc_D)[r] = e.D;
  DW(efc_vel)[r] = e.vel;
  DW(efc_aref)[r] = e.aref;
  DW(efc_pos)[r] = e.pos;
  DW(efc_margin)[r] = e.margin;
  DW(efc_frictionloss)[r] = e.frictionloss;
  DW(efc_type)[r] = type;
  DW(efc_id)[r] = id;
}

// Rows with at most two non-zeros (equality joint, dof friction, joint limit): every lane of the
// warp proposes a row (`active`), rows are numbered by ballot prefix, then each row's Jacobian is
// filled cooperatively.
__device__ __forceinline__ int emit_simple_rows(const KParams& K, int w, int lane, bool active, int nefc, const EfcRow& e, int type, int id,
                                                int c0, float v0, int c1, float v1) {
  const unsigned mask = __ballot_sync(FULL, active);
  const int njmax = K.dsz.njmax, np = K.ms.nv_pad;
  const int r = nefc + __popc(mask & ((1u << lane) - 1u));
  if (active && r < njmax) store_row(K, w, r, e, type, id);
  unsigned mm = mask;
  while (mm) {
    const int src = __ffs(mm) - 1;
    mm &= mm - 1;
    const int rr = __shfl_sync(FULL, r, src);
    const int cc0 = __shfl_sync(FULL, c0, src), cc1 = __shfl_sync(FULL, c1, src);
    const float vv0 = __shfl_sync(FULL, v0, src), vv1 = __shfl_sync(FULL, v1, src);
    if (rr < njmax) {
      float* J = DW(efc_J) + (size_t)rr * np;
      for (int c = lane; c < np; c += 32) J[c] = (c == cc0) ? vv0 : ((c == cc1) ? vv1 : 0.f);
    }
  }
  return nefc + __popc(mask);
}

// constraint.py:125-364 _equality_connect and :793-1108 _equality_weld (between bodies, dense Jacobian): one equality at a
// time (they are few), lane = dof for the 3 / 6 Jacobian rows, lanes 0..5 finish the rows.  Returns the new nefc.
static __device__ __noinline__ int emit_connect_weld(const KParams& K, const unsigned char* mb, int w, int lane, int nefc, int& ne) {
  const int nv = K.ms.nv, np = K.ms.nv_pad, njmax = K.dsz.njmax;
  const float* qvel = DW(qvel);
  const float* scom = DW(subtree_com);
  const float* cdof = DW(cdof);
  for (int k = 0; k < K.ms.neqconnect + K.ms.neqweld; k++) {
    const bool weld = k >= K.ms.neqconnect;
    const int eq = weld ? MI(eq_wld_adr)[k - K.ms.neqconnect] : MI(eq_connect_adr)[k];
    if (DW(eq_active)[eq] == 0) continue;  // warp-uniform
    const int nrow = weld ? 6 : 3, r0 = nefc;
    ne += nrow;
    nefc += nrow;
    if (r0 >= njmax - nrow) continue;  // :186 / :855 (sic)
    const float* data = MR(eq_data) + 11 * eq;
    int b1 = MI(eq_obj1id)[eq], b2 = MI(eq_obj2id)[eq];
    const bool site = K.ms.nsite > 0 && MI(eq_objtype)[eq] == 6;  // mjOBJ_SITE: the site frames are constrained (:203-208, :868-877)
    const float torquescale = weld ? data[10] : 0.f;
    V3 pos1, pos2;
    Q4 quat{1.f, 0.f, 0.f, 0.f}, quat1{1.f, 0.f, 0.f, 0.f};
    if (site) {
      const int s1 = b1, s2 = b2;
      pos1 = ld3(DW(site_xpos) + 3 * s1);
      pos2 = ld3(DW(site_xpos) + 3 * s2);
      b1 = MI(site_bodyid)[s1];
      b2 = MI(site_bodyid)[s2];
      if (weld) {
        quat = mul_quat(ld4(DW(xquat) + 4 * b1), ld4(MR(site_quat) + 4 * s1));
        const Q4 q2 = mul_quat(ld4(DW(xquat) + 4 * b2), ld4(MR(site_quat) + 4 * s2));
        quat1 = Q4{q2.w, -q2.x, -q2.y, -q2.z};
      }
    } else {
      // connect: anchor1 = data[0:3] in body1, data[3:6] in body2.  weld: data[3:6] in body1, data[0:3] in body2.
      pos1 = ld3(DW(xpos) + 3 * b1) + mat_vec(ldm3(DW(xmat) + 9 * b1), ld3(data + (weld ? 3 : 0)));
      pos2 = ld3(DW(xpos) + 3 * b2) + mat_vec(ldm3(DW(xmat) + 9 * b2), ld3(data + (weld ? 0 : 3)));
      if (weld) {
        quat = mul_quat(ld4(DW(xquat) + 4 * b1), ld4(data + 6));
        const Q4 q2 = ld4(DW(xquat) + 4 * b2);
        quat1 = Q4{q2.w, -q2.x, -q2.y, -q2.z};
      }
    }
    const V3 cpos = pos1 - pos2;
    const V3 off1 = pos1 - ld3(scom + 3 * MI(body_rootid)[b1]), off2 = pos2 - ld3(scom + 3 * MI(body_rootid)[b2]);
    float S[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int d0 = 0; d0 < np; d0 += 32) {
      const int dof = d0 + lane;
      float J[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (dof < nv) {  // support.jac_dof (support.py:397): zero unless the dof's body is an ancestor-or-self of the body
        const int db = MI(dof_bodyid)[dof], dend = MI(body_subtreeend)[db];
        const bool in1 = db <= b1 && b1 < dend, in2 = db <= b2 && b2 < dend;
        const S6 cd = ld6(cdof + 6 * dof);
        const V3 ang = top(cd), lin = bot(cd), zero{0.f, 0.f, 0.f};
        const V3 jp = (in1 ? lin + cross(ang, off1) : zero) - (in2 ? lin + cross(ang, off2) : zero);
        J[0] = jp.x; J[1] = jp.y; J[2] = jp.z;
        if (weld) {
          const V3 dr = ((in1 ? ang : zero) - (in2 ? ang : zero)) * torquescale;
          const Q4 qa{-quat1.x * dr.x - quat1.y * dr.y - quat1.z * dr.z, quat1.w * dr.x + quat1.y * dr.z - quat1.z * dr.y,
                      quat1.w * dr.y + quat1.z * dr.x - quat1.x * dr.z, quat1.w * dr.z + quat1.x * dr.y - quat1.y * dr.x};  // math.py:34
          const Q4 qb = mul_quat(qa, quat);
          J[3] = 0.5f * qb.x; J[4] = 0.5f * qb.y; J[5] = 0.5f * qb.z;
        }
        const float qv = qvel[dof];
#pragma unroll
        for (int a = 0; a < 6; a++) S[a] += J[a] * qv;
      }
      if (dof < np) {
#pragma unroll
        for (int a = 0; a < 6; a++)
          if (a < nrow) DW(efc_J)[(size_t)(r0 + a) * np + dof] = J[a];
      }
    }
#pragma unroll
    for (int a = 0; a < 6; a++) S[a] = warp_sum(S[a]);
    V3 crot{0.f, 0.f, 0.f};
    if (weld) {
      const Q4 cq = mul_quat(quat1, quat);
      crot = V3{cq.x, cq.y, cq.z} * torquescale;
    }
    const float pos_imp = weld ? sqrtf(dot(cpos, cpos) + dot(crot, crot)) : norm(cpos);
    if (lane < nrow) {
      const float err[6] = {cpos.x, cpos.y, cpos.z, crot.x, crot.y, crot.z};
      float e_pos = err[0], e_vel = S[0];
#pragma unroll
      for (int a = 1; a < 6; a++) { e_pos = (lane == a) ? err[a] : e_pos; e_vel = (lane == a) ? S[a] : e_vel; }
      const int comp = lane < 3 ? 0 : 1;  // translational / rotational inverse weight
      const float invweight = MR(body_invweight0)[2 * b1 + comp] + MR(body_invweight0)[2 * b2 + comp];
      const EfcRow e = efc_row(K, e_pos, pos_imp, invweight, MR(eq_solref) + 2 * eq, MR(eq_solimp) + 5 * eq, 0.f, e_vel, 0.f);
      store_row(K, w, r0 + lane, e, CT_EQUALITY, eq);
    }
  }
  return nefc;
}

// ELL: elliptic friction cones (constraint.py:1940 _contact_elliptic) instead of pyramids (:1669).
template <bool ELL>
__device__ __forceinline__ void s_make_constraint(const KParams& K, const unsigned char* mb, int w, int lane) {
  const int nv = K.ms.nv, np = K.ms.nv_pad, njmax = K.dsz.njmax;
  int nefc = 0, ne = 0, nf = 0, nl = 0;
  const float* qpos = DW(qpos);
  const float* qvel = DW(qvel);
  const int dsbl = K.o.disableflags;
  if (!(dsbl & DSBL_CONSTRAINT)) {
    if (!(dsbl & DSBL_EQUALITY)) {
      if (K.ms.neqconnect + K.ms.neqweld > 0) nefc = emit_connect_weld(K, mb, w, lane, nefc, ne);
      for (int k0 = 0; k0 < K.ms.neqjnt; k0 += 32) {  // constraint.py:368-497
        const int k = k0 + lane;
        bool active = false;
        EfcRow e = {};
        int eq = 0, c0 = -1, c1 = -1;
        float v0 = 0.f, v1 = 0.f;
        if (k < K.ms.neqjnt) {
          eq = MI(eq_jnt_adr)[k];
          active = DW(eq_active)[eq] != 0;
          if (active) {
            const int j1 = MI(eq_obj1id)[eq], j2 = MI(eq_obj2id)[eq];
            const float* data = MR(eq_data) + 11 * eq;
            const int da1 = MI(jnt_dofadr)[j1], qa1 = MI(jnt_qposadr)[j1];
            float pos, Jqvel, invweight;
            c0 = da1;
            v0 = 1.0f;
            if (j2 > -1) {
              const int qa2 = MI(jnt_qposadr)[j2], da2 = MI(jnt_dofadr)[j2];
              const float dif = qpos[qa2] - MR(qpos0)[qa2];
              const float rhs = data[0] + dif * (data[1] + dif * (data[2] + dif * (data[3] + dif * data[4])));
              const float deriv2 = data[1] + dif * (2.0f * data[2] + dif * (3.0f * data[3] + dif * 4.0f * data[4]));
              pos = qpos[qa1] - MR(qpos0)[qa1] - rhs;
              Jqvel = qvel[da1] - qvel[da2] * deriv2;
              invweight = MR(dof_invweight0)[da1] + MR(dof_invweight0)[da2];
              c1 = da2;
              v1 = -deriv2;
            } else {
              pos = qpos[qa1] - MR(qpos0)[qa1] - data[0];
              Jqvel = qvel[da1];
              invweight = MR(dof_invweight0)[da1];
            }
            e = efc_row(K, pos, pos, invweight, MR(eq_solref) + 2 * eq, MR(eq_solimp) + 5 * eq, 0.f, Jqvel, 0.f);
          }
        }
        const int n2 = emit_simple_rows(K, w, lane, active, nefc, e, CT_EQUALITY, eq, c0, v0, c1, v1);
        ne += n2 - nefc;
        nefc = n2;
      }
    }
    if (!(dsbl & DSBL_FRICTIONLOSS)) {  // constraint.py:1114
      for (int k0 = 0; k0 < K.ms.nfrictiondof; k0 += 32) {
        const int k = k0 + lane;
        const bool active = k < K.ms.nfrictiondof;
        EfcRow e = {};
        int dof = 0;
        if (active) {
          dof = MI(dof_friction_adr)[k];
          e = efc_row(K, 0.f, 0.f, MR(dof_invweight0)[dof], MR(dof_solref) + 2 * dof, MR(dof_solimp) + 5 * dof, 0.f, qvel[dof],
                      MR(dof_frictionloss)[dof]);
        }
        const int n2 = emit_simple_rows(K, w, lane, active, nefc, e, CT_FRICTION_DOF, dof, dof, 1.0f, -1, 0.f);
        nf += n2 - nefc;
        nefc = n2;
      }
    }
    if (!(dsbl & DSBL_LIMIT)) {  // constraint.py:1317-1418
      for (int k0 = 0; k0 < K.ms.nlimit; k0 += 32) {
        const int k = k0 + lane;
        bool active = false;
        EfcRow e = {};
        int j = 0, da = -1;
        float Jv = 0.f;
        if (k < K.ms.nlimit) {
          j = MI(jnt_limited_adr)[k];
          const float q = qpos[MI(jnt_qposadr)[j]], mg = MR(jnt_margin)[j];
          const float dist_min = q - MR(jnt_range)[2 * j], dist_max = MR(jnt_range)[2 * j + 1] - q;
          const float pos = fminf(dist_min, dist_max) - mg;
          active = pos < 0.f;
          if (active) {
            da = MI(jnt_dofadr)[j];
            Jv = dist_min < dist_max ? 1.0f : -1.0f;
            e = efc_row(K, pos, pos, MR(dof_invweight0)[da], MR(jnt_solref) + 2 * j, MR(jnt_solimp) + 5 * j, mg, Jv * qvel[da], 0.f);
          }
        }
        const int n2 = emit_simple_rows(K, w, lane, active, nefc, e, CT_LIMIT_JOINT, j, da, Jv, -1, 0.f);
        nl += n2 - nefc;
        nefc = n2;
      }
      // ball joints (constraint.py:1422 _limit_ball): the rotation angle against max(range), J = -axis on the joint's three dofs
      for (int k0 = 0; k0 < K.ms.nlimitball; k0 += 32) {
        const int k = k0 + lane;
        bool active = false;
        EfcRow e = {};
        int j = 0, da = 0;
        V3 axis{0.f, 0.f, 0.f};
        if (k < K.ms.nlimitball) {
          j = MI(jnt_limited_ball_adr)[k];
          const V3 aa = quat_to_vel(normalize(ld4(qpos + MI(jnt_qposadr)[j])));
          const float angle = norm(aa), mg = MR(jnt_margin)[j];
          axis = angle == 0.f ? aa : aa * (1.0f / angle);  // math.py:261 normalize_with_norm
          const float pos = fmaxf(MR(jnt_range)[2 * j], MR(jnt_range)[2 * j + 1]) - angle - mg;
          active = pos < 0.f;
          if (active) {
            da = MI(jnt_dofadr)[j];
            const float jqvel = -axis.x * qvel[da] - axis.y * qvel[da + 1] - axis.z * qvel[da + 2];
            e = efc_row(K, pos, pos, MR(dof_invweight0)[da], MR(jnt_solref) + 2 * j, MR(jnt_solimp) + 5 * j, mg, jqvel, 0.f);
          }
        }
        const unsigned mask = __ballot_sync(FULL, active);
        const int r = nefc + __popc(mask & ((1u << lane) - 1u));
        if (active && r < K.dsz.njmax) store_row(K, w, r, e, CT_LIMIT_JOINT, j);
        unsigned mm = mask;
        while (mm) {
          const int src = __ffs(mm) - 1;
          mm &= mm - 1;
          const int rr = __shfl_sync(FULL, r, src), dd = __shfl_sync(FULL, da, src);
          const float ax = __shfl_sync(FULL, axis.x, src), ay = __shfl_sync(FULL, axis.y, src), az = __shfl_sync(FULL, axis.z, src);
          if (rr < K.dsz.njmax) {
            float* J = DW(efc_J) + (size_t)rr * K.ms.nv_pad;
            for (int c = lane; c < K.ms.nv_pad; c += 32) J[c] = c == dd ? -ax : (c == dd + 1 ? -ay : (c == dd + 2 ? -az : 0.f));
          }
        }
        nl += __popc(mask);
        nefc += __popc(mask);
      }
    }
    if (!(dsbl & DSBL_CONTACT)) {  // constraint.py:1669-1936, pyramidal cones
      const int ncon = DW(wcon_n)[0], nmp = K.ms.nmaxpyramid;
      const float* scom = DW(subtree_com);
      const float* cdof = DW(cdof);
      const float imp2 = K.o.impratio_invsqrt * K.o.impratio_invsqrt;
      for (int c0 = 0; c0 < ncon; c0 += 32) {
        // phase A, lane = contact: activity, row count and the velocity-independent row scalars
        // (_efc_row is evaluated once per contact; the rows of a pyramid differ only in J and vel)
        const int ci = c0 + lane;
        int cid = 0, condim = 1, nrow = 0;
        float pos = 0.f, includemargin = 0.f;
        EfcParams ep = {0.f, 0.f, 0.f};
        if (ci < ncon) {
          cid = DW(wcon)[ci];
          condim = K.d.contact_dim[cid];
          includemargin = K.d.contact_includemargin[cid];
          pos = K.d.contact_dist[cid] - includemargin;
          if ((K.d.contact_type[cid] & 1) && pos < 0.f) {
            nrow = condim == 1 ? 1 : (ELL ? condim : 2 * (condim - 1));
            if (!ELL) {
              const int b1 = MI(geom_bodyid)[K.d.contact_geom[2 * cid]], b2 = MI(geom_bodyid)[K.d.contact_geom[2 * cid + 1]];
              float invweight = MR(body_invweight0)[2 * b1] + MR(body_invweight0)[2 * b2];
              if (condim > 1) {
                const float fri0 = K.d.contact_friction[5 * cid];
                invweight = invweight + fri0 * fri0 * invweight;
                invweight = invweight * 2.0f * fri0 * fri0 * imp2;
              }
              ep = efc_params(K, pos, invweight, K.d.contact_solref + 2 * cid, K.d.contact_solimp + 5 * cid);
            }
          }
        }
        const int base = nefc + warp_excl_scan(nrow, lane);
        unsigned act = __ballot_sync(FULL, nrow > 0);
        nefc = __shfl_sync(FULL, base + nrow, 31);
        // phase B: Jacobian rows and their velocities, FOUR active contacts at a time.  Lane = (contact slot, 128-bit group of dofs):
        // the 8 lanes of a slot cover 32 dofs per pass, four dofs each, and write the rows with 128-bit stores; the row velocities
        // J . qvel are summed over the 8 lanes of the slot.  (One contact at a time with lane = dof spent ~650 warp instructions per
        // contact, most of them per-row address arithmetic, scalar stores and full-warp reductions: MJB_CON_QUAD = 0 keeps that version.)
#if MJB_CON_QUAD
        const int slot = lane >> 3, kq = lane & 7, np4 = np >> 2;
        while (act) {
          int srcl = -1;
          {
            unsigned t = act;
#pragma unroll
            for (int k = 0; k < 4; k++) {
              const int l = t ? __ffs(t) - 1 : -1;
              t &= t - 1;  // 0 stays 0
              if (slot == k) srcl = l;
            }
            act = t;
          }
          const bool valid = srcl >= 0;
          const int sl = valid ? srcl : 0;
          const int ccid = __shfl_sync(FULL, cid, sl), cdim = __shfl_sync(FULL, condim, sl);
          const int cnrow_any = __shfl_sync(FULL, nrow, sl);
          const int cnrow = valid ? cnrow_any : 0;
          const int row0 = __shfl_sync(FULL, base, sl);
          const float cpos_ = __shfl_sync(FULL, pos, sl), cmargin = __shfl_sync(FULL, includemargin, sl);
          const float cD = __shfl_sync(FULL, ep.D, sl), cki = __shfl_sync(FULL, ep.ki, sl), cb = __shfl_sync(FULL, ep.b, sl);
          float S[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // S[a] = sum_dof proj[a] * qvel[dof] (this lane's dofs, then the slot's)
          int wb1 = 0, wb2 = 0;
          const float* fri = K.d.contact_friction + 5 * (valid ? ccid : 0);
          if (valid) {
            const int b1 = MI(geom_bodyid)[K.d.contact_geom[2 * ccid]], b2 = MI(geom_bodyid)[K.d.contact_geom[2 * ccid + 1]];
            wb1 = MI(body_weldid)[b1];
            wb2 = MI(body_weldid)[b2];
            const V3 cpos = ld3(K.d.contact_pos + 3 * ccid);
            const M3 frame = ldm3(K.d.contact_frame + 9 * ccid);
            const V3 off1 = cpos - ld3(scom + 3 * MI(body_rootid)[wb1]), off2 = cpos - ld3(scom + 3 * MI(body_rootid)[wb2]);
            // rows of this contact: address = -1 if beyond njmax (constraint.py:1749)
            for (int r = kq; r < cnrow; r += 8) K.d.contact_efc_address[ccid * nmp + r] = (row0 + r < njmax) ? row0 + r : -1;
            for (int q0 = 0; q0 < np4; q0 += 8) {
              const int qd = q0 + kq;
              if (qd >= np4) continue;
              // per-dof Jacobian differences projected on the frame axes (support.jac_dof, support.py:397)
              // pr[0] = normal, pr[1..2] = tangents (translation), pr[3..5] = frame rows . rotation; [k] = dof 4 qd + k
              float pr[6][4];
#pragma unroll
              for (int a2 = 0; a2 < 6; a2++)
#pragma unroll
                for (int k = 0; k < 4; k++) pr[a2][k] = 0.f;
#pragma unroll
              for (int k = 0; k < 4; k++) {
                const int dof = 4 * qd + k;
                if (dof < nv) {
                  const int db = MI(dof_bodyid)[dof];
                  // dof's body is an ancestor-or-self of body  <=>  db <= body < subtreeend[db]
                  const int dend = MI(body_subtreeend)[db];
                  const bool in1 = wb1 != 0 && db <= wb1 && wb1 < dend, in2 = wb2 != 0 && db <= wb2 && wb2 < dend;
                  if (in1 != in2) {  // common ancestors contribute exactly zero
                    const S6 cd = ld6(cdof + 6 * dof);
                    const V3 ang = top(cd), lin = bot(cd);
                    V3 dp = lin + cross(ang, in2 ? off2 : off1), dr = ang;
                    if (in1) { dp = V3{0.f, 0.f, 0.f} - dp; dr = V3{0.f, 0.f, 0.f} - dr; }
                    const float qv = qvel[dof];
#pragma unroll
                    for (int a2 = 0; a2 < 3; a2++) {
                      pr[a2][k] = frame.m[3 * a2] * dp.x + frame.m[3 * a2 + 1] * dp.y + frame.m[3 * a2 + 2] * dp.z;
                      pr[3 + a2][k] = frame.m[3 * a2] * dr.x + frame.m[3 * a2 + 1] * dr.y + frame.m[3 * a2 + 2] * dr.z;
                      S[a2] += pr[a2][k] * qv;
                      S[3 + a2] += pr[3 + a2][k] * qv;
                    }
                  }
                }
              }
              float* Jp = DW(efc_J) + (size_t)row0 * np + 4 * qd;
              const int rleft = njmax - row0;  // rows [0, rleft) of this contact exist
              if (ELL || cdim == 1) {  // row a2 is the frame axis itself: normal, tangents, then the rotational rows
#pragma unroll
                for (int a2 = 0; a2 < 6; a2++)
                  if (a2 < cnrow && a2 < rleft) *reinterpret_cast<float4*>(Jp + (size_t)a2 * np) = make_float4(pr[a2][0], pr[a2][1], pr[a2][2], pr[a2][3]);
              } else {  // pyramid edges: rows 2 (a2 - 1), 2 (a2 - 1) + 1 = normal +- friction[a2 - 1] * axis a2
#pragma unroll
                for (int a2 = 1; a2 < 6; a2++) {
                  if (a2 < cdim) {
                    const float f = fri[a2 - 1];
                    const int ra = 2 * (a2 - 1);
                    if (ra < rleft)
                      *reinterpret_cast<float4*>(Jp + (size_t)ra * np) =
                          make_float4(pr[0][0] + f * pr[a2][0], pr[0][1] + f * pr[a2][1], pr[0][2] + f * pr[a2][2], pr[0][3] + f * pr[a2][3]);
                    if (ra + 1 < rleft)
                      *reinterpret_cast<float4*>(Jp + (size_t)(ra + 1) * np) =
                          make_float4(pr[0][0] - f * pr[a2][0], pr[0][1] - f * pr[a2][1], pr[0][2] - f * pr[a2][2], pr[0][3] - f * pr[a2][3]);
                  }
                }
              }
            }
          }
          // efc_vel = J . qvel: sums over the 8 lanes of the slot (all lanes take part; idle slots add zeros)
#pragma unroll
          for (int a2 = 0; a2 < 6; a2++) {
            if (a2 < K.ms.nmaxcondim) {  // warp-uniform
              S[a2] += __shfl_xor_sync(FULL, S[a2], 1);
              S[a2] += __shfl_xor_sync(FULL, S[a2], 2);
              S[a2] += __shfl_xor_sync(FULL, S[a2], 4);
            }
          }
          for (int r = kq; r < cnrow; r += 8) {  // the row scalars, one lane per row
            if (row0 + r >= njmax) continue;
            if (ELL) {  // constraint.py:2146-2206: per-row impedance parameters
              float vel = S[0];
#pragma unroll
              for (int a2 = 1; a2 < 6; a2++) vel = (r == a2) ? S[a2] : vel;
              float invweight = MR(body_invweight0)[2 * wb1] + MR(body_invweight0)[2 * wb2];  // after the weld mapping (:2049)
              const float* ref = K.d.contact_solref + 2 * ccid;
              if (r > 0) {
                const float* srf = K.d.contact_solreffriction + 2 * ccid;
                if (srf[0] != 0.f || srf[1] != 0.f) ref = srf;
                invweight *= imp2;
                if (r > 1) invweight *= fri[0] * fri[0] / (fri[r - 1] * fri[r - 1]);
              }
              const EfcParams p = efc_params(K, cpos_, invweight, ref, K.d.contact_solimp + 5 * ccid);
              const float pos_aref = r == 0 ? cpos_ : 0.f;
              EfcRow e;
              e.D = p.D;
              e.vel = vel;
              e.aref = -p.ki * pos_aref - p.b * vel;
              e.pos = pos_aref + cmargin;
              e.margin = cmargin;
              e.frictionloss = 0.f;
              store_row(K, w, row0 + r, e, cdim == 1 ? CT_CONTACT_FRICTIONLESS : CT_CONTACT_ELLIPTIC, ccid);
            } else {
              float vel = S[0];
              if (cdim > 1) {
                const int dimid2 = r / 2 + 1;
                float Si = S[1];
#pragma unroll
                for (int a2 = 2; a2 < 6; a2++) Si = (dimid2 == a2) ? S[a2] : Si;
                const float frii = fri[dimid2 - 1];
                vel += (r & 1) ? -Si * frii : Si * frii;
              }
              EfcRow e;
              e.D = cD;
              e.vel = vel;
              e.aref = -cki * cpos_ - cb * vel;
              e.pos = cpos_ + cmargin;
              e.margin = cmargin;
              e.frictionloss = 0.f;
              store_row(K, w, row0 + r, e, cdim == 1 ? CT_CONTACT_FRICTIONLESS : CT_CONTACT_PYRAMIDAL, ccid);
            }
          }
        }
#else
        while (act) {
          const int src = __ffs(act) - 1;
          act &= act - 1;
          const int ccid = __shfl_sync(FULL, cid, src), cdim = __shfl_sync(FULL, condim, src), cnrow = __shfl_sync(FULL, nrow, src);
          const int row0 = __shfl_sync(FULL, base, src);
          const float cpos_ = __shfl_sync(FULL, pos, src), cmargin = __shfl_sync(FULL, includemargin, src);
          const float cD = __shfl_sync(FULL, ep.D, src), cki = __shfl_sync(FULL, ep.ki, src), cb = __shfl_sync(FULL, ep.b, src);
          const int b1 = MI(geom_bodyid)[K.d.contact_geom[2 * ccid]], b2 = MI(geom_bodyid)[K.d.contact_geom[2 * ccid + 1]];
          const int wb1 = MI(body_weldid)[b1], wb2 = MI(body_weldid)[b2];
          const V3 cpos = ld3(K.d.contact_pos + 3 * ccid);
          const M3 frame = ldm3(K.d.contact_frame + 9 * ccid);
          const float* fri = K.d.contact_friction + 5 * ccid;
          const V3 off1 = cpos - ld3(scom + 3 * MI(body_rootid)[wb1]), off2 = cpos - ld3(scom + 3 * MI(body_rootid)[wb2]);
          // rows of this contact: address = -1 if beyond njmax (constraint.py:1749)
          if (lane < cnrow) K.d.contact_efc_address[ccid * nmp + lane] = (row0 + lane < njmax) ? row0 + lane : -1;
          // per-dof Jacobian differences projected on the frame axes (support.jac_dof, support.py:397)
          // proj[0] = normal, proj[1..2] = tangents (translation), proj[3..5] = frame rows . rotation
          float S[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // S[a] = sum_dof proj[a] * qvel[dof]
          for (int d0 = 0; d0 < np; d0 += 32) {
            const int dof = d0 + lane;
            float proj[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (dof < nv) {
              const int db = MI(dof_bodyid)[dof];
              // dof's body is an ancestor-or-self of body  <=>  db <= body < subtreeend[db]
              const int dend = MI(body_subtreeend)[db];
              const bool in1 = wb1 != 0 && db <= wb1 && wb1 < dend, in2 = wb2 != 0 && db <= wb2 && wb2 < dend;
              if (in1 != in2) {  // common ancestors contribute exactly zero
                const S6 cd = ld6(cdof + 6 * dof);
                const V3 ang = top(cd), lin = bot(cd);
                V3 dp = lin + cross(ang, in2 ? off2 : off1), dr = ang;
                if (in1) { dp = V3{0.f, 0.f, 0.f} - dp; dr = V3{0.f, 0.f, 0.f} - dr; }
                const float qv = qvel[dof];
#pragma unroll
                for (int a = 0; a < 3; a++) {
                  proj[a] = frame.m[3 * a] * dp.x + frame.m[3 * a + 1] * dp.y + frame.m[3 * a + 2] * dp.z;
                  proj[3 + a] = frame.m[3 * a] * dr.x + frame.m[3 * a + 1] * dr.y + frame.m[3 * a + 2] * dr.z;
                  S[a] += proj[a] * qv;
                  S[3 + a] += proj[3 + a] * qv;
                }
              }
            }
            if (dof < np) {
              for (int dimid = 0; dimid < cnrow; dimid++) {
                const int r = row0 + dimid;
                float Jv = proj[0];
                if (ELL) {  // row dimid is the frame axis itself: normal, tangents, then the rotational rows
#pragma unroll
                  for (int a = 1; a < 6; a++) Jv = (dimid == a) ? proj[a] : Jv;
                } else if (cdim > 1) {
                  const int dimid2 = dimid / 2 + 1;
                  float Ji = proj[1];
#pragma unroll
                  for (int a = 2; a < 6; a++) Ji = (dimid2 == a) ? proj[a] : Ji;
                  const float frii = fri[dimid2 - 1];
                  Jv += (dimid & 1) ? -Ji * frii : Ji * frii;
                }
                if (r < njmax) DW(efc_J)[(size_t)r * np + dof] = Jv;
              }
            }
          }
          // efc_vel = J . qvel of row dimid = S[0] +- friction * S[dimid / 2 + 1]
#pragma unroll
          for (int a = 0; a < 6; a++)
            if (a < cdim) S[a] = warp_sum(S[a]);
          if (ELL && lane < cnrow && row0 + lane < njmax) {  // constraint.py:2146-2206: per-row impedance parameters
            float vel = S[0];
#pragma unroll
            for (int a = 1; a < 6; a++) vel = (lane == a) ? S[a] : vel;
            float invweight = MR(body_invweight0)[2 * wb1] + MR(body_invweight0)[2 * wb2];  // after the weld mapping (:2049)
            const float* ref = K.d.contact_solref + 2 * ccid;
            if (lane > 0) {
              const float* srf = K.d.contact_solreffriction + 2 * ccid;
              if (srf[0] != 0.f || srf[1] != 0.f) ref = srf;
              invweight *= imp2;
              if (lane > 1) invweight *= fri[0] * fri[0] / (fri[lane - 1] * fri[lane - 1]);
            }
            const EfcParams p = efc_params(K, cpos_, invweight, ref, K.d.contact_solimp + 5 * ccid);
            const float pos_aref = lane == 0 ? cpos_ : 0.f;
            EfcRow e;
            e.D = p.D;
            e.vel = vel;
            e.aref = -p.ki * pos_aref - p.b * vel;
            e.pos = pos_aref + cmargin;
            e.margin = cmargin;
            e.frictionloss = 0.f;
            store_row(K, w, row0 + lane, e, cdim == 1 ? CT_CONTACT_FRICTIONLESS : CT_CONTACT_ELLIPTIC, ccid);
          }
          if (!ELL && lane < cnrow && row0 + lane < njmax) {
            float vel = S[0];
            if (cdim > 1) {
              const int dimid2 = lane / 2 + 1;
              float Si = S[1];
#pragma unroll
              for (int a = 2; a < 6; a++) Si = (dimid2 == a) ? S[a] : Si;
              const float frii = fri[dimid2 - 1];
              vel += (lane & 1) ? -Si * frii : Si * frii;
            }
            EfcRow e;
            e.D = cD;
            e.vel = vel;
            e.aref = -cki * cpos_ - cb * vel;
            e.pos = cpos_ + cmargin;
            e.margin = cmargin;
            e.frictionloss = 0.f;
            store_row(K, w, row0 + lane, e, cdim == 1 ? CT_CONTACT_FRICTIONLESS : CT_CONTACT_PYRAMIDAL, ccid);
          }
        }
#endif  // MJB_CON_QUAD
      }
    }
  }
  if (lane == 0) {
    DW(ne)[0] = ne;
    DW(nf)[0] = nf;
    DW(nl)[0] = nl;
    DW(nefc)[0] = nefc;
    if (nefc > njmax) DW(status)[0] |= MJB_STATUS_NEFC_OVERFLOW;
  }
  __syncwarp();
}
