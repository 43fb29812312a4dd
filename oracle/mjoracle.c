/* mjoracle.c -- CPU restatement (float64, one world at a time) of MuJoCo Warp's batched step.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity oracle for libmjb200.so: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 * The product path (mujoco_warp_b200/) never imports, links or calls anything in oracle/.
 *
 * PARITY PINNED to outputs of the reference itself: the reference's unmodified source files are executed here in a float64
 * numpy interpreter of the Warp API (tests/golden/warp_interp.py; `warp-lang` and `mujoco` themselves are not installable),
 * and the resulting golden vectors are committed: tests/golden/reference_step.npz (put_model -> make_data -> forward / step
 * on 19 cases, every Data field), primitive_core.npz (pair functions), ccd_pairs.npz (GJK / EPA).  The oracle agrees with
 * them to <= 1e-9 (tests/test_oracle_reference_golden.py, test_oracle_golden.py, test_oracle_ccd_golden.py).  Also pinned
 * by the reference's closed-form known-answer tests (tests/test_oracle_kat.py) and physics invariants
 * (tests/test_oracle_invariants.py).  Not pinned: the MJCF compiler's derived constants (no MuJoCo to compare with).
 *
 * Every function cites the reference file:line (relative to /root/reference/mujoco_warp/_src/)
 * whose algorithm it restates.  Ordering choices where the reference is nondeterministic
 * (atomic appends): contacts are emitted in nxn pair order, efc rows in MuJoCo-C order
 * (equality, friction, limits by joint id, contacts by contact index then pyramid edge).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/mjb_fields.h"

typedef double real;

/* convex narrowphase (GJK / EPA): one source, compiled here in float64 and by the CUDA library in float32 */
#define CCD_REAL double
#define CCD_FN static inline
#define CCD_SQRT sqrt
#define CCD_FABS fabs
#include <stdio.h>
/* #define CCD_DEBUG_PRINT(...) fprintf(stderr, __VA_ARGS__)   -- traces the GJK / EPA of height-field x mesh pairs (mjb_ccd_core.h) */
#include "../mujoco_warp_b200/csrc/mjb_ccd_core.h"

#define MJ_MINVAL 1e-15
#define MJ_MAXVAL 1e10
#define MJ_MINIMP 1e-4
#define MJ_MAXIMP 0.9999
#define MJ_MINMU 1e-5

enum { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum { GEOM_PLANE = 0, GEOM_HFIELD, GEOM_SPHERE, GEOM_CAPSULE, GEOM_ELLIPSOID, GEOM_CYLINDER, GEOM_BOX, GEOM_MESH };
enum { DSBL_CONSTRAINT = 1, DSBL_EQUALITY = 2, DSBL_FRICTIONLOSS = 4, DSBL_LIMIT = 8, DSBL_CONTACT = 16,
       DSBL_SPRING = 32, DSBL_DAMPER = 64, DSBL_GRAVITY = 128, DSBL_CLAMPCTRL = 256, DSBL_WARMSTART = 512,
       DSBL_ACTUATION = 2048, DSBL_REFSAFE = 4096, DSBL_EULERDAMP = 32768, DSBL_NATIVECCD = 131072 };
enum { ENBL_ENERGY = 2 };
enum { OBJ_BODY = 1, OBJ_SITE = 6 }; /* mjtObj values of eq_objtype */ /* types.py EnableBit.ENERGY */
enum { CT_EQUALITY = 0, CT_FRICTION_DOF = 1, CT_LIMIT_JOINT = 3, CT_CONTACT_FRICTIONLESS = 5,
       CT_CONTACT_PYRAMIDAL = 6, CT_CONTACT_ELLIPTIC = 7 };
enum { ST_SATISFIED = 0, ST_QUADRATIC = 1, ST_LINEARNEG = 2, ST_LINEARPOS = 3, ST_CONE = 4 };
enum { SOLVER_CG = 1, SOLVER_NEWTON = 2 };
enum { BF_PLANE = 1, BF_SPHERE = 2, BF_AABB = 4, BF_OBB = 8 };

typedef struct {
#define X(n) int n;
  MJB_MODEL_SIZES(X)
#undef X
#define X(n) const int* n;
  MJB_MODEL_INT_FIELDS(X)
#undef X
#define X(n) const real* n;
  MJB_MODEL_REAL_FIELDS(X)
#undef X
} OModel;

typedef struct {
#define X(n) int n;
  MJB_OPT_INT_FIELDS(X)
#undef X
#define X(n) real n;
  MJB_OPT_REAL_FIELDS(X)
#undef X
} OOpt;

typedef struct {
#define X(n) int n;
  MJB_DATA_SIZES(X)
#undef X
#define X(n, c) int* n;
  MJB_DATA_INT_FIELDS(X)
#undef X
#define X(n, c) real* n;
  MJB_DATA_REAL_FIELDS(X)
#undef X
} OData;

/* ------------------------------------------------------------------------------------------ */
/* math (reference math.py:24-334)                                                             */
/* ------------------------------------------------------------------------------------------ */
static inline real dot3(const real* a, const real* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline void cross3(real* r, const real* a, const real* b) {
  real x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
static inline real norm3(const real* a) { return sqrt(dot3(a, a)); }
/* wp.normalize semantics: zero vector stays zero */
static inline void normalize3(real* a) {
  real n = norm3(a);
  if (n > 0) { a[0] /= n; a[1] /= n; a[2] /= n; } else { a[0] = a[1] = a[2] = 0; }
}
static inline void normalize4(real* q) {
  real n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n > 0) { q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n; } else { q[0] = q[1] = q[2] = q[3] = 0; }
}
/* math.py:24 */
static inline void mul_quat(real* r, const real* u, const real* v) {
  real a = u[0] * v[0] - u[1] * v[1] - u[2] * v[2] - u[3] * v[3];
  real b = u[0] * v[1] + u[1] * v[0] + u[2] * v[3] - u[3] * v[2];
  real c = u[0] * v[2] - u[1] * v[3] + u[2] * v[0] + u[3] * v[1];
  real d = u[0] * v[3] + u[1] * v[2] - u[2] * v[1] + u[3] * v[0];
  r[0] = a; r[1] = b; r[2] = c; r[3] = d;
}
/* math.py:45 */
static inline void rot_vec_quat(real* r, const real* v, const real* q) {
  real s = q[0]; const real* u = q + 1;
  real uv = dot3(u, v), uu = dot3(u, u), c[3];
  cross3(c, u, v);
  for (int i = 0; i < 3; i++) r[i] = 2.0 * uv * u[i] + (s * s - uu) * v[i] + 2.0 * s * c[i];
}
/* math.py:53 */
static inline void axis_angle_to_quat(real* q, const real* axis, real angle) {
  real s = sin(angle * 0.5), c = cos(angle * 0.5);
  q[0] = c; q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
/* math.py:60 (row-major 3x3) */
static inline void quat_to_mat(real* m, const real* q) {
  real q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
  real q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3];
  real q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
  m[0] = q00 + q11 - q22 - q33; m[1] = 2.0 * (q12 - q03); m[2] = 2.0 * (q13 + q02);
  m[3] = 2.0 * (q12 + q03); m[4] = q00 - q11 + q22 - q33; m[5] = 2.0 * (q23 - q01);
  m[6] = 2.0 * (q13 - q02); m[7] = 2.0 * (q23 + q01); m[8] = q00 - q11 - q22 + q33;
}
/* math.py:121 mju_mulInertVec */
static inline void inert_vec(real* r, const real* i, const real* v) {
  real o[6];
  o[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  o[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  o[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  o[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  o[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  o[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
  memcpy(r, o, sizeof(o));
}
/* math.py:134 */
static inline void motion_cross(real* r, const real* u, const real* v) {
  real a[3], b[3], c[3];
  cross3(a, u, v);
  cross3(b, u + 3, v);
  cross3(c, u, v + 3);
  r[0] = a[0]; r[1] = a[1]; r[2] = a[2];
  r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
/* math.py:148 */
static inline void motion_cross_force(real* r, const real* v, const real* f) {
  real a[3], b[3], c[3];
  cross3(a, v, f);
  cross3(b, v + 3, f + 3);
  cross3(c, v, f + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2];
  r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}
/* math.py:162 */
static inline void quat_to_vel(real* r, const real* q) {
  real s = norm3(q + 1);
  if (s == 0.0) { r[0] = r[1] = r[2] = 0; return; }
  real speed = 2.0 * atan2(s, q[0]);
  if (speed > M_PI) speed -= 2.0 * M_PI;
  for (int i = 0; i < 3; i++) r[i] = q[1 + i] * speed / s;
}
/* math.py:178 */
static inline void quat_sub(real* r, const real* qa, const real* qb) {
  real qn[4] = {qb[0], -qb[1], -qb[2], -qb[3]}, qd[4];
  mul_quat(qd, qn, qa);
  quat_to_vel(r, qd);
}
/* math.py:189 */
static inline void quat_integrate(real* q, const real* v, real dt) {
  real n = norm3(v), ax[3] = {v[0], v[1], v[2]}, qr[4], qq[4] = {q[0], q[1], q[2], q[3]};
  normalize3(ax);
  axis_angle_to_quat(qr, ax, dt * n);
  normalize4(qq);
  mul_quat(q, qq, qr);
  normalize4(q);
}
static inline real safe_div(real x, real y) { return x / (y != 0.0 ? y : MJ_MINVAL); }
/* wp.clamp(x, lo, hi) = min(max(lo, x), hi): when lo > hi (a mixed solimp with dmin > dmax, constraint.py:108) the result is hi */
static inline real clampr(real x, real lo, real hi) { const real t = x < lo ? lo : x; return t > hi ? hi : t; }
/* math.py:203,247 make_frame: rows (a, b, c) */
static void make_frame(real* f, const real* a_in) {
  real a[3] = {a_in[0], a_in[1], a_in[2]}, b[3], c[3];
  normalize3(a);
  if (-0.5 < a[1] && a[1] < 0.5) { b[0] = 0; b[1] = 1; b[2] = 0; } else { b[0] = 0; b[1] = 0; b[2] = 1; }
  real ab = dot3(a, b);
  for (int i = 0; i < 3; i++) b[i] -= a[i] * ab;
  normalize3(b);
  if (norm3(a) == 0.0) b[0] = b[1] = b[2] = 0;
  cross3(c, a, b);
  for (int i = 0; i < 3; i++) { f[i] = a[i]; f[3 + i] = b[i]; f[6 + i] = c[i]; }
}

/* ------------------------------------------------------------------------------------------ */
/* per-world view                                                                              */
/* ------------------------------------------------------------------------------------------ */
static void world_view(const OModel* m, const OData* d, int w, OData* v) {
#define X(n) const int n = m->n; (void)n;
  MJB_MODEL_SIZES(X)
#undef X
#define X(n) const int n = d->n; (void)n;
  MJB_DATA_SIZES(X)
#undef X
  *v = *d;
#define X(n, c) v->n = d->n ? d->n + (size_t)w * (size_t)(c) : 0;
  MJB_DATA_INT_FIELDS(X)
  MJB_DATA_REAL_FIELDS(X)
#undef X
}

/* ------------------------------------------------------------------------------------------ */
/* smooth.kinematics (smooth.py:45-224, 358)                                                   */
/* ------------------------------------------------------------------------------------------ */
static void o_kinematics(const OModel* m, OData* d) {
  const real* qpos = d->qpos;
  for (int b = 1; b < m->nbody; b++) {
    int pid = m->body_parentid[b], jadr = m->body_jntadr[b], jnum = m->body_jntnum[b];
    real xpos[3], xquat[4];
    if (jnum == 1 && m->jnt_type[jadr] == JNT_FREE) { /* smooth.py:83-95 */
      int qa = m->jnt_qposadr[jadr];
      for (int i = 0; i < 3; i++) xpos[i] = qpos[qa + i];
      for (int i = 0; i < 4; i++) xquat[i] = qpos[qa + 3 + i];
      normalize4(xquat);
      memcpy(d->xpos + 3 * b, xpos, sizeof(xpos));
      memcpy(d->xquat + 4 * b, xquat, sizeof(xquat));
      memcpy(d->xanchor + 3 * jadr, xpos, sizeof(xpos));
      memcpy(d->xaxis + 3 * jadr, m->jnt_axis + 3 * jadr, 3 * sizeof(real));
      continue;
    }
    int mocapid = m->body_mocapid[b];
    if (mocapid >= 0) { /* smooth.py:103-106 */
      memcpy(xpos, d->mocap_pos + 3 * mocapid, sizeof(xpos));
      memcpy(xquat, d->mocap_quat + 4 * mocapid, sizeof(xquat));
    } else {
      memcpy(xpos, m->body_pos + 3 * b, sizeof(xpos));
      memcpy(xquat, m->body_quat + 4 * b, sizeof(xquat));
    }
    { /* smooth.py:111-113 (pid >= 0 always) */
      real t[3], q[4];
      rot_vec_quat(t, xpos, d->xquat + 4 * pid);
      for (int i = 0; i < 3; i++) xpos[i] = t[i] + d->xpos[3 * pid + i];
      mul_quat(q, d->xquat + 4 * pid, xquat);
      memcpy(xquat, q, sizeof(q));
    }
    for (int j = jadr; j < jadr + jnum; j++) { /* smooth.py:115-139 */
      int qa = m->jnt_qposadr[j], jt = m->jnt_type[j];
      const real* jaxis = m->jnt_axis + 3 * j;
      real xanchor[3], xaxis[3], t[3];
      rot_vec_quat(t, m->jnt_pos + 3 * j, xquat);
      for (int i = 0; i < 3; i++) xanchor[i] = t[i] + xpos[i];
      rot_vec_quat(xaxis, jaxis, xquat);
      if (jt == JNT_BALL) {
        real ql[4] = {qpos[qa], qpos[qa + 1], qpos[qa + 2], qpos[qa + 3]}, q[4];
        normalize4(ql);
        mul_quat(q, xquat, ql);
        memcpy(xquat, q, sizeof(q));
        rot_vec_quat(t, m->jnt_pos + 3 * j, xquat);
        for (int i = 0; i < 3; i++) xpos[i] = xanchor[i] - t[i];
      } else if (jt == JNT_SLIDE) {
        for (int i = 0; i < 3; i++) xpos[i] += xaxis[i] * (qpos[qa] - m->qpos0[qa]);
      } else if (jt == JNT_HINGE) {
        real ql[4], q[4];
        axis_angle_to_quat(ql, jaxis, qpos[qa] - m->qpos0[qa]);
        mul_quat(q, xquat, ql);
        memcpy(xquat, q, sizeof(q));
        rot_vec_quat(t, m->jnt_pos + 3 * j, xquat);
        for (int i = 0; i < 3; i++) xpos[i] = xanchor[i] - t[i];
      }
      memcpy(d->xanchor + 3 * j, xanchor, sizeof(xanchor));
      memcpy(d->xaxis + 3 * j, xaxis, sizeof(xaxis));
    }
    normalize4(xquat);
    memcpy(d->xpos + 3 * b, xpos, sizeof(xpos));
    memcpy(d->xquat + 4 * b, xquat, sizeof(xquat));
  }
  for (int b = 0; b < m->nbody; b++) { /* smooth.py:147-173 */
    real t[3], q[4];
    quat_to_mat(d->xmat + 9 * b, d->xquat + 4 * b);
    rot_vec_quat(t, m->body_ipos + 3 * b, d->xquat + 4 * b);
    for (int i = 0; i < 3; i++) d->xipos[3 * b + i] = d->xpos[3 * b + i] + t[i];
    mul_quat(q, d->xquat + 4 * b, m->body_iquat + 4 * b);
    quat_to_mat(d->ximat + 9 * b, q);
  }
  for (int g = 0; g < m->ngeom; g++) { /* smooth.py:177-203 */
    if (m->geom_static[g]) continue;
    int b = m->geom_bodyid[g];
    real t[3], q[4];
    rot_vec_quat(t, m->geom_pos + 3 * g, d->xquat + 4 * b);
    for (int i = 0; i < 3; i++) d->geom_xpos[3 * g + i] = d->xpos[3 * b + i] + t[i];
    mul_quat(q, d->xquat + 4 * b, m->geom_quat + 4 * g);
    quat_to_mat(d->geom_xmat + 9 * g, q);
  }
  for (int s = 0; s < m->nsite; s++) { /* smooth.py:207-224 */
    int b = m->site_bodyid[s];
    real t[3], q[4];
    rot_vec_quat(t, m->site_pos + 3 * s, d->xquat + 4 * b);
    for (int i = 0; i < 3; i++) d->site_xpos[3 * s + i] = d->xpos[3 * b + i] + t[i];
    mul_quat(q, d->xquat + 4 * b, m->site_quat + 4 * s);
    quat_to_mat(d->site_xmat + 9 * s, q);
  }
}

/* smooth.com_pos (smooth.py:464-632) */
static void o_com_pos(const OModel* m, OData* d) {
  int nb = m->nbody;
  for (int b = 0; b < nb; b++)
    for (int i = 0; i < 3; i++) d->subtree_com[3 * b + i] = d->xipos[3 * b + i] * m->body_mass[b];
  for (int b = nb - 1; b > 0; b--) { /* children have larger ids than parents */
    int p = m->body_parentid[b];
    for (int i = 0; i < 3; i++) d->subtree_com[3 * p + i] += d->subtree_com[3 * b + i];
  }
  for (int b = 0; b < nb; b++) {
    real mass = m->body_subtreemass[b];
    if (mass != 0.0)
      for (int i = 0; i < 3; i++) d->subtree_com[3 * b + i] /= mass;
  }
  for (int b = 0; b < nb; b++) { /* smooth.py:511-553 */
    const real* mat = d->ximat + 9 * b;
    const real* inert = m->body_inertia + 3 * b;
    real mass = m->body_mass[b], dif[3], tmp[9];
    for (int i = 0; i < 3; i++) dif[i] = d->xipos[3 * b + i] - d->subtree_com[3 * m->body_rootid[b] + i];
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) {
        real s = 0;
        for (int k = 0; k < 3; k++) s += mat[3 * r + k] * inert[k] * mat[3 * c + k];
        tmp[3 * r + c] = s;
      }
    real* res = d->cinert + 10 * b;
    res[0] = tmp[0] + mass * (dif[1] * dif[1] + dif[2] * dif[2]);
    res[1] = tmp[4] + mass * (dif[0] * dif[0] + dif[2] * dif[2]);
    res[2] = tmp[8] + mass * (dif[0] * dif[0] + dif[1] * dif[1]);
    res[3] = tmp[1] - mass * dif[0] * dif[1];
    res[4] = tmp[2] - mass * dif[0] * dif[2];
    res[5] = tmp[5] - mass * dif[1] * dif[2];
    res[6] = mass * dif[0]; res[7] = mass * dif[1]; res[8] = mass * dif[2];
    res[9] = mass;
  }
  for (int j = 0; j < m->njnt; j++) { /* smooth.py:557-598 */
    int b = m->jnt_bodyid[j], da = m->jnt_dofadr[j], jt = m->jnt_type[j];
    const real* xmat = d->xmat + 9 * b;
    real off[3];
    for (int i = 0; i < 3; i++) off[i] = d->subtree_com[3 * m->body_rootid[b] + i] - d->xanchor[3 * j + i];
    real* res = d->cdof + 6 * da;
    if (jt == JNT_FREE || jt == JNT_BALL) {
      if (jt == JNT_FREE) {
        memset(res, 0, 18 * sizeof(real));
        res[3] = 1; res[6 + 4] = 1; res[12 + 5] = 1;
        res += 18;
      }
      for (int a = 0; a < 3; a++) { /* columns of xmat (rows of its transpose) */
        real ax[3] = {xmat[a], xmat[3 + a], xmat[6 + a]}, c[3];
        cross3(c, ax, off);
        for (int i = 0; i < 3; i++) { res[6 * a + i] = ax[i]; res[6 * a + 3 + i] = c[i]; }
      }
    } else if (jt == JNT_SLIDE) {
      for (int i = 0; i < 3; i++) { res[i] = 0; res[3 + i] = d->xaxis[3 * j + i]; }
    } else {
      real c[3];
      cross3(c, d->xaxis + 3 * j, off);
      for (int i = 0; i < 3; i++) { res[i] = d->xaxis[3 * j + i]; res[3 + i] = c[i]; }
    }
  }
}

/* smooth.crb (smooth.py:807-912), dense qM */
static void o_crb(const OModel* m, OData* d) {
  int nb = m->nbody, nv = m->nv, np = m->nv_pad;
  memcpy(d->crb, d->cinert, sizeof(real) * 10 * nb);
  for (int b = nb - 1; b > 0; b--) {
    int p = m->body_parentid[b];
    if (p == 0) continue; /* smooth.py:820 */
    for (int i = 0; i < 10; i++) d->crb[10 * p + i] += d->crb[10 * b + i];
  }
  memset(d->qM, 0, sizeof(real) * np * np);
  for (int i = 0; i < nv; i++) { /* smooth.py:856-885 */
    real buf[6];
    inert_vec(buf, d->crb + 10 * m->dof_bodyid[i], d->cdof + 6 * i);
    real s = m->dof_armature[i];
    for (int k = 0; k < 6; k++) s += d->cdof[6 * i + k] * buf[k];
    d->qM[i * np + i] = s;
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) {
      real v = 0;
      for (int k = 0; k < 6; k++) v += d->cdof[6 * j + k] * buf[k];
      d->qM[i * np + j] += v;
      d->qM[j * np + i] += v;
    }
  }
}

/* dense Cholesky M = L L^T (wp.tile_cholesky; smooth.py:1068-1100).  A has row stride lda. */
static int chol_factor(real* L, int ldl, const real* A, int lda, int n) {
  for (int j = 0; j < n; j++) {
    real s = A[j * lda + j];
    for (int k = 0; k < j; k++) s -= L[j * ldl + k] * L[j * ldl + k];
    if (s <= 0) s = MJ_MINVAL;
    real ljj = sqrt(s);
    L[j * ldl + j] = ljj;
    for (int i = j + 1; i < n; i++) {
      real t = A[i * lda + j];
      for (int k = 0; k < j; k++) t -= L[i * ldl + k] * L[j * ldl + k];
      L[i * ldl + j] = t / ljj;
    }
    for (int i = 0; i < j; i++) L[i * ldl + j] = 0;
  }
  return 0;
}
/* wp.tile_cholesky_solve (smooth.py:2783-2802): x = (L L^T)^-1 y */
static void chol_solve(real* x, const real* L, int ldl, const real* y, int n) {
  for (int i = 0; i < n; i++) {
    real s = y[i];
    for (int k = 0; k < i; k++) s -= L[i * ldl + k] * x[k];
    x[i] = s / L[i * ldl + i];
  }
  for (int i = n - 1; i >= 0; i--) {
    real s = x[i];
    for (int k = i + 1; k < n; k++) s -= L[k * ldl + i] * x[k];
    x[i] = s / L[i * ldl + i];
  }
}
/* smooth.factor_m dense (smooth.py:1091-1109): the per-tree tiles of the reference form a
 * block-diagonal matrix, so one full factorisation gives the same L. */
static void o_factor_m(const OModel* m, OData* d) { chol_factor(d->qLD, m->nv, d->qM, m->nv_pad, m->nv); }

/* support.mul_m dense (support.py:105-129) */
static void o_mul_m(const OModel* m, const OData* d, real* res, const real* vec) {
  for (int i = 0; i < m->nv; i++) {
    real s = 0;
    for (int j = 0; j < m->nv; j++) s += d->qM[i * m->nv_pad + j] * vec[j];
    res[i] = s;
  }
}

/* ------------------------------------------------------------------------------------------ */
/* collision: NXN broadphase + primitive narrowphase                                           */
/* ------------------------------------------------------------------------------------------ */
/* collision_driver.py:91-103 */
static int plane_filter(real s1, real s2, real m1, real m2, const real* p1, const real* p2, const real* r1, const real* r2) {
  real dif[3];
  if (s1 == 0.0) {
    real n[3] = {r1[2], r1[5], r1[8]};
    for (int i = 0; i < 3; i++) dif[i] = p2[i] - p1[i];
    return dot3(dif, n) <= s2 + m1 + m2;
  } else if (s2 == 0.0) {
    real n[3] = {r2[2], r2[5], r2[8]};
    for (int i = 0; i < 3; i++) dif[i] = p1[i] - p2[i];
    return dot3(dif, n) <= s1 + m1 + m2;
  }
  return 1;
}
/* collision_driver.py:107-111 */
static int sphere_filter(real s1, real s2, real m1, real m2, const real* p1, const real* p2) {
  real bound = s1 + s2 + m1 + m2, dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  return dot3(dif, dif) <= bound * bound;
}
static void mat_vec(real* r, const real* mat, const real* v) {
  for (int i = 0; i < 3; i++) r[i] = mat[3 * i] * v[0] + mat[3 * i + 1] * v[1] + mat[3 * i + 2] * v[2];
}
/* collision_driver.py:116-213 */
static int aabb_filter(const real* c1, const real* c2, const real* s1, const real* s2, real m1, real m2,
                       const real* p1, const real* p2, const real* r1, const real* r2) {
  real cen1[3], cen2[3], mx1[3], mn1[3], mx2[3], mn2[3], margin = m1 + m2;
  mat_vec(cen1, r1, c1); mat_vec(cen2, r2, c2);
  for (int i = 0; i < 3; i++) { cen1[i] += p1[i]; cen2[i] += p2[i]; mx1[i] = mx2[i] = -MJ_MAXVAL; mn1[i] = mn2[i] = MJ_MAXVAL; }
  for (int k = 0; k < 8; k++) {
    real sg[3] = {(k & 4) ? 1.0 : -1.0, (k & 2) ? 1.0 : -1.0, (k & 1) ? 1.0 : -1.0};
    real a[3] = {sg[0] * s1[0], sg[1] * s1[1], sg[2] * s1[2]}, b[3] = {sg[0] * s2[0], sg[1] * s2[1], sg[2] * s2[2]}, q1[3], q2[3];
    mat_vec(q1, r1, a); mat_vec(q2, r2, b);
    for (int i = 0; i < 3; i++) {
      if (q1[i] > mx1[i]) mx1[i] = q1[i];
      if (q1[i] < mn1[i]) mn1[i] = q1[i];
      if (q2[i] > mx2[i]) mx2[i] = q2[i];
      if (q2[i] < mn2[i]) mn2[i] = q2[i];
    }
  }
  for (int i = 0; i < 3; i++) {
    if (cen1[i] + mx1[i] + margin < cen2[i] + mn2[i]) return 0;
    if (cen2[i] + mx2[i] + margin < cen1[i] + mn1[i]) return 0;
  }
  return 1;
}
/* collision_driver.py:218-271 */
static int obb_filter(const real* c1, const real* c2, const real* s1, const real* s2, real m1, real m2,
                      const real* p1, const real* p2, const real* r1, const real* r2) {
  real margin = m1 + m2, xc[2][3], nrm[6][3];
  mat_vec(xc[0], r1, c1); mat_vec(xc[1], r2, c2);
  for (int i = 0; i < 3; i++) { xc[0][i] += p1[i]; xc[1][i] += p2[i]; }
  for (int k = 0; k < 3; k++)
    for (int i = 0; i < 3; i++) { nrm[k][i] = r1[3 * i + k]; nrm[3 + k][i] = r2[3 * i + k]; }
  for (int j = 0; j < 2; j++)
    for (int k = 0; k < 3; k++) {
      real proj[2], rad[2];
      for (int i = 0; i < 2; i++) {
        const real* size = i == 0 ? s1 : s2;
        proj[i] = dot3(xc[i], nrm[3 * j + k]);
        rad[i] = fabs(size[0] * dot3(nrm[3 * i], nrm[3 * j + k])) + fabs(size[1] * dot3(nrm[3 * i + 1], nrm[3 * j + k])) +
                 fabs(size[2] * dot3(nrm[3 * i + 2], nrm[3 * j + k]));
      }
      if (rad[0] + rad[1] + margin < fabs(proj[1] - proj[0])) return 0;
    }
  return 1;
}
/* collision_driver.py:274-321 */
static int broadphase_filter(const OModel* m, const OData* d, int filter, int g1, int g2) {
  const real *a1 = m->geom_aabb + 6 * g1, *a2 = m->geom_aabb + 6 * g2;
  real rb1 = m->geom_rbound[g1], rb2 = m->geom_rbound[g2], m1 = m->geom_margin[g1], m2 = m->geom_margin[g2];
  const real *p1 = d->geom_xpos + 3 * g1, *p2 = d->geom_xpos + 3 * g2, *r1 = d->geom_xmat + 9 * g1, *r2 = d->geom_xmat + 9 * g2;
  if (rb1 == 0.0 || rb2 == 0.0) {
    if (filter & BF_PLANE) return plane_filter(rb1, rb2, m1, m2, p1, p2, r1, r2);
  } else {
    if ((filter & BF_SPHERE) && !sphere_filter(rb1, rb2, m1, m2, p1, p2)) return 0;
    if ((filter & BF_AABB) && !aabb_filter(a1, a2, a1 + 3, a2 + 3, m1, m2, p1, p2, r1, r2)) return 0;
    if ((filter & BF_OBB) && !obb_filter(a1, a2, a1 + 3, a2 + 3, m1, m2, p1, p2, r1, r2)) return 0;
  }
  return 1;
}

typedef struct {
  real dist, pos[3], frame[9];
  int id;
} OContactGeom;

/* collision_primitive_core.py:107-111 */
static void c_plane_sphere(const real* n, const real* ppos, const real* spos, real r, real* dist, real* pos) {
  real dif[3] = {spos[0] - ppos[0], spos[1] - ppos[1], spos[2] - ppos[2]};
  *dist = dot3(dif, n) - r;
  for (int i = 0; i < 3; i++) pos[i] = spos[i] - n[i] * (r + 0.5 * (*dist));
}
/* collision_primitive_core.py:115-143 */
static void c_sphere_sphere(const real* p1, real r1, const real* p2, real r2, real* dist, real* pos, real* n) {
  real dir[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  real dd = norm3(dir);
  if (dd == 0.0) { n[0] = 1; n[1] = 0; n[2] = 0; } else { for (int i = 0; i < 3; i++) n[i] = dir[i] / dd; }
  *dist = dd - (r1 + r2);
  for (int i = 0; i < 3; i++) pos[i] = p1[i] + n[i] * (r1 + 0.5 * (*dist));
}
/* collision_primitive_core.py:41-45 */
static void closest_segment_point(real* r, const real* a, const real* b, const real* pt) {
  real ab[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, pa[3] = {pt[0] - a[0], pt[1] - a[1], pt[2] - a[2]};
  real t = clampr(dot3(pa, ab) / (dot3(ab, ab) + 1e-6), 0.0, 1.0);
  for (int i = 0; i < 3; i++) r[i] = a[i] + t * ab[i];
}

/* helpers of box_box: 3x3 row-major products and the six face re-orientations (collision_primitive_core.py:616-644:
 * a signed permutation that turns face `f` of a box (0..2 = +x,+y,+z, 3..5 = -x,-y,-z) into the +z face) */
static void m3_mul(real* out, const real* a, const real* b) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) out[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
}
static void m3_tr(real* out, const real* a) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) out[3 * i + j] = a[3 * j + i];
}
static void face_rotation(real* m, int f) {
  static const signed char tbl[6][3][2] = {/* per row: (column, sign) */
      {{2, -1}, {1, 1}, {0, 1}}, {{0, 1}, {2, -1}, {1, 1}}, {{0, 1}, {1, 1}, {2, 1}},
      {{2, 1}, {1, 1}, {0, -1}}, {{0, 1}, {2, 1}, {1, -1}}, {{0, -1}, {1, 1}, {2, -1}}};
  memset(m, 0, 9 * sizeof(real));
  for (int i = 0; i < 3; i++) m[3 * i + tbl[f][i][0]] = tbl[f][i][1];
}

/* collision_primitive_core.py:648-1099 box_box (after MuJoCo C's box-box): separating-axis search over the 6 face
 * normals and the 9 edge cross products, then clipping in the frame where the reference face is the +z face:
 * face case -- the incident box's nearest face (or edge) is clipped against the reference rectangle;
 * edge case -- the nearest face of box 2 is projected along the separating axis and clipped likewise.
 * Returns the number of contacts (<= 8); all share one normal. */
static int c_box_box(const real* p1, const real* R1, const real* s1, const real* p2, const real* R2, const real* s2, real margin,
                     real* dist, real* pos /* [8][3] */, real* nrm) {
  real R1t[9], R2t[9], rot21[9], rot12[9], a21[9], a12[9], pos21[3], pos12[3], d21[3], d12[3], plen1[3], plen2[3];
  m3_tr(R1t, R1); m3_tr(R2t, R2);
  for (int i = 0; i < 3; i++) { d21[i] = p2[i] - p1[i]; d12[i] = p1[i] - p2[i]; }
  mat_vec(pos21, R1t, d21); mat_vec(pos12, R2t, d12);
  m3_mul(rot21, R1t, R2); m3_tr(rot12, rot21);
  for (int i = 0; i < 9; i++) { a21[i] = fabs(rot21[i]); a12[i] = fabs(rot12[i]); }
  mat_vec(plen2, a21, s2); mat_vec(plen1, a12, s1);
  real separation = margin + 3.0 * (s1[0] + s2[0]) + 3.0 * (s1[1] + s2[1]) + 3.0 * (s1[2] + s2[2]);
  int code = -1;
  for (int i = 0; i < 3; i++) { /* face normals of both boxes */
    real c1 = -fabs(pos21[i]) + s1[i] + plen2[i], c2 = -fabs(pos12[i]) + s2[i] + plen1[i];
    if (c1 < -margin || c2 < -margin) return 0;
    if (c1 < separation) { separation = c1; code = i + 3 * (pos21[i] < 0); }
    if (c2 < separation) { separation = c2; code = i + 3 * (pos12[i] < 0) + 6; }
  }
  real clnorm[3] = {0, 0, 0};
  int inv = 0, cle1 = 0, cle2 = 0;
  for (int i = 0; i < 3; i++) { /* edge i of box 1 x edge j of box 2 */
    for (int j = 0; j < 3; j++) {
      const real* a = rot12 + 3 * j; /* axis j of box 2 in the frame of box 1 */
      real cr[3];
      if (i == 0) { cr[0] = 0; cr[1] = -a[2]; cr[2] = a[1]; }
      else if (i == 1) { cr[0] = a[2]; cr[1] = 0; cr[2] = -a[0]; }
      else { cr[0] = -a[1]; cr[1] = a[0]; cr[2] = 0; }
      real len = norm3(cr);
      if (len < MJ_MINVAL) continue;
      for (int k = 0; k < 3; k++) cr[k] /= len;
      real bd = dot3(pos21, cr), c3 = 0;
      for (int k = 0; k < 3; k++) {
        if (k != i) c3 += s1[k] * fabs(cr[k]);
        if (k != j) c3 += s2[k] * a21[3 * i + 3 - k - j] / len;
      }
      c3 -= fabs(bd);
      if (c3 < -margin) return 0;
      if (c3 < separation * (1.0 - 1e-12)) {
        separation = c3;
        cle1 = cle2 = 0;
        for (int k = 0; k < 3; k++) {
          if (k != i && ((cr[k] > 0) ^ (bd < 0))) cle1 += 1 << k;
          if (k != j && ((rot21[3 * i + 3 - k - j] > 0) ^ (bd < 0) ^ ((k - j + 3) % 3 == 1))) cle2 += 1 << k;
        }
        code = 12 + 3 * i + j;
        memcpy(clnorm, cr, sizeof(cr));
        inv = bd < 0;
      }
    }
  }
  if (code == -1) return 0;
  real pts[8][3], depth[8], rw[9], pw[3], rotmore[9], rmt[9], hz;
  int n = 0;
  memset(pts, 0, sizeof(pts));
  if (code < 12) { /* a face of box `bi` is the reference face; the other box is incident */
    int bi = code / 6;
    face_rotation(rotmore, code % 6);
    real r[9], rt[9], p[3], ss[3], tmp[3];
    m3_mul(r, rotmore, bi ? rot12 : rot21);
    mat_vec(p, rotmore, bi ? pos12 : pos21);
    mat_vec(tmp, rotmore, bi ? s2 : s1);
    for (int i = 0; i < 3; i++) ss[i] = fabs(tmp[i]);
    const real* s = bi ? s1 : s2;
    m3_tr(rt, r);
    real lx = ss[0], ly = ss[1];
    hz = ss[2];
    p[2] -= hz;
    int clcorner = 0, dirs = 0;
    for (int i = 0; i < 3; i++) if (r[6 + i] < 0) clcorner += 1 << i;
    real lp[3] = {p[0], p[1], p[2]}, cn[2][3] = {{0, 0, 0}, {0, 0, 0}};
    for (int i = 0; i < 3; i++)
      for (int k = 0; k < 3; k++) lp[k] += rt[3 * i + k] * s[i] * ((clcorner & (1 << i)) ? 1.0 : -1.0);
    for (int i = 0; i < 3; i++) {
      if (fabs(r[6 + i]) < 0.5) {
        for (int k = 0; k < 3; k++) cn[dirs ? 1 : 0][k] = rt[3 * i + k] * s[i] * ((clcorner & (1 << i)) ? -2.0 : 2.0);
        dirs++;
      }
    }
    for (int i = 0; i < dirs * dirs; i++) { /* incident edges against the sides of the reference rectangle */
      for (int q = 0; q < 2; q++) {
        real lav[3], lbv[3];
        for (int k = 0; k < 3; k++) {
          lav[k] = lp[k] + (i < 2 ? 0.0 : (i == 2 ? cn[0][k] : cn[1][k]));
          lbv[k] = (i == 0 || i == 3) ? cn[0][k] : cn[1][k];
        }
        if (fabs(lbv[q]) > MJ_MINVAL) {
          real br = 1.0 / lbv[q];
          for (int j = -1; j <= 1; j += 2) {
            real l = ss[q] * j, c1 = (l - lav[q]) * br;
            if (c1 < 0 || c1 > 1) continue;
            real c2 = lav[1 - q] + lbv[1 - q] * c1;
            if (fabs(c2) > ss[1 - q]) continue;
            if (n < 8) { for (int k = 0; k < 3; k++) pts[n][k] = lav[k] + c1 * lbv[k]; n++; }
          }
        }
      }
    }
    if (dirs == 2) { /* corners of the reference rectangle inside the incident face */
      real ax = cn[0][0], bx = cn[1][0], ay = cn[0][1], by = cn[1][1], C = safe_div(1.0, ax * by - bx * ay);
      for (int i = 0; i < 4; i++) {
        real llx = (i / 2) ? lx : -lx, lly = (i % 2) ? ly : -ly, x = llx - lp[0], y = lly - lp[1];
        real u = (x * by - y * bx) * C, v = (y * ax - x * ay) * C;
        if (u > 0 && v > 0 && u < 1 && v < 1 && n < 8) {
          pts[n][0] = llx; pts[n][1] = lly; pts[n][2] = lp[2] + u * cn[0][2] + v * cn[1][2];
          n++;
        }
      }
    }
    for (int i = 0; i < (1 << dirs); i++) { /* incident corners inside the reference rectangle */
      real t[3];
      for (int k = 0; k < 3; k++) t[k] = lp[k] + (real)(i & 1) * cn[0][k] + (real)((i & 2) != 0) * cn[1][k];
      if (t[0] > -lx && t[0] < lx && t[1] > -ly && t[1] < ly && n < 8) { memcpy(pts[n], t, sizeof(t)); n++; }
    }
    int m = n;
    n = 0;
    for (int i = 0; i < m; i++) {
      if (pts[i][2] > margin) continue;
      if (i != n) memcpy(pts[n], pts[i], sizeof(pts[i]));
      depth[n] = pts[n][2];
      pts[n][2] *= 0.5;
      n++;
    }
    m3_tr(rmt, rotmore);
    m3_mul(rw, bi ? R2 : R1, rmt);
    memcpy(pw, bi ? p2 : p1, sizeof(pw));
    for (int k = 0; k < 3; k++) nrm[k] = (bi ? -1.0 : 1.0) * rw[3 * k + 2];
  } else { /* edge e1 of box 1 against edge e2 of box 2 */
    int e1 = (code - 12) / 3, e2 = (code - 12) % 3;
    int ax1 = 1 - (e2 & 1), ax2 = 2 - (e2 & 2), pax1 = 1 - (e1 & 1), pax2 = 2 - (e1 & 2);
    if (a21[3 * e1 + ax1] < a21[3 * e1 + ax2]) { int t = ax1; ax1 = ax2; ax2 = t; }
    if (a12[3 * e2 + pax1] < a12[3 * e2 + pax2]) { int t = pax1; pax1 = pax2; pax2 = t; }
    (void)pax1;
    face_rotation(rotmore, (cle1 & (1 << pax2)) ? pax2 : pax2 + 3);
    m3_tr(rmt, rotmore);
    real p[3], rn[3], r[9], rt[9], s[3], tmp[3];
    mat_vec(p, rotmore, pos21); mat_vec(rn, rotmore, clnorm); m3_mul(r, rotmore, rot21); m3_tr(rt, r);
    mat_vec(tmp, rmt, s1);
    for (int i = 0; i < 3; i++) s[i] = fabs(tmp[i]);
    real lx = s[0], ly = s[1];
    hz = s[2];
    p[2] -= hz;
    real sg1 = (cle2 & (1 << ax1)) ? 1.0 : -1.0, sg2 = (cle2 & (1 << ax2)) ? 1.0 : -1.0, pu[4][3];
    for (int k = 0; k < 3; k++) { /* the face of box 2 nearest to box 1: two edges parallel to e2 */
      real e = rt[3 * e2 + k] * s2[e2];
      real q0 = p[k] + rt[3 * ax1 + k] * s2[ax1] * sg1 + rt[3 * ax2 + k] * s2[ax2] * sg2;
      real q2 = p[k] - rt[3 * ax1 + k] * s2[ax1] * sg1 + rt[3 * ax2 + k] * s2[ax2] * sg2;
      pts[0][k] = q0 + e; pts[1][k] = q0 - e; pts[2][k] = q2 + e; pts[3][k] = q2 - e;
    }
    real alp[3], acn1[3], acn2[3];
    for (int k = 0; k < 3; k++) { alp[k] = pts[0][k]; acn1[k] = pts[1][k] - pts[0][k]; acn2[k] = pts[2][k] - pts[0][k]; }
    if (fabs(rn[2]) < MJ_MINVAL) return 0;
    real sgn = inv ? -1.0 : 1.0, innorm = sgn / rn[2];
    for (int i = 0; i < 4; i++) { /* project along the separating axis onto the reference plane */
      memcpy(pu[i], pts[i], sizeof(pts[i]));
      real sc = pts[i][2] * sgn * innorm;
      for (int k = 0; k < 3; k++) pts[i][k] -= rn[k] * sc;
    }
    real plp[3], pcn1[3], pcn2[3];
    for (int k = 0; k < 3; k++) { plp[k] = pts[0][k]; pcn1[k] = pts[1][k] - pts[0][k]; pcn2[k] = pts[2][k] - pts[0][k]; }
    for (int i = 0; i < 4; i++) {
      for (int q = 0; q < 2; q++) {
        real la = plp[q] + (i < 2 ? 0.0 : (i == 2 ? pcn1[q] : pcn2[q])), lb = (i == 0 || i == 3) ? pcn1[q] : pcn2[q];
        real lc = plp[1 - q] + (i < 2 ? 0.0 : (i == 2 ? pcn1[1 - q] : pcn2[1 - q])), ld = (i == 0 || i == 3) ? pcn1[1 - q] : pcn2[1 - q];
        real lua[3], lub[3];
        for (int k = 0; k < 3; k++) {
          lua[k] = alp[k] + (i < 2 ? 0.0 : (i == 2 ? acn1[k] : acn2[k]));
          lub[k] = (i == 0 || i == 3) ? acn1[k] : acn2[k];
        }
        if (fabs(lb) > MJ_MINVAL) {
          real br = 1.0 / lb;
          for (int j = -1; j <= 1; j += 2) {
            if (n == 8) break;
            real l = s[q] * j, c1 = (l - la) * br;
            if (c1 < 0 || c1 > 1) continue;
            real c2 = lc + ld * c1;
            if (fabs(c2) > s[1 - q]) continue;
            if ((lua[2] + lub[2] * c1) * innorm > margin) continue;
            for (int k = 0; k < 3; k++) pts[n][k] = lua[k] * 0.5 + c1 * lub[k] * 0.5;
            pts[n][q] += 0.5 * l;
            pts[n][1 - q] += 0.5 * c2;
            depth[n] = pts[n][2] * innorm * 2.0;
            n++;
          }
        }
      }
    }
    int nl = n;
    real ax = pcn1[0], bx = pcn2[0], ay = pcn1[1], by = pcn2[1], C = safe_div(1.0, ax * by - bx * ay);
    for (int i = 0; i < 4; i++) { /* corners of the reference rectangle under the projected face */
      if (n == 8) break;
      real llx = (i / 2) ? lx : -lx, lly = (i % 2) ? ly : -ly, x = llx - plp[0], y = lly - plp[1];
      real u = (x * by - y * bx) * C, v = (y * ax - x * ay) * C;
      if (nl == 0) { if ((u < 0 || u > 1) && (v < 0 || v > 1)) continue; }
      else if (u < 0 || v < 0 || u > 1 || v > 1) continue;
      u = clampr(u, 0.0, 1.0); v = clampr(v, 0.0, 1.0);
      real w = 1.0 - u - v, vt[3], tc1 = 0, corner[3] = {llx, lly, 0.0};
      for (int k = 0; k < 3; k++) { vt[k] = pu[0][k] * w + pu[1][k] * u + pu[2][k] * v; tc1 += (corner[k] - vt[k]) * (corner[k] - vt[k]); }
      if (vt[2] > 0 && tc1 > margin * margin) continue;
      for (int k = 0; k < 3; k++) pts[n][k] = 0.5 * (corner[k] + vt[k]);
      depth[n] = sqrt(tc1) * (vt[2] < 0 ? -1.0 : 1.0);
      n++;
    }
    int nf = n;
    for (int i = 0; i < 4; i++) { /* corners of the face of box 2 over the reference rectangle */
      if (n >= 8) break;
      real x = pu[i][0], y = pu[i][1];
      if (nl == 0 && nf != 0) { if ((x < -lx || x > lx) && (y < -ly || y > ly)) continue; }
      else if (x < -lx || x > lx || y < -ly || y > ly) continue;
      real c1 = 0, tp[3] = {pu[i][0], pu[i][1], 0.0};
      for (int j = 0; j < 2; j++) {
        if (pu[i][j] < -s[j]) { c1 += (pu[i][j] + s[j]) * (pu[i][j] + s[j]); tp[j] = -s[j] * 0.5; }
        else if (pu[i][j] > s[j]) { c1 += (pu[i][j] - s[j]) * (pu[i][j] - s[j]); tp[j] = s[j] * 0.5; }
      }
      c1 += pu[i][2] * innorm * pu[i][2] * innorm;
      if (pu[i][2] > 0 && c1 > margin * margin) continue;
      for (int k = 0; k < 3; k++) pts[n][k] = (tp[k] + pu[i][k]) * 0.5;
      depth[n] = sqrt(c1) * (pu[i][2] < 0 ? -1.0 : 1.0);
      n++;
    }
    m3_mul(rw, R1, rmt);
    memcpy(pw, p1, sizeof(pw));
    real t[3];
    mat_vec(t, rw, rn);
    for (int k = 0; k < 3; k++) nrm[k] = sgn * t[k];
  }
  for (int i = 0; i < n; i++) {
    pts[i][2] += hz;
    mat_vec(pos + 3 * i, rw, pts[i]);
    for (int k = 0; k < 3; k++) pos[3 * i + k] += pw[k];
    dist[i] = depth[i];
  }
  return n;
}

/* collision_primitive_core.py:1103-1154.  Sphere (geom 1) against box (geom 2): the centre is taken into the box
 * frame and clamped to the box; outside, the contact lies between the clamped point and the deepest sphere point;
 * inside, the nearest face decides (faces scanned in the order -x,+x,-y,+y,-z,+z, strict improvement). */
static void c_sphere_box(const real* sp, real r, const real* bp, const real* br, const real* bs, real* dist, real* pos, real* nrm) {
  real c[3], q[3], dl[3], nl[3] = {0, 0, 0}, loc[3], len;
  for (int i = 0; i < 3; i++) c[i] = br[i] * (sp[0] - bp[0]) + br[3 + i] * (sp[1] - bp[1]) + br[6 + i] * (sp[2] - bp[2]);
  for (int i = 0; i < 3; i++) { q[i] = clampr(c[i], -bs[i], bs[i]); dl[i] = q[i] - c[i]; }
  len = norm3(dl);
  if (len <= MJ_MINVAL) {
    real closest = 2.0 * (bs[0] + bs[1] + bs[2]);
    int k = 0;
    for (int i = 0; i < 6; i++) {
      real fd = fabs(((i & 1) ? 1.0 : -1.0) * bs[i / 2] - c[i / 2]);
      if (closest > fd) { closest = fd; k = i; }
    }
    nl[k / 2] = (k & 1) ? -1.0 : 1.0;
    for (int i = 0; i < 3; i++) loc[i] = c[i] + nl[i] * (r - closest) / 2.0;
    *dist = -closest - r;
  } else {
    for (int i = 0; i < 3; i++) { nl[i] = dl[i] / len; loc[i] = 0.5 * (q[i] + c[i] + nl[i] * r); }
    *dist = len - r;
  }
  mat_vec(nrm, br, nl);
  mat_vec(pos, br, loc);
  for (int i = 0; i < 3; i++) pos[i] += bp[i];
}

/* collision_primitive_core.py:1158-1491 (after MuJoCo C's capsule-box).  Finds the closest feature pair between the
 * capsule segment and the box (segment end vs face, then segment vs each of the 12 edges), derives from the discrete
 * configuration where a second support point along the segment lies, and collides a sphere of the capsule's radius
 * at both segment points with the box.  dist[1] = MJ_MAXVAL when there is no second point. */
static void c_capsule_box(const real* cp, const real* caxis, real r, real hl, const real* bp, const real* br, const real* bs,
                          real* dist, real* pos, real* nrm) {
  real pl[3], ax[3], half[3];
  for (int i = 0; i < 3; i++) {
    pl[i] = br[i] * (cp[0] - bp[0]) + br[3 + i] * (cp[1] - bp[1]) + br[6 + i] * (cp[2] - bp[2]);
    ax[i] = br[i] * caxis[0] + br[3 + i] * caxis[1] + br[6 + i] * caxis[2];
    half[i] = ax[i] * hl;
  }
  const int axisdir = (half[0] > 0.0) + 2 * (half[1] > 0.0) + 4 * (half[2] > 0.0);
  real best = 1.0e32, bestseg = -12.0, bestbox = 0.0;
  int cltype = -4, clface = -12, clcorner = -123, cledge = -123;
  dist[0] = dist[1] = MJ_MAXVAL;
  memset(pos, 0, 6 * sizeof(real)); memset(nrm, 0, 6 * sizeof(real));
  for (int sg = -1; sg <= 1; sg += 2) { /* a segment end closest to a face (or inside the box) */
    real tip[3], dd = 0.0;
    int nout = 0, axout = -1;
    for (int j = 0; j < 3; j++) {
      tip[j] = pl[j] + sg * half[j];
      real cj = tip[j];
      if (cj < -bs[j]) { nout++; axout = j; cj = -bs[j]; } else if (cj > bs[j]) { nout++; axout = j; cj = bs[j]; }
      dd += (cj - tip[j]) * (cj - tip[j]);
    }
    if (nout > 1) continue;
    if (dd < best) { best = dd; bestseg = sg; cltype = -2 + sg; clface = axout; }
  }
  for (int i = 0; i < 8; i++) { /* segment against the edge that leaves corner i along axis j */
    for (int j = 0; j < 3; j++) {
      if (i & (1 << j)) continue;
      real e[3] = {(i & 1) ? bs[0] : -bs[0], (i & 2) ? bs[1] : -bs[1], (i & 4) ? bs[2] : -bs[2]}, dif[3];
      e[j] = 0.0;
      for (int k = 0; k < 3; k++) dif[k] = e[k] - pl[k];
      real u = -bs[j] * dif[j], v = dot3(half, dif), ma = bs[j] * bs[j], mb = -bs[j] * half[j], mc = hl * hl;
      real det = ma * mc - mb * mb;
      if (fabs(det) < MJ_MINVAL) continue;
      real x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
      int s1 = 1, s2 = 1;
      if (x1 > 1.0) { x1 = 1.0; s1 = 2; x2 = safe_div(v - mb, mc); }
      else if (x1 < -1.0) { x1 = -1.0; s1 = 0; x2 = safe_div(v + mb, mc); }
      if (x2 > 1.0 || x2 < -1.0) {
        if (x2 > 1.0) { x2 = 1.0; s2 = 2; x1 = safe_div(u - mb, ma); } else { x2 = -1.0; s2 = 0; x1 = safe_div(u + mb, ma); }
        if (x1 > 1.0) { x1 = 1.0; s1 = 2; } else if (x1 < -1.0) { x1 = -1.0; s1 = 0; }
      }
      for (int k = 0; k < 3; k++) dif[k] -= half[k] * x2;
      dif[j] += bs[j] * x1;
      real dd = dot3(dif, dif);
      if (dd < best - MJ_MINVAL) {
        best = dd; bestseg = x2; bestbox = x1;
        cltype = s1 * 3 + s2;
        clcorner = i + (1 << j) * (cltype / 6);
        cledge = j;
      }
    }
  }
  if (cltype == -4) return;
  real second = -4.0;
  if (cltype >= 0 && cltype / 3 != 1) { /* a box corner is closest */
    int c1 = axisdir ^ clcorner;
    if (c1 != 0 && c1 != 7) {
      int mul = 1;
      if (!(c1 == 1 || c1 == 2 || c1 == 4)) { mul = -1; c1 = 7 - c1; }
      int a = c1 == 1 ? 0 : (c1 == 2 ? 1 : 2), a1 = (a + 1) % 3, a2 = (a + 2) % 3;
      if (ax[a] * ax[a] > 0.5) {
        real mm = 2.0 * safe_div(bs[a], fabs(half[a]));
        second = fmin(1.0 - mul * bestseg, mm);
      } else {
        real mm = 2.0 * fmin(safe_div(bs[a1], fabs(half[a1])), safe_div(bs[a2], fabs(half[a2])));
        second = -fmin(1.0 + mul * bestseg, mm);
      }
      second *= mul;
    }
  } else if (cltype >= 0) { /* the interior of a box edge is closest */
    int c1 = (axisdir ^ clcorner) & (7 - (1 << cledge));
    if (c1 == 1 || c1 == 2 || c1 == 4) {
      int a = cledge, a1 = (a + 1) % 3, a2 = (a + 2) % 3, mul;
      if (fabs(ax[a1]) > fabs(ax[a2])) a1 = a2;
      a2 = 3 - a - a1;
      if (c1 & (1 << a2)) { mul = 1; second = 1.0 - bestseg; } else { mul = -1; second = 1.0 + bestseg; }
      second = fmin(2.0 * safe_div(bs[a2], fabs(half[a2])), second);
      real e2 = (((axisdir >> a) & 1) == ((c1 >> a2) & 1)) ? 1.0 - bestbox : 1.0 + bestbox;
      second = fmin(bs[a] * safe_div(e2, fabs(half[a])), second);
      second *= mul;
    }
  } else if (clface != -1) { /* a segment end is closest to a face, from outside */
    int mul = cltype == -3 ? 1 : -1;
    real t1[3];
    second = 2.0;
    for (int i = 0; i < 3; i++) t1[i] = pl[i] - half[i] * mul;
    for (int i = 0; i < 3; i++) {
      if (i == clface) continue;
      real har = safe_div((real)mul, half[i]);
      real e1 = (bs[i] - t1[i]) * har;
      if (0.0 < e1 && e1 < second) second = e1;
      e1 = (-bs[i] - t1[i]) * har;
      if (0.0 < e1 && e1 < second) second = e1;
    }
    second *= mul;
  }
  for (int k = 0; k < 2; k++) {
    if (k == 1 && !(second > -3.0)) break;
    real t = k == 0 ? bestseg : second + bestseg, loc[3], glob[3];
    for (int i = 0; i < 3; i++) loc[i] = pl[i] + half[i] * t;
    mat_vec(glob, br, loc);
    for (int i = 0; i < 3; i++) glob[i] += bp[i];
    c_sphere_box(glob, r, bp, br, bs, dist + k, pos + 3 * k, nrm + 3 * k);
  }
}

/* Computes the contact geometries of one geom pair; returns the number of candidates
 * (collision_primitive.py wrappers :280-1280 + cores).  Candidates with dist >= margin are
 * dropped by the caller (collision_core.py:199-203). */
/* collision_driver.py:43 MJ_COLLISION_TABLE: pair types routed to the convex (GJK / EPA) narrowphase.  box-box stays on the
 * primitive only under DisableBit.NATIVECCD (the reference's default sends it to ccd + multicontact, :74, :736; see narrowphase_pair). */
static int is_convex_pair(int t1, int t2) {
  if (t1 == GEOM_SPHERE || t1 == GEOM_CAPSULE) return t2 == GEOM_ELLIPSOID || (t1 == GEOM_CAPSULE && t2 == GEOM_CYLINDER);
  if (t1 == GEOM_ELLIPSOID) return t2 == GEOM_ELLIPSOID || t2 == GEOM_CYLINDER || t2 == GEOM_BOX;
  if (t1 == GEOM_CYLINDER) return t2 == GEOM_CYLINDER || t2 == GEOM_BOX;
  return 0;
}
static int is_convex_pair_mesh(int t1, int t2) { return (t2 == GEOM_MESH && t1 >= GEOM_SPHERE) || is_convex_pair(t1, t2); }

/* collision_convex.py:712 eval_ccd_write_contact (single contact, no collision sensor): returns 1 candidate whose dist is the
 * surface distance (>= margin when the margin-inflated shapes do not overlap, which write_contact then rejects) */
static int ccd_pair(const real* p1, const real* r1, const real* s1, int t1, const real* v1, int nv1, const int* gr1, const real* p2,
                    const real* r2, const real* s2, int t2, const real* v2, int nv2, const int* gr2, real margin, real tolerance,
                    int iterations, OContactGeom* out) {
  CcdGeom g1, g2;
  g1.vert = v1; g1.nvert = nv1; g2.vert = v2; g2.nvert = nv2;
  g1.graph = gr1; g2.graph = gr2;
  static _Thread_local CcdPolytope pt;
  CcdGjk res;
  memcpy(g1.pos, p1, sizeof(g1.pos)); memcpy(g1.rot, r1, sizeof(g1.rot)); memcpy(g1.size, s1, sizeof(g1.size));
  memcpy(g2.pos, p2, sizeof(g2.pos)); memcpy(g2.rot, r2, sizeof(g2.rot)); memcpy(g2.size, s2, sizeof(g2.size));
  g1.margin = g2.margin = margin; g1.type = t1; g2.type = t2;
  if (iterations > CCD_MAX_ITER) iterations = CCD_MAX_ITER;
  real dist, x1[3], x2[3], n[3];
  int face_idx = -1;
  int nc = ccd_run(tolerance, 0.0, iterations, iterations, g1, g2, &pt, &res, &dist, x1, x2, &face_idx);
  const int ncand = (t1 == GEOM_BOX && t2 == GEOM_BOX) ? 4 : 1; /* box-box: up to 4 contacts by face clipping (collision_gjk.py:1930) */
  for (int k = 0; k < ncand; k++) { out[k].id = k; out[k].dist = MJ_MAXVAL; memset(out[k].pos, 0, sizeof(out[k].pos)); memset(out[k].frame, 0, sizeof(out[k].frame)); }
  if (nc == 0 || dist >= 0.0) return ncand;
  real w1[4][3], w2[4][3];
  memcpy(w1[0], x1, sizeof(x1)); memcpy(w2[0], x2, sizeof(x2));
  int ncon = 1;
  if (ncand == 4 && face_idx > -1) { /* collision_convex.py:809-846 */
    static _Thread_local CcdMulti mc;
    ncon = ccd_multicontact_box(&mc, &pt, pt.face[face_idx], &g1, &g2, w1, w2);
  }
  for (int i = 0; i < 3; i++) n[i] = w1[0][i] - w2[0][i];
  for (int k = 0; k < ncon; k++) {
    out[k].dist = dist + margin;
    for (int i = 0; i < 3; i++) out[k].pos[i] = 0.5 * (w1[k][i] + w2[k][i]);
    make_frame(out[k].frame, n);
  }
  return ncand;
}

/* collision_convex.py:56 _hfield_filter + :158 ccd_hfield_kernel (see MuJoCo mjc_ConvexHField): the geom is brought into the height
 * field's frame, the grid cells under its bounding box are walked, every cell contributes two triangular prisms (bottom at
 * -z_bottom, top at the elevations + margin) that run GJK / EPA against the geom, and up to four of the resulting contacts are kept:
 * the deepest, the one furthest from it, the one furthest from their line, the one furthest from the other triangle edge. */
#define MJ_MAXCONPAIR 50
static int hfield_pair(const OModel* m, const OData* d, int g1, int g2, real margin, real tolerance, int iterations, OContactGeom* out) {
  for (int k = 0; k < 4; k++) { out[k].id = k; out[k].dist = MJ_MAXVAL; memset(out[k].pos, 0, sizeof(out[k].pos)); memset(out[k].frame, 0, sizeof(out[k].frame)); }
  const int hid = m->geom_dataid[g1];
  const real* size1 = m->hfield_size + 4 * hid;
  const real *pos1 = d->geom_xpos + 3 * g1, *mat1 = d->geom_xmat + 9 * g1, *pos2 = d->geom_xpos + 3 * g2, *mat2 = d->geom_xmat + 9 * g2;
  real dp[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]}, pos[3];
  for (int i = 0; i < 3; i++) pos[i] = mat1[i] * dp[0] + mat1[3 + i] * dp[1] + mat1[6 + i] * dp[2];
  const real r2 = m->geom_rbound[g2], fm = m->geom_margin[g1] + m->geom_margin[g2];
  for (int i = 0; i < 2; i++)
    if (size1[i] < pos[i] - r2 - fm || -size1[i] > pos[i] + r2 + fm) return 4;
  if (size1[2] < pos[2] - r2 - fm) return 4;
  if (-size1[3] > pos[2] + r2 + fm) return 4;
  CcdGeom ga, gb;
  memset(&ga, 0, sizeof(ga)); memset(&gb, 0, sizeof(gb));
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) gb.rot[3 * i + j] = mat1[i] * mat2[j] + mat1[3 + i] * mat2[3 + j] + mat1[6 + i] * mat2[6 + j];
  memcpy(gb.pos, pos, sizeof(pos));
  memcpy(gb.size, m->geom_size + 3 * g2, 3 * sizeof(real));
  gb.margin = 0.0; gb.type = m->geom_type[g2];
  if (gb.type == GEOM_MESH) {
    int id = m->geom_dataid[g2];
    gb.vert = m->mesh_vert + 3 * m->mesh_vertadr[id];
    gb.nvert = m->mesh_vertnum[id];
    gb.graph = m->mesh_graphadr[id] >= 0 ? m->mesh_graph + m->mesh_graphadr[id] : 0;
  }
  real ext[6]; /* xmax xmin ymax ymin zmax zmin by support queries */
  for (int a = 0; a < 3; a++)
    for (int sgn = 0; sgn < 2; sgn++) {
      real dir[3] = {0, 0, 0}, pnt[3];
      int cache = -1;
      dir[a] = sgn ? -1.0 : 1.0;
      ccd_support(&gb, dir, pnt, &cache);
      ext[2 * a + sgn] = pnt[a];
    }
  const real xmax = ext[0], xmin = ext[1], ymax = ext[2], ymin = ext[3], zmax = ext[4], zmin = ext[5];
  if (xmin - fm > size1[0] || xmax + fm < -size1[0] || ymin - fm > size1[1] || ymax + fm < -size1[1] || zmin - fm > size1[2] || zmax + fm < -size1[3])
    return 4;
  ga.type = GEOM_HFIELD; ga.margin = 0.0;
  ga.rot[0] = ga.rot[4] = ga.rot[8] = 1.0;
  const int nrow = m->hfield_nrow[hid], ncol = m->hfield_ncol[hid], adr = m->hfield_adr[hid];
  const real x_scale = 0.5 * (real)(ncol - 1) / size1[0], y_scale = 0.5 * (real)(nrow - 1) / size1[1];
  int cmin = (int)floor((xmin + size1[0]) * x_scale), cmax = (int)ceil((xmax + size1[0]) * x_scale);
  int rmin = (int)floor((ymin + size1[1]) * y_scale), rmax = (int)ceil((ymax + size1[1]) * y_scale);
  if (cmin < 0) cmin = 0;
  if (cmax > ncol - 1) cmax = ncol - 1;
  if (rmin < 0) rmin = 0;
  if (rmax > nrow - 1) rmax = nrow - 1;
  const real dx = 2.0 * size1[0] / (real)(ncol - 1), dy = 2.0 * size1[1] / (real)(nrow - 1);
  const int dr[2] = {1, 0};
  real prism[6][3];
  memset(prism, 0, sizeof(prism));
  prism[0][2] = prism[1][2] = prism[2][2] = -size1[3];
  real cdist[MJ_MAXCONPAIR], cpos[MJ_MAXCONPAIR][3], cnrm[MJ_MAXCONPAIR][3];
  real min_dist = MJ_MAXVAL, min_normal[3] = {MJ_MAXVAL, MJ_MAXVAL, MJ_MAXVAL}, min_pos[3] = {MJ_MAXVAL, MJ_MAXVAL, MJ_MAXVAL};
  int min_id = -1, count = 0;
  gb.margin = margin;
  static _Thread_local CcdPolytope pt;
  CcdGjk res;
  if (iterations > CCD_MAX_ITER) iterations = CCD_MAX_ITER;
#define HF_ADD_VERT(cc, rr)                                                      \
  do {                                                                           \
    const real x_ = dx * (real)(cc) - size1[0], y_ = dy * (real)(rr) - size1[1]; \
    const real z_ = m->hfield_data[adr + (rr) * ncol + (cc)] * size1[2] + margin; \
    memcpy(prism[0], prism[1], sizeof(prism[0])); memcpy(prism[1], prism[2], sizeof(prism[0])); \
    memcpy(prism[3], prism[4], sizeof(prism[0])); memcpy(prism[4], prism[5], sizeof(prism[0])); \
    prism[2][0] = x_; prism[5][0] = x_; prism[2][1] = y_; prism[5][1] = y_; prism[5][2] = z_; \
  } while (0)
  for (int r = rmin; r < rmax; r++) {
    for (int i = 0; i < 2; i++) HF_ADD_VERT(cmin, r + dr[i]);
    for (int c = cmin + 1; c <= cmax; c++) {
      for (int i = 0; i < 2; i++) {
        if (count >= MJ_MAXCONPAIR) continue; /* the reference reports an overflow and skips (without advancing the prism) */
        HF_ADD_VERT(c, r + dr[i]);
        if (prism[3][2] < zmin && prism[4][2] < zmin && prism[5][2] < zmin) continue;
        memcpy(ga.prism, prism, sizeof(prism));
        for (int k = 0; k < 3; k++) ga.pos[k] = (prism[0][k] + prism[1][k] + prism[2][k] + prism[3][k] + prism[4][k] + prism[5][k]) * (1.0 / 6.0);
        real dist, w1[3], w2[3];
        int face_idx;
        int nc = ccd_run(tolerance, 0.0, iterations, iterations, ga, gb, &pt, &res, &dist, w1, w2, &face_idx);
        if (nc == 0) continue;
        cdist[count] = dist;
        real pl[3] = {0.5 * (w1[0] + w2[0]), 0.5 * (w1[1] + w2[1]), 0.5 * (w1[2] + w2[2])}, nl[3] = {w1[0] - w2[0], w1[1] - w2[1], w1[2] - w2[2]}, fr[9];
        mat_vec(cpos[count], mat1, pl);
        for (int k = 0; k < 3; k++) cpos[count][k] += pos1[k];
        make_frame(fr, nl);
        mat_vec(cnrm[count], mat1, fr);
        if (dist < min_dist) { min_dist = dist; memcpy(min_normal, cnrm[count], sizeof(min_normal)); memcpy(min_pos, cpos[count], sizeof(min_pos)); min_id = count; }
        count++;
      }
    }
  }
#undef HF_ADD_VERT
#define HF_WRITE(k, dd, pp, nn) do { out[k].dist = (dd); memcpy(out[k].pos, (pp), 3 * sizeof(real)); make_frame(out[k].frame, (nn)); } while (0)
  if (min_id < 0) return 4;
  HF_WRITE(0, min_dist, min_pos, min_normal);
  const real MIN_NEXT = 1.0e-3;
  int id1 = -1, id2 = -1, id3 = -1;
  real dist1 = -MJ_MAXVAL, dist12 = -MJ_MAXVAL, dist3 = -MJ_MAXVAL, v[3];
  for (int i = 0; i < count; i++) {
    if (i == min_id) continue;
    for (int k = 0; k < 3; k++) v[k] = cpos[i][k] - min_pos[k];
    real dd = norm3(v);
    if (dd > dist1) { id1 = i; dist1 = dd; }
  }
  if (id1 == -1 || (0.0 < dist1 && dist1 < MIN_NEXT)) return 4;
  HF_WRITE(1, cdist[id1], cpos[id1], cnrm[id1]);
  real d_min1[3], t3[3];
  for (int k = 0; k < 3; k++) t3[k] = min_pos[k] - cpos[id1][k];
  cross3(d_min1, min_normal, t3);
  for (int i = 0; i < count; i++) {
    if (i == min_id || i == id1) continue;
    for (int k = 0; k < 3; k++) v[k] = cpos[i][k] - min_pos[k];
    real dd = fabs(dot3(v, d_min1));
    if (dd > dist12) { id2 = i; dist12 = dd; }
  }
  if (id2 == -1 || (0.0 < dist12 && dist12 < MIN_NEXT)) return 4;
  HF_WRITE(2, cdist[id2], cpos[id2], cnrm[id2]);
  real vec_min2[3], vec_12[3], u3[3];
  for (int k = 0; k < 3; k++) { t3[k] = min_pos[k] - cpos[id2][k]; u3[k] = cpos[id1][k] - cpos[id2][k]; }
  cross3(vec_min2, min_normal, t3);
  cross3(vec_12, min_normal, u3);
  for (int i = 0; i < count; i++) {
    if (i == min_id || i == id1 || i == id2) continue;
    for (int k = 0; k < 3; k++) { v[k] = cpos[i][k] - min_pos[k]; u3[k] = cpos[id1][k] - cpos[i][k]; }
    real dd = fabs(dot3(v, vec_min2)) + fabs(dot3(u3, vec_12));
    if (dd > dist3) { id3 = i; dist3 = dd; }
  }
  if (id3 == -1 || (0.0 < dist3 && dist3 < MIN_NEXT)) return 4;
  HF_WRITE(3, cdist[id3], cpos[id3], cnrm[id3]);
#undef HF_WRITE
  return 4;
}

static int narrowphase_pair(const OModel* m, const OData* d, int g1, int g2, real margin, real ccd_tolerance, int ccd_iterations,
                            int nativeccd, OContactGeom* out) {
  int t1 = m->geom_type[g1], t2 = m->geom_type[g2];
  const real *p1 = d->geom_xpos + 3 * g1, *p2 = d->geom_xpos + 3 * g2, *r1 = d->geom_xmat + 9 * g1, *r2 = d->geom_xmat + 9 * g2;
  const real *s1 = m->geom_size + 3 * g1, *s2 = m->geom_size + 3 * g2;
  real n1[3] = {r1[2], r1[5], r1[8]}, n2[3] = {r2[2], r2[5], r2[8]};
  if (t1 == GEOM_HFIELD) return hfield_pair(m, d, g1, g2, margin, ccd_tolerance, ccd_iterations, out);
  /* collision_driver.py:74, :736: box-box goes through ccd + multicontact unless DisableBit.NATIVECCD */
  if (is_convex_pair_mesh(t1, t2) || (nativeccd && t1 == GEOM_BOX && t2 == GEOM_BOX)) {
    const real *v1 = 0, *v2 = 0;
    const int *gr1 = 0, *gr2 = 0;
    int nv1 = 0, nv2 = 0;
    if (t1 == GEOM_MESH) {
      int id = m->geom_dataid[g1];
      v1 = m->mesh_vert + 3 * m->mesh_vertadr[id];
      nv1 = m->mesh_vertnum[id];
      if (m->mesh_graphadr[id] >= 0) gr1 = m->mesh_graph + m->mesh_graphadr[id];
    }
    if (t2 == GEOM_MESH) {
      int id = m->geom_dataid[g2];
      v2 = m->mesh_vert + 3 * m->mesh_vertadr[id];
      nv2 = m->mesh_vertnum[id];
      if (m->mesh_graphadr[id] >= 0) gr2 = m->mesh_graph + m->mesh_graphadr[id];
    }
    return ccd_pair(p1, r1, s1, t1, v1, nv1, gr1, p2, r2, s2, t2, v2, nv2, gr2, margin, ccd_tolerance, ccd_iterations, out);
  }
  if (t1 == GEOM_PLANE && t2 == GEOM_MESH) { /* collision_primitive.py:52-277 plane_convex, exhaustive branch (no mesh graph) */
    int id = m->geom_dataid[g2], nvert = m->mesh_vertnum[id], idx[4] = {-1, -1, -1, -1};
    const real* vert = m->mesh_vert + 3 * m->mesh_vertadr[id];
    real dp[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]}, pl[3], n[3], HUGE_V = 1e6;
    for (int i = 0; i < 3; i++) { pl[i] = r2[i] * dp[0] + r2[3 + i] * dp[1] + r2[6 + i] * dp[2]; n[i] = r2[i] * n1[0] + r2[3 + i] * n1[1] + r2[6 + i] * n1[2]; }
    for (int k = 0; k < 4; k++) { out[k].dist = MJ_MAXVAL; memset(out[k].pos, 0, sizeof(out[k].pos)); make_frame(out[k].frame, n1); out[k].id = k; }
#define SUPPORT(i) ((pl[0] - vert[3 * (i)]) * n[0] + (pl[1] - vert[3 * (i) + 1]) * n[1] + (pl[2] - vert[3 * (i) + 2]) * n[2])
    real max_support = -HUGE_V, a[3] = {0, 0, 0}, b[3] = {0, 0, 0}, c[3] = {0, 0, 0};
    if (m->mesh_graphadr[id] >= 0 && nvert >= 10) { /* collision_primitive.py:140-255: the same four searches by hill climbing */
      const int* graph = m->mesh_graph + m->mesh_graphadr[id];
      const int numvert = graph[0], *vert_edgeadr = graph + 2, *vert_globalid = graph + 2 + numvert, *edge_localid = graph + 2 + 2 * numvert;
      int imax = 0;
      real threshold = 0, best = -HUGE_V, ab[3] = {0, 0, 0}, ac[3] = {0, 0, 0}, bc[3] = {0, 0, 0}, t[3];
      for (int phase = 0; phase < 5; phase++) { /* 0: deepest vertex, 1: a (deepest above threshold), 2: b, 3: c, 4: d */
        if (phase > 0) best = -HUGE_V;
        for (;;) {
          int prev = imax, i = vert_edgeadr[imax];
          while (edge_localid[i] >= 0) {
            int subidx = edge_localid[i], vi = vert_globalid[subidx];
            real sp = SUPPORT(vi), mask = sp > threshold ? 0.0 : -HUGE_V, val;
            const real* v = vert + 3 * vi;
            if (phase == 0) val = sp;
            else if (phase == 1) val = sp > threshold ? sp : -HUGE_V;
            else if (phase == 2) { for (int j = 0; j < 3; j++) t[j] = a[j] - v[j]; val = dot3(t, t) + mask; }
            else if (phase == 3) { for (int j = 0; j < 3; j++) t[j] = a[j] - v[j]; val = fabs(dot3(t, ab)) + mask; }
            else {
              real bp[3];
              for (int j = 0; j < 3; j++) { t[j] = a[j] - v[j]; bp[j] = b[j] - v[j]; }
              val = (fabs(dot3(t, ac)) + mask) + (fabs(dot3(bp, bc)) + mask);
            }
            if (val > best) { best = val; imax = subidx; }
            i++;
          }
          if (imax == prev) break;
        }
        int gi = vert_globalid[imax];
        if (phase == 0) { max_support = best; threshold = fmax(0.0, max_support - 1e-3); }
        else if (phase == 1) { idx[0] = gi; memcpy(a, vert + 3 * gi, sizeof(a)); }
        else if (phase == 2) {
          idx[1] = gi; memcpy(b, vert + 3 * gi, sizeof(b));
          for (int j = 0; j < 3; j++) t[j] = a[j] - b[j];
          cross3(ab, n, t);
        } else if (phase == 3) {
          idx[2] = gi; memcpy(c, vert + 3 * gi, sizeof(c));
          for (int j = 0; j < 3; j++) t[j] = a[j] - c[j];
          cross3(ac, n, t);
          for (int j = 0; j < 3; j++) t[j] = b[j] - c[j];
          cross3(bc, n, t);
        } else idx[3] = gi;
      }
      goto plane_mesh_collect;
    }
    for (int i = 0; i < nvert; i++) { real sp = SUPPORT(i); if (sp > max_support) { max_support = sp; idx[0] = i; memcpy(a, vert + 3 * i, sizeof(a)); } }
    if (max_support < 0) return 4;
    real threshold = max_support - 1e-3, bd = -HUGE_V, cd = -HUGE_V, dd = -HUGE_V, ab[3], ac[3], bc[3], t[3];
    for (int i = 0; i < nvert; i++) {
      real mask = SUPPORT(i) > threshold ? 0.0 : -HUGE_V, d;
      for (int j = 0; j < 3; j++) t[j] = a[j] - vert[3 * i + j];
      d = dot3(t, t) + mask;
      if (d > bd) { idx[1] = i; bd = d; memcpy(b, vert + 3 * i, sizeof(b)); }
    }
    for (int j = 0; j < 3; j++) t[j] = a[j] - b[j];
    cross3(ab, n, t);
    for (int i = 0; i < nvert; i++) {
      real mask = SUPPORT(i) > threshold ? 0.0 : -HUGE_V, d;
      for (int j = 0; j < 3; j++) t[j] = a[j] - vert[3 * i + j];
      d = fabs(dot3(t, ab)) + mask;
      if (d > cd) { idx[2] = i; cd = d; memcpy(c, vert + 3 * i, sizeof(c)); }
    }
    for (int j = 0; j < 3; j++) t[j] = a[j] - c[j];
    cross3(ac, n, t);
    for (int j = 0; j < 3; j++) t[j] = b[j] - c[j];
    cross3(bc, n, t);
    for (int i = 0; i < nvert; i++) {
      real mask = SUPPORT(i) > threshold ? 0.0 : -HUGE_V, ap[3], bp[3], d;
      for (int j = 0; j < 3; j++) { ap[j] = a[j] - vert[3 * i + j]; bp[j] = b[j] - vert[3 * i + j]; }
      d = fabs(dot3(ap, ac)) + mask + fabs(dot3(bp, bc)) + mask;
      if (d > dd) { idx[3] = i; dd = d; }
    }
  plane_mesh_collect:;
    int count = 0;
    for (int i = 3; i >= 0; i--) { /* :258-275 contacts from the indices that appear once among idx[0..i] */
      int rep = 0;
      for (int j = 0; j <= i; j++) rep += idx[j] == idx[i];
      if (rep != 1) continue;
      real w[3], dist = -SUPPORT(idx[i]);
      mat_vec(w, r2, vert + 3 * idx[i]);
      out[count].dist = dist;
      for (int j = 0; j < 3; j++) out[count].pos[j] = p2[j] + w[j] - 0.5 * dist * n1[j];
      count++;
    }
#undef SUPPORT
    return 4;
  }
  if (t1 == GEOM_PLANE && t2 == GEOM_SPHERE) { /* collision_primitive.py:280 */
    c_plane_sphere(n1, p1, p2, s2[0], &out[0].dist, out[0].pos);
    make_frame(out[0].frame, n1);
    out[0].id = 0;
    return 1;
  }
  if (t1 == GEOM_PLANE && t2 == GEOM_CAPSULE) { /* collision_primitive_core.py:312-361 */
    real b[3], c[3], nb, na = dot3(n1, n2);
    for (int i = 0; i < 3; i++) b[i] = n2[i] - n1[i] * na;
    nb = norm3(b);
    if (nb != 0.0) for (int i = 0; i < 3; i++) b[i] /= nb;
    if (nb < 0.5) {
      if (-0.5 < n1[1] && n1[1] < 0.5) { b[0] = 0; b[1] = 1; b[2] = 0; } else { b[0] = 0; b[1] = 0; b[2] = 1; }
    }
    cross3(c, n1, b);
    for (int k = 0; k < 2; k++) {
      real e[3], sg = k == 0 ? 1.0 : -1.0;
      for (int i = 0; i < 3; i++) e[i] = p2[i] + sg * n2[i] * s2[1];
      c_plane_sphere(n1, p1, e, s2[0], &out[k].dist, out[k].pos);
      for (int i = 0; i < 3; i++) { out[k].frame[i] = n1[i]; out[k].frame[3 + i] = b[i]; out[k].frame[6 + i] = c[i]; }
      out[k].id = k;
    }
    return 2;
  }
  if (t1 == GEOM_PLANE && t2 == GEOM_BOX) { /* collision_primitive_core.py:396-443 */
    real cd, dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
    cd = dot3(dif, n1);
    for (int k = 0; k < 8; k++) {
      real corner[3] = {(k & 1) ? s2[0] : -s2[0], (k & 2) ? s2[1] : -s2[1], (k & 4) ? s2[2] : -s2[2]}, cw[3];
      mat_vec(cw, r2, corner);
      real cdist = cd + dot3(n1, cw);
      out[k].dist = cdist;
      for (int i = 0; i < 3; i++) out[k].pos[i] = cw[i] + p2[i] - 0.5 * n1[i] * cdist;
      make_frame(out[k].frame, n1);
      out[k].id = k;
    }
    return 8;
  }
  if (t1 == GEOM_PLANE && t2 == GEOM_ELLIPSOID) { /* collision_primitive_core.py:365-392 */
    real ld[3], sup[3], loc[3], w[3], nn;
    for (int i = 0; i < 3; i++) ld[i] = (r2[i] * n1[0] + r2[3 + i] * n1[1] + r2[6 + i] * n1[2]) * s2[i];
    nn = norm3(ld);
    for (int i = 0; i < 3; i++) { sup[i] = nn > 0.0 ? -ld[i] / nn : 0.0; loc[i] = sup[i] * s2[i]; }
    mat_vec(w, r2, loc);
    real pe[3] = {p2[0] + w[0], p2[1] + w[1], p2[2] + w[2]}, dif[3] = {pe[0] - p1[0], pe[1] - p1[1], pe[2] - p1[2]};
    out[0].dist = dot3(n1, dif);
    for (int i = 0; i < 3; i++) out[0].pos[i] = pe[i] - n1[i] * out[0].dist * 0.5;
    make_frame(out[0].frame, n1);
    out[0].id = 0;
    return 1;
  }
  if (t1 == GEOM_PLANE && t2 == GEOM_CYLINDER) { /* collision_primitive_core.py:519-612 */
    real axis[3] = {n2[0], n2[1], n2[2]}, prjaxis = dot3(n1, n2), dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]}, vec[3], vec1[3];
    if (prjaxis > 0) { for (int i = 0; i < 3; i++) axis[i] = -axis[i]; prjaxis = -prjaxis; }
    real dist0 = dot3(dif, n1), len_sqr, prjvec, radius = s2[0], hh = s2[1];
    for (int i = 0; i < 3; i++) vec[i] = axis[i] * prjaxis - n1[i];
    len_sqr = dot3(vec, vec);
    if (len_sqr >= 1e-12) { real sc = safe_div(radius, sqrt(len_sqr)); for (int i = 0; i < 3; i++) vec[i] *= sc; }
    else { vec[0] = radius; vec[1] = 0; vec[2] = 0; }
    prjvec = dot3(vec, n1);
    for (int i = 0; i < 3; i++) axis[i] *= hh;
    prjaxis *= hh;
    real d1 = dist0 + prjaxis + prjvec, d2 = dist0 - prjaxis + prjvec, d3 = dist0 + prjaxis - prjvec * 0.5, nv1;
    cross3(vec1, vec, axis);
    nv1 = norm3(vec1);
    for (int i = 0; i < 3; i++) vec1[i] = (nv1 > 0.0 ? vec1[i] / nv1 : 0.0) * (radius * sqrt(3.0) * 0.5);
    for (int i = 0; i < 3; i++) {
      out[0].pos[i] = p2[i] + vec[i] + axis[i] - n1[i] * (d1 * 0.5);
      out[1].pos[i] = p2[i] + vec[i] - axis[i] - n1[i] * (d2 * 0.5);
      out[2].pos[i] = p2[i] + vec1[i] + axis[i] - vec[i] * 0.5 - n1[i] * (d3 * 0.5);
      out[3].pos[i] = p2[i] - vec1[i] + axis[i] - vec[i] * 0.5 - n1[i] * (d3 * 0.5);
    }
    out[0].dist = d1; out[1].dist = d2; out[2].dist = d3; out[3].dist = d3;
    for (int k = 0; k < 4; k++) { make_frame(out[k].frame, n1); out[k].id = k; }
    return 4;
  }
  if (t1 == GEOM_SPHERE && t2 == GEOM_CYLINDER) { /* collision_primitive_core.py:447-515 */
    real vec[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]}, x = dot3(vec, n2), aproj[3], pproj[3], n[3], psq;
    for (int i = 0; i < 3; i++) { aproj[i] = n2[i] * x; pproj[i] = vec[i] - aproj[i]; }
    psq = dot3(pproj, pproj);
    int side = fabs(x) < s2[1], cap = psq < s2[0] * s2[0];
    if (side && cap) { if (s2[1] - fabs(x) < s2[0] - sqrt(psq)) side = 0; else cap = 0; }
    if (side) {
      real tgt[3] = {p2[0] + aproj[0], p2[1] + aproj[1], p2[2] + aproj[2]};
      c_sphere_sphere(p1, s1[0], tgt, s2[0], &out[0].dist, out[0].pos, n);
    } else if (cap) {
      real sg = x > 0.0 ? 1.0 : -1.0, pc[3], pn[3];
      for (int i = 0; i < 3; i++) { pc[i] = p2[i] + sg * n2[i] * s2[1]; pn[i] = sg * n2[i]; }
      c_plane_sphere(pn, pc, p1, s1[0], &out[0].dist, out[0].pos);
      for (int i = 0; i < 3; i++) n[i] = -pn[i];
    } else {
      real inv = safe_div(1.0, sqrt(psq)), sg = x < 0.0 ? -1.0 : 1.0, corner[3];
      for (int i = 0; i < 3; i++) corner[i] = p2[i] + n2[i] * (sg * s2[1]) + pproj[i] * (s2[0] * inv);
      c_sphere_sphere(p1, s1[0], corner, 0.0, &out[0].dist, out[0].pos, n);
    }
    make_frame(out[0].frame, n);
    out[0].id = 0;
    return 1;
  }
  if (t1 == GEOM_SPHERE && t2 == GEOM_SPHERE) { /* collision_primitive.py:352 */
    real n[3];
    c_sphere_sphere(p1, s1[0], p2, s2[0], &out[0].dist, out[0].pos, n);
    make_frame(out[0].frame, n);
    out[0].id = 0;
    return 1;
  }
  if (t1 == GEOM_SPHERE && t2 == GEOM_CAPSULE) { /* collision_primitive_core.py:147-178 */
    real a[3], b[3], pt[3], n[3];
    for (int i = 0; i < 3; i++) { a[i] = p2[i] - n2[i] * s2[1]; b[i] = p2[i] + n2[i] * s2[1]; }
    closest_segment_point(pt, a, b, p1);
    c_sphere_sphere(p1, s1[0], pt, s2[0], &out[0].dist, out[0].pos, n);
    make_frame(out[0].frame, n);
    out[0].id = 0;
    return 1;
  }
  if (t1 == GEOM_SPHERE && t2 == GEOM_BOX) { /* collision_primitive.py:1047 */
    real n[3];
    c_sphere_box(p1, s1[0], p2, r2, s2, &out[0].dist, out[0].pos, n);
    make_frame(out[0].frame, n);
    out[0].id = 0;
    return 1;
  }
  if (t1 == GEOM_BOX && t2 == GEOM_BOX) { /* collision_primitive.py:1202 */
    real dist[8], pos[24], n[3];
    int nc = c_box_box(p1, r1, s1, p2, r2, s2, margin, dist, pos, n);
    for (int k = 0; k < 8; k++) {
      out[k].id = k;
      if (k < nc) { out[k].dist = dist[k]; memcpy(out[k].pos, pos + 3 * k, 3 * sizeof(real)); make_frame(out[k].frame, n); }
      else { out[k].dist = MJ_MAXVAL; memset(out[k].pos, 0, 3 * sizeof(real)); memset(out[k].frame, 0, 9 * sizeof(real)); }
    }
    return 8;
  }
  if (t1 == GEOM_CAPSULE && t2 == GEOM_BOX) { /* collision_primitive.py:1117 */
    real dist[2], pos[6], n[6];
    c_capsule_box(p1, n1, s1[0], s1[1], p2, r2, s2, dist, pos, n);
    for (int k = 0; k < 2; k++) {
      out[k].dist = dist[k]; out[k].id = k;
      memcpy(out[k].pos, pos + 3 * k, 3 * sizeof(real));
      if (dist[k] < MJ_MAXVAL) make_frame(out[k].frame, n + 3 * k); else memset(out[k].frame, 0, 9 * sizeof(real));
    }
    return 2;
  }
  if (t1 == GEOM_CAPSULE && t2 == GEOM_CAPSULE) { /* collision_primitive_core.py:182-308 */
    real ax1[3], ax2[3], dif[3], n[3], v1[3], v2[3], dist, pos[3];
    for (int i = 0; i < 3; i++) { ax1[i] = n1[i] * s1[1]; ax2[i] = n2[i] * s2[1]; dif[i] = p1[i] - p2[i]; }
    real ma = dot3(ax1, ax1), mb = -dot3(ax1, ax2), mc = dot3(ax2, ax2), u = -dot3(ax1, dif), v = dot3(ax2, dif);
    real det = ma * mc - mb * mb;
    out[0].dist = out[1].dist = INFINITY;
    out[0].id = 0; out[1].id = 1;
    memset(out[0].pos, 0, sizeof(real) * 12); memset(out[1].pos, 0, sizeof(real) * 12);
    if (fabs(det) >= MJ_MINVAL) {
      real x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
      if (x1 > 1.0) { x1 = 1.0; x2 = (v - mb) / mc; } else if (x1 < -1.0) { x1 = -1.0; x2 = (v + mb) / mc; }
      if (x2 > 1.0) { x2 = 1.0; x1 = clampr((u - mb) / ma, -1.0, 1.0); }
      else if (x2 < -1.0) { x2 = -1.0; x1 = clampr((u + mb) / ma, -1.0, 1.0); }
      for (int i = 0; i < 3; i++) { v1[i] = p1[i] + ax1[i] * x1; v2[i] = p2[i] + ax2[i] * x2; }
      c_sphere_sphere(v1, s1[0], v2, s2[0], &dist, pos, n);
      if (dist <= margin) { out[0].dist = dist; memcpy(out[0].pos, pos, sizeof(pos)); make_frame(out[0].frame, n); }
    } else {
      int cnt = 0;
      for (int k = 0; k < 4 && cnt < 2; k++) {
        real x1, x2;
        if (k == 0) { x1 = 1.0; x2 = clampr((v - mb) / mc, -1.0, 1.0); }
        else if (k == 1) { x1 = -1.0; x2 = clampr((v + mb) / mc, -1.0, 1.0); }
        else if (k == 2) { x2 = 1.0; x1 = clampr((u - mb) / ma, -1.0, 1.0); }
        else { x2 = -1.0; x1 = clampr((u + mb) / ma, -1.0, 1.0); }
        for (int i = 0; i < 3; i++) { v1[i] = p1[i] + ax1[i] * x1; v2[i] = p2[i] + ax2[i] * x2; }
        c_sphere_sphere(v1, s1[0], v2, s2[0], &dist, pos, n);
        if (dist <= margin) { out[cnt].dist = dist; memcpy(out[cnt].pos, pos, sizeof(pos)); make_frame(out[cnt].frame, n); cnt++; }
      }
    }
    return 2;
  }
  return -1; /* pair type outside the oracle's coverage */
}

/* collision_core.py:236-341 */
static void contact_params(const OModel* m, int g1, int g2, int pairid, real* margin, real* gap, int* condim,
                           real* friction, real* solref, real* solreffriction, real* solimp) {
  if (pairid > -1) {
    *margin = m->pair_margin[pairid]; *gap = m->pair_gap[pairid]; *condim = m->pair_dim[pairid];
    memcpy(friction, m->pair_friction + 5 * pairid, 5 * sizeof(real));
    memcpy(solref, m->pair_solref + 2 * pairid, 2 * sizeof(real));
    memcpy(solreffriction, m->pair_solreffriction + 2 * pairid, 2 * sizeof(real));
    memcpy(solimp, m->pair_solimp + 5 * pairid, 5 * sizeof(real));
  } else {
    real sm1 = m->geom_solmix[g1], sm2 = m->geom_solmix[g2], mix, f[3];
    int p1 = m->geom_priority[g1], p2 = m->geom_priority[g2];
    const real *f1 = m->geom_friction + 3 * g1, *f2 = m->geom_friction + 3 * g2;
    if (p1 > p2) { mix = 1.0; *condim = m->geom_condim[g1]; memcpy(f, f1, sizeof(f)); }
    else if (p2 > p1) { mix = 0.0; *condim = m->geom_condim[g2]; memcpy(f, f2, sizeof(f)); }
    else {
      mix = safe_div(sm1, sm1 + sm2);
      if (sm1 < MJ_MINVAL && sm2 < MJ_MINVAL) mix = 0.5;
      if (sm1 < MJ_MINVAL && sm2 >= MJ_MINVAL) mix = 0.0;
      if (sm1 >= MJ_MINVAL && sm2 < MJ_MINVAL) mix = 1.0;
      *condim = m->geom_condim[g1] > m->geom_condim[g2] ? m->geom_condim[g1] : m->geom_condim[g2];
      for (int i = 0; i < 3; i++) f[i] = f1[i] > f2[i] ? f1[i] : f2[i];
    }
    friction[0] = f[0]; friction[1] = f[0]; friction[2] = f[1]; friction[3] = f[2]; friction[4] = f[2];
    const real *sr1 = m->geom_solref + 2 * g1, *sr2 = m->geom_solref + 2 * g2;
    if (sr1[0] > 0.0 && sr2[0] > 0.0) { for (int i = 0; i < 2; i++) solref[i] = mix * sr1[i] + (1.0 - mix) * sr2[i]; }
    else { for (int i = 0; i < 2; i++) solref[i] = sr1[i] < sr2[i] ? sr1[i] : sr2[i]; }
    solreffriction[0] = solreffriction[1] = 0.0;
    for (int i = 0; i < 5; i++) solimp[i] = mix * m->geom_solimp[5 * g1 + i] + (1.0 - mix) * m->geom_solimp[5 * g2 + i];
    *margin = m->geom_margin[g1] + m->geom_margin[g2];
    *gap = m->geom_gap[g1] + m->geom_gap[g2];
  }
  for (int i = 0; i < 5; i++) friction[i] = friction[i] > MJ_MINMU ? friction[i] : MJ_MINMU;
}

/* collision_driver.sap_broadphase (collision_driver.py:554-644): bounding spheres (radius rbound + margin, planes
 * unbounded) are projected on one fixed direction (:573-574), sorted by their lower end (:458-482), and every geom
 * is paired with the geoms that follow it in sorted order up to and INCLUDING the first one whose lower end lies
 * beyond its upper end (_sap_range :411-432: limit = min(ngeom - 1, first index with lower > upper)).  The pairs
 * then go through the same static table and geometric filter as the NXN path (:485-551).  cand[p] marks entries of
 * the filtered pair list that the sweep reaches. */
static void sap_candidates(const OModel* m, const OData* d, unsigned char* cand) {
  const int ng = m->ngeom;
  real dir[3] = {0.5935, 0.7790, 0.1235};
  real nrm = sqrt(dot3(dir, dir));
  for (int i = 0; i < 3; i++) dir[i] /= nrm;
  real* lower = (real*)malloc(sizeof(real) * 2 * ng);
  real* upper = lower + ng;
  int* order = (int*)malloc(sizeof(int) * ng);
  for (int g = 0; g < ng; g++) {
    real rb = m->geom_rbound[g];
    if (rb == 0.0) rb = MJ_MAXVAL;
    real radius = rb + m->geom_margin[g], centre = dot3(dir, d->geom_xpos + 3 * g);
    if (isnan(centre)) { lower[g] = upper[g] = MJ_MAXVAL; } else { lower[g] = centre - radius; upper[g] = centre + radius; }
    order[g] = g;
  }
  for (int i = 1; i < ng; i++) { /* stable insertion sort by lower end */
    int g = order[i], j = i - 1;
    while (j >= 0 && lower[order[j]] > lower[g]) { order[j + 1] = order[j]; j--; }
    order[j + 1] = g;
  }
  memset(cand, 0, m->nxn);
  for (int i = 0; i < ng; i++) {
    int limit = i + 1;
    while (limit < ng && !(lower[order[limit]] > upper[order[i]])) limit++;
    if (limit > ng - 1) limit = ng - 1;
    for (int j = i + 1; j <= limit; j++) {
      int a = order[i], b = order[j];
      if (b < a) { int t = a; a = b; b = t; }
      int p = m->nxn_lookup[(a * (2 * ng - a - 3)) / 2 + b - 1];
      if (p >= 0) cand[p] = 1;
    }
  }
  free(lower); free(order);
}

#define O_MAXWCON 512
typedef struct {
  OContactGeom g;
  real includemargin, friction[5], solref[2], solreffriction[2], solimp[5];
  int dim, geom[2], type;
} OContactFull;

/* collision_driver.collision (collision_driver.py:755): NXN broadphase (:647) and primitive
 * narrowphase (collision_primitive.py:1300).  Returns <0 if an unsupported pair type is met. */
static int o_collision(const OModel* m, const OOpt* o, OData* d, const OData* glob, int w) {
  int unsupported = 0, ncol = 0, ncon = 0;
  OContactFull* buf = (OContactFull*)malloc(sizeof(OContactFull) * O_MAXWCON);
  d->wcon_n[0] = 0;
  if (glob->naconmax == 0 || (o->disableflags & (DSBL_CONSTRAINT | DSBL_CONTACT))) { free(buf); return 0; }
  unsigned char* cand = NULL;
  if (o->broadphase != 0 && m->nxn > 0) { /* BroadphaseType.SAP_TILE / SAP_SEGMENTED (types.py) differ only in the sort */
    cand = (unsigned char*)malloc(m->nxn);
    sap_candidates(m, d, cand);
  }
  for (int p = 0; p < m->nxn; p++) {
    int ga = m->nxn_pair[2 * p], gb = m->nxn_pair[2 * p + 1];
    int pid0 = m->nxn_pairid[2 * p], pid1 = m->nxn_pairid[2 * p + 1];
    if (cand && !cand[p]) continue;
    if (!(broadphase_filter(m, d, o->broadphase_filter, ga, gb) || pid1 >= 0)) continue;
    ncol++;
    int g1 = ga, g2 = gb;
    if (m->geom_type[ga] > m->geom_type[gb]) { g1 = gb; g2 = ga; } /* collision_driver.py:351 */
    real margin, gap, fr[5], sr[2], srf[2], si[5];
    int condim;
    contact_params(m, g1, g2, pid0, &margin, &gap, &condim, fr, sr, srf, si);
    OContactGeom cg[8];
    int n = narrowphase_pair(m, d, g1, g2, margin, o->ccd_tolerance, o->ccd_iterations, !(o->disableflags & DSBL_NATIVECCD), cg);
    if (n < 0) { unsupported = 1; continue; }
    for (int k = 0; k < n; k++) { /* write_contact, collision_core.py:160-232 */
      int active = cg[k].dist < margin;
      if ((pid0 == -2 || !active) && pid1 == -1) continue;
      int type = 0;
      if (pid0 >= -1 && active) type |= 1;
      if (pid1 >= 0) type |= 2;
      if (ncon < O_MAXWCON) {
        OContactFull* c = buf + ncon;
        c->g = cg[k]; c->includemargin = margin - gap; c->dim = condim; c->geom[0] = g1; c->geom[1] = g2; c->type = type;
        memcpy(c->friction, fr, sizeof(fr)); memcpy(c->solref, sr, sizeof(sr));
        memcpy(c->solreffriction, srf, sizeof(srf)); memcpy(c->solimp, si, sizeof(si));
      }
      ncon++;
    }
  }
  free(cand);
  if (ncon > O_MAXWCON) ncon = O_MAXWCON;
  int base, cbase;
#pragma omp atomic capture
  { base = glob->nacon[0]; glob->nacon[0] += ncon; }
#pragma omp atomic capture
  { cbase = glob->ncollision[0]; glob->ncollision[0] += ncol; }
  (void)cbase;
  int nw = 0;
  for (int k = 0; k < ncon; k++) {
    int cid = base + k;
    if (cid >= glob->naconmax) { d->status[0] |= MJB_STATUS_NCON_OVERFLOW; break; }
    const OContactFull* c = buf + k;
    glob->contact_dist[cid] = c->g.dist;
    memcpy(glob->contact_pos + 3 * cid, c->g.pos, 3 * sizeof(real));
    memcpy(glob->contact_frame + 9 * cid, c->g.frame, 9 * sizeof(real));
    glob->contact_includemargin[cid] = c->includemargin;
    memcpy(glob->contact_friction + 5 * cid, c->friction, 5 * sizeof(real));
    memcpy(glob->contact_solref + 2 * cid, c->solref, 2 * sizeof(real));
    memcpy(glob->contact_solreffriction + 2 * cid, c->solreffriction, 2 * sizeof(real));
    memcpy(glob->contact_solimp + 5 * cid, c->solimp, 5 * sizeof(real));
    glob->contact_dim[cid] = c->dim;
    glob->contact_geom[2 * cid] = c->geom[0]; glob->contact_geom[2 * cid + 1] = c->geom[1];
    glob->contact_worldid[cid] = w;
    glob->contact_type[cid] = c->type;
    glob->contact_geomcollisionid[cid] = c->g.id;
    for (int i = 0; i < m->nmaxpyramid; i++) glob->contact_efc_address[cid * m->nmaxpyramid + i] = -1;
    if (nw < glob->nwconmax) d->wcon[nw++] = cid; else d->status[0] |= MJB_STATUS_WCON_OVERFLOW;
  }
  d->wcon_n[0] = nw;
  free(buf);
  return unsupported ? -1 : 0;
}

/* ------------------------------------------------------------------------------------------ */
/* constraint.make_constraint (constraint.py:2210)                                             */
/* ------------------------------------------------------------------------------------------ */
/* constraint.py:53-121 */
static void efc_row(const OOpt* o, OData* d, int r, real pos_aref, real pos_imp, real invweight, const real* solref,
                    const real* solimp, real margin, real vel, real frictionloss, int type, int id) {
  real timeconst = solref[0], dampratio = solref[1];
  real dmin = solimp[0], dmax = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  if (!(o->disableflags & DSBL_REFSAFE)) timeconst = fmax(timeconst, 2.0 * o->timestep);
  dmin = clampr(dmin, MJ_MINIMP, MJ_MAXIMP);
  dmax = clampr(dmax, MJ_MINIMP, MJ_MAXIMP);
  width = fmax(MJ_MINVAL, width);
  mid = clampr(mid, MJ_MINIMP, MJ_MAXIMP);
  power = fmax(1.0, power);
  real dmax_sq = dmax * dmax;
  real k = 1.0 / (dmax_sq * timeconst * timeconst * dampratio * dampratio);
  real b = 2.0 / (dmax * timeconst);
  if (solref[0] <= 0) k = -solref[0] / dmax_sq;
  if (solref[1] <= 0) b = -solref[1] / dmax;
  real imp_x = fabs(pos_imp) / width;
  real imp_a = (1.0 / pow(mid, power - 1.0)) * pow(imp_x, power);
  real imp_b = 1.0 - (1.0 / pow(1.0 - mid, power - 1.0)) * pow(1.0 - imp_x, power);
  real imp_y = imp_x < mid ? imp_a : imp_b;
  real imp = clampr(dmin + imp_y * (dmax - dmin), dmin, dmax);
  if (imp_x > 1.0) imp = dmax;
  d->efc_D[r] = 1.0 / fmax(invweight * (1.0 - imp) / imp, MJ_MINVAL);
  d->efc_vel[r] = vel;
  d->efc_aref[r] = -k * imp * pos_aref - b * vel;
  d->efc_pos[r] = pos_aref + margin;
  d->efc_margin[r] = margin;
  d->efc_frictionloss[r] = frictionloss;
  d->efc_type[r] = type;
  d->efc_id[r] = id;
}

/* support.jac_dof (support.py:397-432) */
static void jac_dof(const OModel* m, const OData* d, const real* point, int bodyid, int dofid, real* jacp, real* jacr) {
  int db = m->dof_bodyid[dofid], in_tree = (db == 0), p = bodyid;
  while (p != 0) {
    if (p == db) { in_tree = 1; break; }
    p = m->body_parentid[p];
  }
  if (!in_tree) { for (int i = 0; i < 3; i++) jacp[i] = jacr[i] = 0; return; }
  real off[3];
  for (int i = 0; i < 3; i++) off[i] = point[i] - d->subtree_com[3 * m->body_rootid[bodyid] + i];
  const real* cd = d->cdof + 6 * dofid;
  real c[3];
  cross3(c, cd, off);
  for (int i = 0; i < 3; i++) { jacp[i] = cd[3 + i] + c[i]; jacr[i] = cd[i]; }
}

/* sensor.py:2854 energy_pos: potential energy = -sum m g.xipos + joint springs (:2704-2812) */
static void o_energy_pos(const OModel* m, const OOpt* o, OData* d) {
  real e = 0;
  if (!(o->disableflags & DSBL_GRAVITY)) {
    real g[3] = {o->gravity_x, o->gravity_y, o->gravity_z};
    for (int b = 1; b < m->nbody; b++) e -= m->body_mass[b] * dot3(g, d->xipos + 3 * b);
  }
  if (!(o->disableflags & DSBL_SPRING)) {
    for (int j = 0; j < m->njnt; j++) {
      real k = m->jnt_stiffness[j];
      if (k == 0.0) continue;
      int pa = m->jnt_qposadr[j], t = m->jnt_type[j];
      const real *q = d->qpos + pa, *qs = m->qpos_spring + pa;
      if (t == JNT_FREE || t == JNT_BALL) {
        real s = 0, quat[4], dif[3];
        int qo = 0;
        if (t == JNT_FREE) { for (int i = 0; i < 3; i++) s += (q[i] - qs[i]) * (q[i] - qs[i]); qo = 3; }
        memcpy(quat, q + qo, sizeof(quat));
        normalize4(quat);
        quat_sub(dif, quat, qs + qo);
        e += 0.5 * k * (s + dot3(dif, dif));
      } else {
        e += 0.5 * k * (q[0] - qs[0]) * (q[0] - qs[0]);
      }
    }
  }
  d->energy[0] = e;
  d->energy[1] = 0;
}
/* sensor.py:2922 energy_vel: kinetic energy = 0.5 qvel^T M qvel */
static void o_energy_vel(const OModel* m, OData* d) {
  real e = 0;
  for (int i = 0; i < m->nv; i++) {
    real s = 0;
    for (int j = 0; j < m->nv; j++) s += d->qM[i * m->nv_pad + j] * d->qvel[j];
    e += d->qvel[i] * s;
  }
  d->energy[1] = 0.5 * e;
}

static void o_make_constraint(const OModel* m, const OOpt* o, OData* d, const OData* glob) {
  int nv = m->nv, np = m->nv_pad, njmax = glob->njmax, nefc = 0;
  d->ne[0] = d->nf[0] = d->nl[0] = d->nefc[0] = 0;
  if (o->disableflags & DSBL_CONSTRAINT) return;
  real ts = o->timestep;
  (void)ts;
  if (!(o->disableflags & DSBL_EQUALITY)) {
    for (int k = 0; k < m->neqconnect; k++) { /* constraint.py:125-364 connect (bodies) */
      int eq = m->eq_connect_adr[k];
      if (!d->eq_active[eq]) continue;
      d->ne[0] += 3;
      int r = nefc;
      nefc += 3;
      if (r >= njmax - 3) continue; /* :186 (sic) */
      const real* data = m->eq_data + 11 * eq;
      int b1 = m->eq_obj1id[eq], b2 = m->eq_obj2id[eq];
      real pos1[3], pos2[3], pos[3], Jqvel[3] = {0, 0, 0}, t[3];
      if (m->nsite > 0 && m->eq_objtype[eq] == OBJ_SITE) { /* :203-208 */
        memcpy(pos1, d->site_xpos + 3 * b1, sizeof(pos1));
        memcpy(pos2, d->site_xpos + 3 * b2, sizeof(pos2));
        b1 = m->site_bodyid[b1];
        b2 = m->site_bodyid[b2];
      } else {
        mat_vec(t, d->xmat + 9 * b1, data);
        for (int i = 0; i < 3; i++) pos1[i] = d->xpos[3 * b1 + i] + t[i];
        mat_vec(t, d->xmat + 9 * b2, data + 3);
        for (int i = 0; i < 3; i++) pos2[i] = d->xpos[3 * b2 + i] + t[i];
      }
      for (int i = 0; i < 3; i++) pos[i] = pos1[i] - pos2[i];
      for (int dof = 0; dof < nv; dof++) {
        real jp1[3], jp2[3], jr[3];
        jac_dof(m, d, pos1, b1, dof, jp1, jr);
        jac_dof(m, d, pos2, b2, dof, jp2, jr);
        for (int i = 0; i < 3; i++) {
          d->efc_J[(size_t)(r + i) * np + dof] = jp1[i] - jp2[i];
          Jqvel[i] += (jp1[i] - jp2[i]) * d->qvel[dof];
        }
      }
      real invweight = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2], pos_imp = norm3(pos);
      for (int i = 0; i < 3; i++)
        efc_row(o, d, r + i, pos[i], pos_imp, invweight, m->eq_solref + 2 * eq, m->eq_solimp + 5 * eq, 0.0, Jqvel[i], 0.0, CT_EQUALITY, eq);
    }
    for (int k = 0; k < m->neqweld; k++) { /* constraint.py:793-1108 weld (bodies) */
      int eq = m->eq_wld_adr[k];
      if (!d->eq_active[eq]) continue;
      d->ne[0] += 6;
      int r = nefc;
      nefc += 6;
      if (r >= njmax - 6) continue; /* :855 (sic) */
      const real* data = m->eq_data + 11 * eq;
      int b1 = m->eq_obj1id[eq], b2 = m->eq_obj2id[eq];
      real torquescale = data[10], pos1[3], pos2[3], cpos[3], t[3], quat[4], quat1[4], Jqp[3] = {0, 0, 0}, Jqr[3] = {0, 0, 0};
      if (m->nsite > 0 && m->eq_objtype[eq] == OBJ_SITE) { /* :868-877: the site frames are welded */
        int s1 = b1, s2 = b2;
        real q2[4];
        memcpy(pos1, d->site_xpos + 3 * s1, sizeof(pos1));
        memcpy(pos2, d->site_xpos + 3 * s2, sizeof(pos2));
        b1 = m->site_bodyid[s1];
        b2 = m->site_bodyid[s2];
        mul_quat(quat, d->xquat + 4 * b1, m->site_quat + 4 * s1);
        mul_quat(q2, d->xquat + 4 * b2, m->site_quat + 4 * s2);
        quat1[0] = q2[0];
        for (int i = 1; i < 4; i++) quat1[i] = -q2[i];
      } else {
        mat_vec(t, d->xmat + 9 * b1, data + 3); /* anchor2 = data[3:6] lives in body1 */
        for (int i = 0; i < 3; i++) pos1[i] = d->xpos[3 * b1 + i] + t[i];
        mat_vec(t, d->xmat + 9 * b2, data);
        for (int i = 0; i < 3; i++) pos2[i] = d->xpos[3 * b2 + i] + t[i];
        mul_quat(quat, d->xquat + 4 * b1, data + 6);
        quat1[0] = d->xquat[4 * b2];
        for (int i = 1; i < 4; i++) quat1[i] = -d->xquat[4 * b2 + i];
      }
      for (int i = 0; i < 3; i++) cpos[i] = pos1[i] - pos2[i];
      for (int dof = 0; dof < nv; dof++) {
        real jp1[3], jp2[3], jr1[3], jr2[3], dr[3], qa[4], qb[4];
        jac_dof(m, d, pos1, b1, dof, jp1, jr1);
        jac_dof(m, d, pos2, b2, dof, jp2, jr2);
        for (int i = 0; i < 3; i++) dr[i] = (jr1[i] - jr2[i]) * torquescale;
        qa[0] = -quat1[1] * dr[0] - quat1[2] * dr[1] - quat1[3] * dr[2]; /* math.py:34 quat_mul_axis */
        qa[1] = quat1[0] * dr[0] + quat1[2] * dr[2] - quat1[3] * dr[1];
        qa[2] = quat1[0] * dr[1] + quat1[3] * dr[0] - quat1[1] * dr[2];
        qa[3] = quat1[0] * dr[2] + quat1[1] * dr[1] - quat1[2] * dr[0];
        mul_quat(qb, qa, quat);
        for (int i = 0; i < 3; i++) {
          real jp = jp1[i] - jp2[i], jr = 0.5 * qb[1 + i];
          d->efc_J[(size_t)(r + i) * np + dof] = jp;
          d->efc_J[(size_t)(r + 3 + i) * np + dof] = jr;
          Jqp[i] += jp * d->qvel[dof];
          Jqr[i] += jr * d->qvel[dof];
        }
      }
      real crotq[4], crot[3];
      mul_quat(crotq, quat1, quat);
      for (int i = 0; i < 3; i++) crot[i] = crotq[1 + i] * torquescale;
      real pos_imp = sqrt(dot3(cpos, cpos) + dot3(crot, crot));
      real iw_t = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2], iw_r = m->body_invweight0[2 * b1 + 1] + m->body_invweight0[2 * b2 + 1];
      for (int i = 0; i < 3; i++) {
        efc_row(o, d, r + i, cpos[i], pos_imp, iw_t, m->eq_solref + 2 * eq, m->eq_solimp + 5 * eq, 0.0, Jqp[i], 0.0, CT_EQUALITY, eq);
        efc_row(o, d, r + 3 + i, crot[i], pos_imp, iw_r, m->eq_solref + 2 * eq, m->eq_solimp + 5 * eq, 0.0, Jqr[i], 0.0, CT_EQUALITY, eq);
      }
    }
    /* constraint.py:368-497 joint equality */
    for (int k = 0; k < m->neqjnt; k++) {
      int eq = m->eq_jnt_adr[k];
      if (!d->eq_active[eq]) continue;
      d->ne[0]++;
      int r = nefc++;
      if (r >= njmax) continue;
      int j1 = m->eq_obj1id[eq], j2 = m->eq_obj2id[eq];
      const real* data = m->eq_data + 11 * eq;
      int da1 = m->jnt_dofadr[j1], qa1 = m->jnt_qposadr[j1];
      real* J = d->efc_J + (size_t)r * np;
      for (int i = 0; i < nv; i++) J[i] = 0;
      J[da1] = 1.0;
      real pos, Jqvel, invweight;
      if (j2 > -1) {
        int qa2 = m->jnt_qposadr[j2], da2 = m->jnt_dofadr[j2];
        real dif = d->qpos[qa2] - m->qpos0[qa2];
        real rhs = data[0] + dif * (data[1] + dif * (data[2] + dif * (data[3] + dif * data[4])));
        real deriv2 = data[1] + dif * (2.0 * data[2] + dif * (3.0 * data[3] + dif * 4.0 * data[4]));
        pos = d->qpos[qa1] - m->qpos0[qa1] - rhs;
        Jqvel = d->qvel[da1] - d->qvel[da2] * deriv2;
        invweight = m->dof_invweight0[da1] + m->dof_invweight0[da2];
        J[da2] = -deriv2;
      } else {
        pos = d->qpos[qa1] - m->qpos0[qa1] - data[0];
        Jqvel = d->qvel[da1];
        invweight = m->dof_invweight0[da1];
      }
      efc_row(o, d, r, pos, pos, invweight, m->eq_solref + 2 * eq, m->eq_solimp + 5 * eq, 0.0, Jqvel, 0.0, CT_EQUALITY, eq);
    }
  }
  if (!(o->disableflags & DSBL_FRICTIONLOSS)) { /* constraint.py:1114 friction dof */
    for (int k = 0; k < m->nfrictiondof; k++) {
      int dof = m->dof_friction_adr[k];
      d->nf[0]++;
      int r = nefc++;
      if (r >= njmax) continue;
      real* J = d->efc_J + (size_t)r * np;
      for (int i = 0; i < nv; i++) J[i] = 0;
      J[dof] = 1.0;
      efc_row(o, d, r, 0.0, 0.0, m->dof_invweight0[dof], m->dof_solref + 2 * dof, m->dof_solimp + 5 * dof, 0.0, d->qvel[dof],
              m->dof_frictionloss[dof], CT_FRICTION_DOF, dof);
    }
  }
  if (!(o->disableflags & DSBL_LIMIT)) { /* constraint.py:1317-1418 */
    for (int k = 0; k < m->nlimit; k++) {
      int j = m->jnt_limited_adr[k];
      real q = d->qpos[m->jnt_qposadr[j]], lo = m->jnt_range[2 * j], hi = m->jnt_range[2 * j + 1], mg = m->jnt_margin[j];
      real dmin = q - lo, dmax = hi - q, pos = fmin(dmin, dmax) - mg;
      if (!(pos < 0)) continue;
      d->nl[0]++;
      int r = nefc++;
      if (r >= njmax) continue;
      int da = m->jnt_dofadr[j];
      real Jv = (dmin < dmax) ? 1.0 : -1.0;
      real* J = d->efc_J + (size_t)r * np;
      for (int i = 0; i < nv; i++) J[i] = 0;
      J[da] = Jv;
      efc_row(o, d, r, pos, pos, m->dof_invweight0[da], m->jnt_solref + 2 * j, m->jnt_solimp + 5 * j, mg, Jv * d->qvel[da], 0.0,
              CT_LIMIT_JOINT, j);
    }
    for (int k = 0; k < m->nlimitball; k++) { /* constraint.py:1422-1545 _limit_ball */
      int j = m->jnt_limited_ball_adr[k];
      real q[4], aa[3], axis[3] = {0, 0, 0};
      for (int i = 0; i < 4; i++) q[i] = d->qpos[m->jnt_qposadr[j] + i];
      normalize4(q);
      quat_to_vel(aa, q);
      real angle = sqrt(aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2]), mg = m->jnt_margin[j];
      for (int i = 0; i < 3; i++) axis[i] = angle == 0 ? aa[i] : aa[i] / angle; /* math.py:261 normalize_with_norm */
      real pos = fmax(m->jnt_range[2 * j], m->jnt_range[2 * j + 1]) - angle - mg;
      if (!(pos < 0)) continue;
      d->nl[0]++;
      int r = nefc++;
      if (r >= njmax) continue;
      int da = m->jnt_dofadr[j];
      real* J = d->efc_J + (size_t)r * np;
      for (int i = 0; i < nv; i++) J[i] = 0;
      real jqvel = 0;
      for (int i = 0; i < 3; i++) { J[da + i] = -axis[i]; jqvel -= axis[i] * d->qvel[da + i]; }
      efc_row(o, d, r, pos, pos, m->dof_invweight0[da], m->jnt_solref + 2 * j, m->jnt_solimp + 5 * j, mg, jqvel, 0.0, CT_LIMIT_JOINT, j);
    }
  }
  if (!(o->disableflags & DSBL_CONTACT)) { /* constraint.py:1669-1936 pyramidal */
    int nmp = m->nmaxpyramid;
    for (int k = 0; k < d->wcon_n[0]; k++) {
      int cid = d->wcon[k];
      if (!(glob->contact_type[cid] & 1)) continue;
      int condim = glob->contact_dim[cid];
      real includemargin = glob->contact_includemargin[cid];
      real pos = glob->contact_dist[cid] - includemargin;
      if (!(pos < 0)) continue;
      int elliptic = o->cone == 1; /* ConeType.ELLIPTIC */
      int nrow = condim == 1 ? 1 : (elliptic ? condim : 2 * (condim - 1));
      int g1 = glob->contact_geom[2 * cid], g2 = glob->contact_geom[2 * cid + 1];
      int b1 = m->geom_bodyid[g1], b2 = m->geom_bodyid[g2];
      const real *cpos = glob->contact_pos + 3 * cid, *frame = glob->contact_frame + 9 * cid, *fri = glob->contact_friction + 5 * cid;
      real invw0 = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2];
      int wb1 = m->body_weldid[b1], wb2 = m->body_weldid[b2];
      if (elliptic) { /* constraint.py:1940-2206: rows = normal, then one per friction direction, in contact-frame axes */
        real invw = m->body_invweight0[2 * wb1] + m->body_invweight0[2 * wb2]; /* after the weld mapping (:2049-2050, :2146) */
        for (int dimid = 0; dimid < nrow; dimid++) {
          int r = nefc++;
          if (r >= njmax) { glob->contact_efc_address[cid * nmp + dimid] = -1; continue; }
          glob->contact_efc_address[cid * nmp + dimid] = r;
          real* J = d->efc_J + (size_t)r * np;
          real Jqvel = 0;
          for (int dof = 0; dof < nv; dof++) {
            real j1p[3], j1r[3], j2p[3], j2r[3], Jd = 0;
            jac_dof(m, d, cpos, wb1, dof, j1p, j1r);
            jac_dof(m, d, cpos, wb2, dof, j2p, j2r);
            for (int x = 0; x < 3; x++)
              Jd += dimid < 3 ? frame[3 * dimid + x] * (j2p[x] - j1p[x]) : frame[3 * (dimid - 3) + x] * (j2r[x] - j1r[x]);
            J[dof] = Jd;
            Jqvel += Jd * d->qvel[dof];
          }
          const real* ref = glob->contact_solref + 2 * cid;
          real invweight = invw, pos_aref = pos;
          if (dimid > 0) {
            const real* srf = glob->contact_solreffriction + 2 * cid;
            if (srf[0] != 0.0 || srf[1] != 0.0) ref = srf;
            invweight *= o->impratio_invsqrt * o->impratio_invsqrt;
            if (dimid > 1) invweight *= fri[0] * fri[0] / (fri[dimid - 1] * fri[dimid - 1]);
            pos_aref = 0.0;
          }
          efc_row(o, d, r, pos_aref, pos, invweight, ref, glob->contact_solimp + 5 * cid, includemargin, Jqvel, 0.0,
                  condim == 1 ? CT_CONTACT_FRICTIONLESS : CT_CONTACT_ELLIPTIC, cid);
        }
        continue;
      }
      for (int dimid = 0; dimid < nrow; dimid++) {
        int r = nefc++;
        if (r >= njmax) { glob->contact_efc_address[cid * nmp + dimid] = -1; continue; }
        glob->contact_efc_address[cid * nmp + dimid] = r;
        real invweight = invw0, frii = 0;
        int dimid2 = 0;
        if (condim > 1) {
          dimid2 = dimid / 2 + 1;
          real fri0 = fri[0];
          frii = fri[dimid2 - 1];
          invweight = invweight + fri0 * fri0 * invweight;
          invweight = invweight * 2.0 * fri0 * fri0 * o->impratio_invsqrt * o->impratio_invsqrt;
        }
        real* J = d->efc_J + (size_t)r * np;
        real Jqvel = 0;
        for (int dof = 0; dof < nv; dof++) {
          real j1p[3], j1r[3], j2p[3], j2r[3], Jn = 0, Ji = 0;
          jac_dof(m, d, cpos, wb1, dof, j1p, j1r);
          jac_dof(m, d, cpos, wb2, dof, j2p, j2r);
          for (int x = 0; x < 3; x++) {
            real dp = j2p[x] - j1p[x];
            Jn += frame[x] * dp;
            if (condim > 1) {
              if (dimid2 < 3) Ji += frame[3 * dimid2 + x] * dp;
              else Ji += frame[3 * (dimid2 - 3) + x] * (j2r[x] - j1r[x]);
            }
          }
          if (condim > 1) Jn += (dimid % 2 == 0) ? Ji * frii : -Ji * frii;
          J[dof] = Jn;
          Jqvel += Jn * d->qvel[dof];
        }
        efc_row(o, d, r, pos, pos, invweight, glob->contact_solref + 2 * cid, glob->contact_solimp + 5 * cid, includemargin, Jqvel,
                0.0, condim == 1 ? CT_CONTACT_FRICTIONLESS : CT_CONTACT_PYRAMIDAL, cid);
      }
    }
  }
  d->nefc[0] = nefc;
  if (nefc > njmax) d->status[0] |= MJB_STATUS_NEFC_OVERFLOW;
}

/* smooth.transmission (smooth.py:2042-2149), JOINT / JOINTINPARENT */
static void o_transmission(const OModel* m, OData* d) {
  int adr = 0;
  for (int a = 0; a < m->nu; a++) {
    int j = m->actuator_trnid[2 * a], jt = m->jnt_type[j], qa = m->jnt_qposadr[j], va = m->jnt_dofadr[j];
    const real* gear = m->actuator_gear + 6 * a;
    int trn = m->actuator_trntype[a];
    d->moment_rowadr[a] = adr;
    if (jt == JNT_FREE) {
      d->moment_rownnz[a] = 6;
      d->actuator_length[a] = 0;
      real g[6];
      memcpy(g, gear, sizeof(g));
      if (trn == 1) {
        real q[4] = {d->qpos[qa + 3], d->qpos[qa + 4], d->qpos[qa + 5], d->qpos[qa + 6]};
        normalize4(q);
        q[1] = -q[1]; q[2] = -q[2]; q[3] = -q[3];
        rot_vec_quat(g + 3, gear + 3, q);
      }
      for (int i = 0; i < 6; i++) { d->moment_colind[adr + i] = va + i; d->actuator_moment[adr + i] = g[i]; }
      adr += 6;
    } else if (jt == JNT_BALL) {
      real q[4] = {d->qpos[qa], d->qpos[qa + 1], d->qpos[qa + 2], d->qpos[qa + 3]}, aa[3], g[3] = {gear[0], gear[1], gear[2]};
      normalize4(q);
      quat_to_vel(aa, q);
      if (trn == 1) {
        real qn[4] = {q[0], -q[1], -q[2], -q[3]}, t[3];
        rot_vec_quat(t, g, qn);
        memcpy(g, t, sizeof(t));
      }
      d->actuator_length[a] = dot3(aa, g);
      d->moment_rownnz[a] = 3;
      for (int i = 0; i < 3; i++) { d->moment_colind[adr + i] = va + i; d->actuator_moment[adr + i] = g[i]; }
      adr += 3;
    } else {
      d->actuator_length[a] = d->qpos[qa] * gear[0];
      d->moment_rownnz[a] = 1;
      d->moment_colind[adr] = va;
      d->actuator_moment[adr] = gear[0];
      adr += 1;
    }
  }
}

/* forward.fwd_velocity (forward.py:593): actuator_velocity :541, com_vel smooth.py:1936-2038,
 * passive passive.py:71-180,536-565, rne smooth.py:1113-1274 */
static void o_fwd_velocity(const OModel* m, const OOpt* o, OData* d) {
  int nb = m->nbody, nv = m->nv;
  for (int a = 0; a < m->nu; a++) {
    real v = 0;
    for (int i = 0; i < d->moment_rownnz[a]; i++) {
      int s = d->moment_rowadr[a] + i;
      v += d->actuator_moment[s] * d->qvel[d->moment_colind[s]];
    }
    d->actuator_velocity[a] = v;
  }
  /* com_vel */
  for (int i = 0; i < 6; i++) d->cvel[i] = 0;
  for (int b = 1; b < nb; b++) {
    real cvel[6];
    memcpy(cvel, d->cvel + 6 * m->body_parentid[b], sizeof(cvel));
    int dof = m->body_dofadr[b];
    for (int j = m->body_jntadr[b]; j < m->body_jntadr[b] + m->body_jntnum[b]; j++) {
      int jt = m->jnt_type[j];
      if (jt == JNT_FREE) {
        for (int k = 0; k < 3; k++)
          for (int i = 0; i < 6; i++) cvel[i] += d->cdof[6 * (dof + k) + i] * d->qvel[dof + k];
        for (int k = 3; k < 6; k++) motion_cross(d->cdof_dot + 6 * (dof + k), cvel, d->cdof + 6 * (dof + k));
        for (int k = 0; k < 3; k++) memset(d->cdof_dot + 6 * (dof + k), 0, 6 * sizeof(real));
        for (int k = 3; k < 6; k++)
          for (int i = 0; i < 6; i++) cvel[i] += d->cdof[6 * (dof + k) + i] * d->qvel[dof + k];
        dof += 6;
      } else if (jt == JNT_BALL) {
        for (int k = 0; k < 3; k++) motion_cross(d->cdof_dot + 6 * (dof + k), cvel, d->cdof + 6 * (dof + k));
        for (int k = 0; k < 3; k++)
          for (int i = 0; i < 6; i++) cvel[i] += d->cdof[6 * (dof + k) + i] * d->qvel[dof + k];
        dof += 3;
      } else {
        motion_cross(d->cdof_dot + 6 * dof, cvel, d->cdof + 6 * dof);
        for (int i = 0; i < 6; i++) cvel[i] += d->cdof[6 * dof + i] * d->qvel[dof];
        dof += 1;
      }
    }
    memcpy(d->cvel + 6 * b, cvel, sizeof(cvel));
  }
  /* passive: spring / damper */
  int dsbl_spring = o->disableflags & DSBL_SPRING, dsbl_damper = o->disableflags & DSBL_DAMPER;
  for (int i = 0; i < nv; i++) d->qfrc_spring[i] = d->qfrc_damper[i] = d->qfrc_gravcomp[i] = d->qfrc_fluid[i] = 0;
  if (!(dsbl_spring && dsbl_damper)) {
    for (int j = 0; j < m->njnt; j++) {
      int da = m->jnt_dofadr[j], qa = m->jnt_qposadr[j], jt = m->jnt_type[j];
      real k = m->jnt_stiffness[j], c = m->dof_damping[da];
      int hs = k != 0.0 && !dsbl_spring, hd = c != 0.0 && !dsbl_damper;
      int n = jt == JNT_FREE ? 6 : (jt == JNT_BALL ? 3 : 1);
      if (hs) {
        if (jt == JNT_FREE) {
          for (int i = 0; i < 3; i++) d->qfrc_spring[da + i] = -k * (d->qpos[qa + i] - m->qpos_spring[qa + i]);
          real rot[4] = {d->qpos[qa + 3], d->qpos[qa + 4], d->qpos[qa + 5], d->qpos[qa + 6]}, dif[3];
          normalize4(rot);
          quat_sub(dif, rot, m->qpos_spring + qa + 3);
          for (int i = 0; i < 3; i++) d->qfrc_spring[da + 3 + i] = -k * dif[i];
        } else if (jt == JNT_BALL) {
          real rot[4] = {d->qpos[qa], d->qpos[qa + 1], d->qpos[qa + 2], d->qpos[qa + 3]}, dif[3];
          normalize4(rot);
          quat_sub(dif, rot, m->qpos_spring + qa);
          for (int i = 0; i < 3; i++) d->qfrc_spring[da + i] = -k * dif[i];
        } else {
          d->qfrc_spring[da] = -k * (d->qpos[qa] - m->qpos_spring[qa]);
        }
      }
      if (hd) for (int i = 0; i < n; i++) d->qfrc_damper[da + i] = -c * d->qvel[da + i];
    }
  }
  /* passive.py:246-272 _gravity_force: gravity compensation of bodies with gravcomp != 0, through the Jacobian of their CoM */
  int gravcomp = m->ngravcomp && !(o->disableflags & DSBL_GRAVITY) && !(dsbl_spring && dsbl_damper);
  if (gravcomp) {
    for (int b = 1; b < m->nbody; b++) {
      real gc = m->body_gravcomp[b];
      if (gc == 0.0) continue;
      real force[3] = {-o->gravity_x * m->body_mass[b] * gc, -o->gravity_y * m->body_mass[b] * gc, -o->gravity_z * m->body_mass[b] * gc};
      for (int dof = 0; dof < nv; dof++) {
        real jp[3], jr[3];
        jac_dof(m, d, d->xipos + 3 * b, b, dof, jp, jr);
        d->qfrc_gravcomp[dof] += dot3(jp, force);
      }
    }
  }
  for (int i = 0; i < nv; i++) { /* passive.py:536-565 */
    real q = (dsbl_spring && dsbl_damper) ? 0.0 : d->qfrc_spring[i] + d->qfrc_damper[i];
    if (gravcomp && !m->jnt_actgravcomp[m->dof_jntid[i]]) q += d->qfrc_gravcomp[i];
    d->qfrc_passive[i] = q;
  }
  /* rne (flg_acc = False) */
  for (int i = 0; i < 6; i++) d->cacc[i] = 0;
  if (!(o->disableflags & DSBL_GRAVITY)) { d->cacc[3] = -o->gravity_x; d->cacc[4] = -o->gravity_y; d->cacc[5] = -o->gravity_z; }
  for (int b = 1; b < nb; b++) {
    real acc[6];
    memcpy(acc, d->cacc + 6 * m->body_parentid[b], sizeof(acc));
    for (int k = 0; k < m->body_dofnum[b]; k++) {
      int dof = m->body_dofadr[b] + k;
      for (int i = 0; i < 6; i++) acc[i] += d->cdof_dot[6 * dof + i] * d->qvel[dof];
    }
    memcpy(d->cacc + 6 * b, acc, sizeof(acc));
  }
  for (int i = 0; i < 6; i++) d->cfrc_int[i] = 0;
  for (int b = 1; b < nb; b++) {
    real f[6], t[6], u[6];
    inert_vec(f, d->cinert + 10 * b, d->cacc + 6 * b);
    inert_vec(t, d->cinert + 10 * b, d->cvel + 6 * b);
    motion_cross_force(u, d->cvel + 6 * b, t);
    for (int i = 0; i < 6; i++) d->cfrc_int[6 * b + i] = f[i] + u[i];
  }
  for (int b = nb - 1; b > 0; b--) {
    int p = m->body_parentid[b];
    for (int i = 0; i < 6; i++) d->cfrc_int[6 * p + i] += d->cfrc_int[6 * b + i];
  }
  for (int i = 0; i < nv; i++) {
    real s = 0;
    for (int k = 0; k < 6; k++) s += d->cdof[6 * i + k] * d->cfrc_int[6 * m->dof_bodyid[i] + k];
    d->qfrc_bias[i] = s;
  }
}

/* support.next_act (support.py:38-64) */
static real next_act(real dt, int dyntype, const real* dynprm, const real* actrange, real act, real act_dot, real scale, int clamp) {
  if (dyntype == 3) {
    real tau = fmax(MJ_MINVAL, dynprm[0]);
    act = act + scale * act_dot * tau * (1.0 - exp(-dt / tau));
  } else if (dyntype == 5) {
    return act;
  } else {
    act = act + scale * act_dot * dt;
  }
  if (clamp) act = clampr(act, actrange[0], actrange[1]);
  return act;
}

/* forward.fwd_actuation (forward.py:617-927), no muscles / tendons */
static void o_fwd_actuation(const OModel* m, const OOpt* o, OData* d) {
  int nv = m->nv;
  if (!m->nu || (o->disableflags & DSBL_ACTUATION)) {
    for (int i = 0; i < m->na; i++) d->act_dot[i] = 0;
    for (int i = 0; i < nv; i++) d->qfrc_actuator[i] = 0;
    return;
  }
  for (int u = 0; u < m->nu; u++) {
    real ctrl = d->ctrl[u];
    if (m->actuator_ctrllimited[u] && !(o->disableflags & DSBL_CLAMPCTRL))
      ctrl = clampr(ctrl, m->actuator_ctrlrange[2 * u], m->actuator_ctrlrange[2 * u + 1]);
    real ctrl_act = ctrl;
    int af = m->actuator_actadr[u];
    if (m->na && af >= 0) {
      int al = af + m->actuator_actnum[u] - 1, dyn = m->actuator_dyntype[u];
      const real* dynprm = m->actuator_dynprm + 10 * u;
      real act_dot = 0, act = 0;
      if (dyn == 1) act_dot = ctrl;
      else if (dyn == 2 || dyn == 3) { act = d->act[al]; act_dot = (ctrl - act) / fmax(dynprm[0], MJ_MINVAL); }
      d->act_dot[al] = act_dot;
      if (m->actuator_actearly[u]) {
        if (dyn == 1 || dyn == 0) act = d->act[al];
        ctrl_act = next_act(o->timestep, dyn, dynprm, m->actuator_actrange + 2 * u, act, act_dot, 1.0, m->actuator_actlimited[u]);
      } else {
        ctrl_act = d->act[al];
      }
    }
    real len = d->actuator_length[u], vel = d->actuator_velocity[u], gain = 0, bias = 0;
    const real *gp = m->actuator_gainprm + 10 * u, *bp = m->actuator_biasprm + 10 * u;
    if (m->actuator_gaintype[u] == 0) gain = gp[0];
    else if (m->actuator_gaintype[u] == 1) gain = gp[0] + gp[1] * len + gp[2] * vel;
    if (m->actuator_biastype[u] == 1) bias = bp[0] + bp[1] * len + bp[2] * vel;
    real force = gain * ctrl_act + bias;
    if (m->actuator_forcelimited[u]) force = clampr(force, m->actuator_forcerange[2 * u], m->actuator_forcerange[2 * u + 1]);
    d->actuator_force[u] = force;
  }
  for (int i = 0; i < nv; i++) d->qfrc_actuator[i] = 0;
  for (int u = 0; u < m->nu; u++)
    for (int i = 0; i < d->moment_rownnz[u]; i++) {
      int s = d->moment_rowadr[u] + i;
      d->qfrc_actuator[d->moment_colind[s]] += d->actuator_moment[s] * d->actuator_force[u];
    }
  for (int i = 0; i < nv; i++) { /* forward.py:806-833 */
    int j = m->dof_jntid[i];
    real q = d->qfrc_actuator[i];
    if (m->ngravcomp && m->jnt_actgravcomp[j]) q += d->qfrc_gravcomp[i];
    if (m->jnt_actfrclimited[j]) q = clampr(q, m->jnt_actfrcrange[2 * j], m->jnt_actfrcrange[2 * j + 1]);
    d->qfrc_actuator[i] = q;
  }
}

/* forward.fwd_acceleration (forward.py:931-969) + support.apply_ft (support.py:175-216) */
static void o_fwd_acceleration(const OModel* m, OData* d, int factorize) {
  int nv = m->nv, nb = m->nbody;
  for (int i = 0; i < nv; i++) d->qfrc_smooth[i] = d->qfrc_passive[i] - d->qfrc_bias[i] + d->qfrc_actuator[i] + d->qfrc_applied[i];
  for (int i = 0; i < nv; i++) {
    const real* cd = d->cdof + 6 * i;
    int db = m->dof_bodyid[i];
    real acc = 0;
    for (int b = db; b < nb; b++) {
      const real* ft = d->xfrc_applied + 6 * b;
      if (ft[0] == 0 && ft[1] == 0 && ft[2] == 0 && ft[3] == 0 && ft[4] == 0 && ft[5] == 0) continue;
      int p = b;
      while (p != 0 && p != db) p = m->body_parentid[p];
      if (p == 0) continue;
      real off[3], c[3];
      for (int k = 0; k < 3; k++) off[k] = d->xipos[3 * b + k] - d->subtree_com[3 * m->body_rootid[b] + k];
      cross3(c, cd, off);
      acc += cd[3] * ft[0] + cd[4] * ft[1] + cd[5] * ft[2] + cd[0] * ft[3] + cd[1] * ft[4] + cd[2] * ft[5] + dot3(c, ft);
    }
    d->qfrc_smooth[i] += acc;
  }
  if (factorize) o_factor_m(m, d);
  chol_solve(d->qacc_smooth, d->qLD, nv, d->qfrc_smooth, nv);
}

/* ------------------------------------------------------------------------------------------ */
/* solver.solve (solver.py:3297): Newton / CG with exact iterative linesearch, pyramidal cones  */
/* ------------------------------------------------------------------------------------------ */
typedef struct { real c, g, h; } pt3;
static inline pt3 pt_add(pt3 a, pt3 b) { pt3 r = {a.c + b.c, a.g + b.g, a.h + b.h}; return r; }
/* solver.py:160 */
static inline pt3 eval_pt_direct(real jaref, real jv, real dd, real alpha) {
  real x = jaref + alpha * jv, jvD = jv * dd;
  pt3 r = {0.5 * dd * x * x, jvD * x, jv * jvD};
  return r;
}
/* solver.py:224 */
static inline pt3 eval_frictionloss_pt(real x, real f, real rf, real jv, real dd) {
  pt3 r;
  if (-rf < x && x < rf) { real jvD = jv * dd; r.c = 0.5 * dd * x * x; r.g = jvD * x; r.h = jv * jvD; }
  else if (x <= -rf) { r.c = f * (-0.5 * rf - x); r.g = -f * jv; r.h = 0; }
  else { r.c = f * (-0.5 * rf + x); r.g = f * jv; r.h = 0; }
  return r;
}
/* solver.py:571-598 */
static inline pt3 efc_eval_pt(int r, real alpha, int ne, int nf, real D, real floss, real jaref, real jv) {
  pt3 z = {0, 0, 0};
  if (r >= ne + nf) {
    real x = jaref + alpha * jv;
    if (x < 0.0) return eval_pt_direct(jaref, jv, D, alpha);
    return z;
  }
  if (r >= ne) {
    real x = jaref + alpha * jv, rf = safe_div(floss, D);
    return eval_frictionloss_pt(x, floss, rf, jv, D);
  }
  return eval_pt_direct(jaref, jv, D, alpha);
}
/* solver.py:198 */
static inline pt3 eval_pt_quad(const real* q, real alpha) {
  real aq2 = alpha * q[2];
  pt3 r = {alpha * aq2 + alpha * q[1] + q[0], 2.0 * aq2 + q[1], 2.0 * q[2]};
  return r;
}
static inline int in_bracket(pt3 x, pt3 y) { return (x.g < y.g && y.g < 0.0) || (x.g > y.g && y.g > 0.0); }

typedef struct {
  int nefc, ne, nf;
  real *Jaref, *jv, *grad, *Mgrad, *search, *mv, *H, *L, *prev_grad, *prev_Mgrad;
  real gauss, cost, prev_cost, search_dot, grad_dot;
  const OData* glob; /* pooled contact arrays (elliptic cones read friction / dim / efc_address) */
  int nmp, elliptic;
  real impratio_invsqrt;
} OSolver;

/* solver.py:263-319 _eval_elliptic: cost, gradient, curvature of one elliptic contact along the search direction.
 * quad = summed quadratic of its rows; q1 = (u0, v0, uu); q2 = (uv, vv, dm) as prepared at solver.py:986-1050. */
static pt3 eval_pt_quad(const real* q, real alpha);
static pt3 eval_elliptic(real mu, const real* quad, const real* q1, const real* q2, real alpha) {
  pt3 z = {0, 0, 0};
  real u0 = q1[0], v0 = q1[1], uu = q1[2], uv = q2[0], vv = q2[1], dm = q2[2];
  real N = u0 + alpha * v0, Tsqr = uu + alpha * (2.0 * uv + alpha * vv);
  if (Tsqr <= 0.0) return N < 0.0 ? eval_pt_quad(quad, alpha) : z;
  real T = sqrt(Tsqr);
  if (N >= mu * T) return z;
  if (mu * N + T <= 0.0) return eval_pt_quad(quad, alpha);
  real N1 = v0, T1 = (uv + alpha * vv) / T, T2 = vv / T - (uv + alpha * vv) * T1 / (T * T), e = N - mu * T, e1 = N1 - mu * T1;
  pt3 r = {0.5 * dm * e * e, dm * e * e1, dm * (e1 * e1 + e * (-mu * T2))};
  return r;
}
/* rows of one elliptic contact: returns dim and fills adr[dim]; 0 if any row is missing (njmax overflow) */
static int elliptic_rows(const OSolver* s, int conid, int* adr) {
  int dim = s->glob->contact_dim[conid];
  for (int j = 0; j < dim; j++) {
    adr[j] = s->glob->contact_efc_address[conid * s->nmp + j];
    if (adr[j] < 0) return 0;
  }
  return dim;
}
/* evaluation of row r of an elliptic contact at alpha (solver.py:602-628, 986-1050): only the primary row contributes */
static pt3 elliptic_row_eval(const OData* d, const OSolver* s, int r, real alpha) {
  pt3 z = {0, 0, 0};
  int conid = d->efc_id[r], adr[6];
  int dim = elliptic_rows(s, conid, adr);
  if (!dim || adr[0] != r) return z;
  const real* fri = s->glob->contact_friction + 5 * conid;
  real mu = fri[0] * s->impratio_invsqrt, D0 = d->efc_D[r], ja = s->Jaref[r], jv = s->jv[r];
  real quad[3] = {0.5 * ja * ja * D0, jv * D0 * ja, 0.5 * jv * jv * D0}, q1[3] = {ja * mu, jv * mu, 0}, q2[3] = {0, 0, D0 / (mu * mu * (1.0 + mu * mu))};
  for (int j = 1; j < dim; j++) {
    real jaj = s->Jaref[adr[j]], jvj = s->jv[adr[j]], dj = d->efc_D[adr[j]], uj = jaj * fri[j - 1], vj = jvj * fri[j - 1];
    quad[0] += 0.5 * jaj * dj * jaj; quad[1] += jvj * dj * jaj; quad[2] += 0.5 * jvj * dj * jvj;
    q1[2] += uj * uj; q2[0] += uj * vj; q2[1] += vj * vj;
  }
  return eval_elliptic(mu, quad, q1, q2, alpha);
}

/* total 1-D cost model at step alpha (sum over rows + Gauss quadratic) */
static pt3 ls_eval(const OData* d, const OSolver* s, const real* quad_gauss, real alpha) {
  pt3 t = eval_pt_quad(quad_gauss, alpha);
  for (int r = 0; r < s->nefc; r++) {
    if (s->elliptic && r >= s->ne + s->nf && d->efc_type[r] == CT_CONTACT_ELLIPTIC) t = pt_add(t, elliptic_row_eval(d, s, r, alpha));
    else t = pt_add(t, efc_eval_pt(r, alpha, s->ne, s->nf, d->efc_D[r], d->efc_frictionloss[r], s->Jaref[r], s->jv[r]));
  }
  return t;
}

/* solver.py:887-1341 linesearch_iterative (pyramidal) */
static real o_linesearch(const OModel* m, const OOpt* o, OData* d, OSolver* s) {
  int nv = m->nv, np = m->nv_pad;
  o_mul_m(m, d, s->mv, s->search);
  for (int r = 0; r < s->nefc; r++) {
    real v = 0;
    for (int i = 0; i < nv; i++) v += d->efc_J[(size_t)r * np + i] * s->search[i];
    s->jv[r] = v;
  }
  real snorm = sqrt(s->search_dot), scale = o->meaninertia * (real)nv;
  real gtol = fmax(o->tolerance * o->ls_tolerance * snorm * scale, 1e-6);
  real qg[3] = {s->gauss, 0, 0};
  for (int i = 0; i < nv; i++) {
    qg[1] += s->search[i] * (d->efc_Ma[i] - d->qfrc_smooth[i]);
    qg[2] += 0.5 * s->search[i] * s->mv[i];
  }
  pt3 p0 = {qg[0], qg[1], 2.0 * qg[2]};
  for (int r = 0; r < s->nefc; r++) {
    if (s->elliptic && r >= s->ne + s->nf && d->efc_type[r] == CT_CONTACT_ELLIPTIC) p0 = pt_add(p0, elliptic_row_eval(d, s, r, 0.0));
    else p0 = pt_add(p0, efc_eval_pt(r, 0.0, s->ne, s->nf, d->efc_D[r], d->efc_frictionloss[r], s->Jaref[r], s->jv[r]));
  }
  real lo_alpha_in = -safe_div(p0.g, p0.h);
  pt3 lo_in = ls_eval(d, s, qg, lo_alpha_in);
  real alpha;
  if (fabs(lo_in.g) < gtol && lo_in.c < p0.c) {
    alpha = lo_alpha_in;
  } else {
    alpha = 0.0;
    int lo_less = lo_in.g < p0.g;
    pt3 lo = lo_less ? lo_in : p0, hi = lo_less ? p0 : lo_in;
    real lo_alpha = lo_less ? lo_alpha_in : 0.0, hi_alpha = lo_less ? 0.0 : lo_alpha_in;
    for (int it = 0; it < o->ls_iterations; it++) {
      real lo_next_alpha = lo_alpha - safe_div(lo.g, lo.h);
      real hi_next_alpha = hi_alpha - safe_div(hi.g, hi.h);
      real mid_alpha = 0.5 * (lo_alpha + hi_alpha);
      pt3 lo_next = ls_eval(d, s, qg, lo_next_alpha), hi_next = ls_eval(d, s, qg, hi_next_alpha), mid = ls_eval(d, s, qg, mid_alpha);
      int s1 = in_bracket(lo, lo_next);
      if (s1) { lo = lo_next; lo_alpha = lo_next_alpha; }
      int s2 = in_bracket(lo, mid);
      if (s2) { lo = mid; lo_alpha = mid_alpha; }
      int s3 = in_bracket(lo, hi_next);
      if (s3) { lo = hi_next; lo_alpha = hi_next_alpha; }
      int swap_lo = s1 || s2 || s3;
      int h1 = in_bracket(hi, hi_next);
      if (h1) { hi = hi_next; hi_alpha = hi_next_alpha; }
      int h2 = in_bracket(hi, mid);
      if (h2) { hi = mid; hi_alpha = mid_alpha; }
      int h3 = in_bracket(hi, lo_next);
      if (h3) { hi = lo_next; hi_alpha = lo_next_alpha; }
      int swap_hi = h1 || h2 || h3;
      int ls_done = (!swap_lo && !swap_hi) || (lo.g < 0.0 && lo.g > -gtol) || (hi.g > 0.0 && hi.g < gtol);
      int improved = lo.c < p0.c || hi.c < p0.c, lo_better = lo.c < hi.c;
      if (improved && lo_better) alpha = lo_alpha;
      if (improved && !lo_better) alpha = hi_alpha;
      if (ls_done) break;
    }
  }
  for (int i = 0; i < nv; i++) { d->qacc[i] += alpha * s->search[i]; d->efc_Ma[i] += alpha * s->mv[i]; }
  for (int r = 0; r < s->nefc; r++) s->Jaref[r] += alpha * s->jv[r];
  return alpha;
}

/* solver.py:2154-2219 _update_constraint (pyramidal / non-elliptic rows) */
static void o_update_constraint(const OModel* m, OData* d, OSolver* s) {
  int nv = m->nv, np = m->nv_pad;
  s->prev_cost = s->cost;
  real cost = 0;
  for (int r = 0; r < s->nefc; r++) {
    real D = d->efc_D[r], ja = s->Jaref[r];
    int st;
    if (r < s->ne) { d->efc_force[r] = -D * ja; st = ST_QUADRATIC; cost += 0.5 * D * ja * ja; }
    else if (r < s->ne + s->nf) {
      real f = d->efc_frictionloss[r], rf = safe_div(f, D);
      if (ja <= -rf) { d->efc_force[r] = f; st = ST_LINEARNEG; cost += -f * (0.5 * rf + ja); }
      else if (ja >= rf) { d->efc_force[r] = -f; st = ST_LINEARPOS; cost += -f * (0.5 * rf - ja); }
      else { d->efc_force[r] = -D * ja; st = ST_QUADRATIC; cost += 0.5 * D * ja * ja; }
    } else if (!(s->elliptic && d->efc_type[r] == CT_CONTACT_ELLIPTIC)) {
      if (ja >= 0.0) { d->efc_force[r] = 0; st = ST_SATISFIED; }
      else { d->efc_force[r] = -D * ja; st = ST_QUADRATIC; cost += 0.5 * D * ja * ja; }
    } else { /* elliptic friction cone contact, solver.py:1886-1943 */
      int conid = d->efc_id[r], adr[6];
      int dim = elliptic_rows(s, conid, adr);
      if (!dim) continue;
      const real* fri = s->glob->contact_friction + 5 * conid;
      real mu = fri[0] * s->impratio_invsqrt, N = s->Jaref[adr[0]] * mu, TT = 0, ufr = 0;
      for (int j = 1; j < dim; j++) {
        real uj = s->Jaref[adr[j]] * fri[j - 1];
        TT += uj * uj;
        if (adr[j] == r) ufr = uj * fri[j - 1];
      }
      real T = TT <= 0.0 ? 0.0 : sqrt(TT);
      if (N >= mu * T || (T <= 0.0 && N >= 0.0)) { d->efc_force[r] = 0; st = ST_SATISFIED; }
      else if (mu * N + T <= 0.0 || (T <= 0.0 && N < 0.0)) { d->efc_force[r] = -D * ja; st = ST_QUADRATIC; cost += 0.5 * D * ja * ja; }
      else {
        real dm = safe_div(d->efc_D[adr[0]], mu * mu * (1.0 + mu * mu)), nmt = N - mu * T, force = -dm * nmt * mu;
        if (r == adr[0]) { d->efc_force[r] = force; cost += 0.5 * dm * nmt * nmt; }
        else d->efc_force[r] = -safe_div(force, T) * ufr;
        st = ST_CONE;
      }
    }
    d->efc_state[r] = st;
  }
  for (int i = 0; i < nv; i++) {
    real q = 0;
    for (int r = 0; r < s->nefc; r++) q += d->efc_J[(size_t)r * np + i] * d->efc_force[r];
    d->qfrc_constraint[i] = q;
  }
  real g = 0;
  for (int i = 0; i < nv; i++) g += (d->efc_Ma[i] - d->qfrc_smooth[i]) * (d->qacc[i] - d->qacc_smooth[i]);
  s->gauss = 0.5 * g;
  s->cost = cost + 0.5 * g;
}

/* solver.py:2879-3008 _update_gradient (full Hessian rebuild every iteration; the reference's
 * incremental update, solver.py:3011, is algebraically the same matrix) */
static void o_update_gradient(const OModel* m, const OOpt* o, OData* d, OSolver* s) {
  int nv = m->nv, np = m->nv_pad;
  real gd = 0;
  for (int i = 0; i < nv; i++) {
    s->grad[i] = d->efc_Ma[i] - d->qfrc_smooth[i] - d->qfrc_constraint[i];
    gd += s->grad[i] * s->grad[i];
  }
  s->grad_dot = gd;
  if (o->solver == SOLVER_CG) {
    chol_solve(s->Mgrad, d->qLD, nv, s->grad, nv);
  } else {
    for (int i = 0; i < nv; i++)
      for (int j = 0; j < nv; j++) s->H[i * nv + j] = d->qM[i * np + j];
    for (int r = 0; r < s->nefc; r++) {
      if (d->efc_state[r] != ST_QUADRATIC) continue;
      const real* J = d->efc_J + (size_t)r * np;
      real D = d->efc_D[r];
      for (int i = 0; i < nv; i++) {
        if (J[i] == 0) continue;
        for (int j = 0; j <= i; j++) s->H[i * nv + j] += D * J[i] * J[j];
      }
    }
    if (s->elliptic) { /* solver.py:2592-2730 update_gradient_JTCJ_dense: cone Hessian of contacts in the middle zone */
      for (int k = 0; k < d->wcon_n[0]; k++) {
        int conid = d->wcon[k], adr[6];
        int dim = s->glob->contact_dim[conid];
        if (dim == 1 || !(s->glob->contact_dist[conid] - s->glob->contact_includemargin[conid] < 0.0)) continue;
        if (!(s->glob->contact_type[conid] & 1)) continue;
        if (!elliptic_rows(s, conid, adr) || adr[0] >= s->nefc || d->efc_state[adr[0]] != ST_CONE) continue;
        const real* fri = s->glob->contact_friction + 5 * conid;
        real mu = fri[0] * s->impratio_invsqrt, mu2 = mu * mu, dm = safe_div(d->efc_D[adr[0]], mu2 * (1.0 + mu2));
        if (dm == 0.0) continue;
        real n = s->Jaref[adr[0]] * mu, u[6] = {n, 0, 0, 0, 0, 0}, tt = 0, C[6][6];
        for (int j = 1; j < dim; j++) { u[j] = s->Jaref[adr[j]] * fri[j - 1]; tt += u[j] * u[j]; }
        real t = fmax(tt <= 0.0 ? 0.0 : sqrt(tt), MJ_MINVAL), ttt = fmax(t * t * t, MJ_MINVAL);
        for (int a = 0; a < dim; a++) {
          for (int b = 0; b < dim; b++) {
            real h;
            if (a == 0 && b == 0) h = 1.0;
            else if (a == 0) h = -safe_div(mu, t) * u[b];
            else if (b == 0) h = -safe_div(mu, t) * u[a];
            else { h = mu * safe_div(n, ttt) * u[a] * u[b]; if (a == b) h += mu2 - mu * safe_div(n, t); }
            C[a][b] = h * dm * (a == 0 ? mu : fri[a - 1]) * (b == 0 ? mu : fri[b - 1]);
          }
        }
        for (int i = 0; i < nv; i++) {
          for (int j = 0; j <= i; j++) {
            real h = 0;
            for (int a = 0; a < dim; a++)
              for (int b = 0; b < dim; b++) h += C[a][b] * d->efc_J[(size_t)adr[a] * np + i] * d->efc_J[(size_t)adr[b] * np + j];
            s->H[i * nv + j] += h;
          }
        }
      }
    }
    chol_factor(s->L, nv, s->H, nv, nv);
    chol_solve(s->Mgrad, s->L, nv, s->grad, nv);
  }
}

static void o_solve(const OModel* m, const OOpt* o, OData* d, const OData* glob) {
  int nv = m->nv, np = m->nv_pad, njmax = glob->njmax;
  if (njmax == 0 || nv == 0) {
    memcpy(d->qacc, d->qacc_smooth, sizeof(real) * nv);
    d->solver_niter[0] = 0;
    return;
  }
  OSolver s;
  memset(&s, 0, sizeof(s));
  s.nefc = d->nefc[0] < njmax ? d->nefc[0] : njmax;
  s.ne = d->ne[0]; s.nf = d->nf[0];
  s.glob = glob; s.nmp = m->nmaxpyramid; s.elliptic = o->cone == 1; s.impratio_invsqrt = o->impratio_invsqrt;
  real* ws = (real*)calloc((size_t)(2 * njmax + 8 * nv + 2 * nv * nv + 16), sizeof(real));
  s.Jaref = ws; s.jv = s.Jaref + njmax; s.grad = s.jv + njmax; s.Mgrad = s.grad + nv; s.search = s.Mgrad + nv;
  s.mv = s.search + nv; s.prev_grad = s.mv + nv; s.prev_Mgrad = s.prev_grad + nv; s.H = s.prev_Mgrad + nv; s.L = s.H + nv * nv;
  memcpy(d->qacc, (o->disableflags & DSBL_WARMSTART) ? d->qacc_smooth : d->qacc_warmstart, sizeof(real) * nv);
  /* init_context, solver.py:3257 */
  s.cost = MJ_MAXVAL;
  d->solver_niter[0] = 0;
  for (int r = 0; r < s.nefc; r++) {
    real v = 0;
    for (int i = 0; i < nv; i++) v += d->efc_J[(size_t)r * np + i] * d->qacc[i];
    s.Jaref[r] = v - d->efc_aref[r];
  }
  o_mul_m(m, d, d->efc_Ma, d->qacc);
  o_update_constraint(m, d, &s);
  o_update_gradient(m, o, d, &s);
  s.search_dot = 0;
  for (int i = 0; i < nv; i++) { s.search[i] = -s.Mgrad[i]; s.search_dot += s.search[i] * s.search[i]; }
  real scl = 1.0 / (o->meaninertia * (real)nv);
  for (int it = 0; it < o->iterations; it++) { /* _solver_iteration, solver.py:3188 */
    o_linesearch(m, o, d, &s);
    if (o->solver == SOLVER_CG) { memcpy(s.prev_grad, s.grad, sizeof(real) * nv); memcpy(s.prev_Mgrad, s.Mgrad, sizeof(real) * nv); }
    o_update_constraint(m, d, &s);
    o_update_gradient(m, o, d, &s);
    real beta = 0;
    if (o->solver == SOLVER_CG) { /* solver.py:3081 */
      real num = 0, den = 0;
      for (int i = 0; i < nv; i++) { num += s.grad[i] * (s.Mgrad[i] - s.prev_Mgrad[i]); den += s.prev_grad[i] * s.prev_Mgrad[i]; }
      beta = fmax(0.0, num / fmax(MJ_MINVAL, den));
    }
    s.search_dot = 0;
    for (int i = 0; i < nv; i++) {
      real v = -s.Mgrad[i];
      if (o->solver == SOLVER_CG) v += beta * s.search[i];
      s.search[i] = v;
      s.search_dot += v * v;
    }
    d->solver_niter[0]++;
    real improvement = (s.prev_cost - s.cost) * scl, gradient = sqrt(s.grad_dot) * scl;
    if (improvement < o->tolerance || gradient < o->tolerance) break;
  }
  free(ws);
}

/* forward.euler (forward.py:327-354) + _advance (:213-274) */
static void o_advance(const OModel* m, const OOpt* o, OData* d, const real* qacc) {
  real dt = o->timestep;
  for (int u = 0; u < m->nu; u++)
    for (int j = m->actuator_actadr[u]; j >= 0 && j < m->actuator_actadr[u] + m->actuator_actnum[u]; j++)
      d->act[j] = next_act(dt, m->actuator_dyntype[u], m->actuator_dynprm + 10 * u, m->actuator_actrange + 2 * u, d->act[j],
                           d->act_dot[j], 1.0, m->actuator_actlimited[u]);
  for (int i = 0; i < m->nv; i++) d->qvel[i] += qacc[i] * dt;
  for (int j = 0; j < m->njnt; j++) { /* forward.py:52-112 */
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j], jt = m->jnt_type[j];
    if (jt == JNT_FREE) {
      for (int i = 0; i < 3; i++) d->qpos[qa + i] += dt * d->qvel[da + i];
      quat_integrate(d->qpos + qa + 3, d->qvel + da + 3, dt);
    } else if (jt == JNT_BALL) {
      quat_integrate(d->qpos + qa, d->qvel + da, dt);
    } else {
      d->qpos[qa] += dt * d->qvel[da];
    }
  }
  d->time[0] += dt;
  memcpy(d->qacc_warmstart, d->qacc, sizeof(real) * m->nv);
}
static void o_euler(const OModel* m, const OOpt* o, OData* d) {
  int nv = m->nv, np = m->nv_pad;
  if (!(o->disableflags & (DSBL_EULERDAMP | DSBL_DAMPER))) {
    real* A = (real*)malloc(sizeof(real) * (2 * nv * nv + nv));
    real *L = A + nv * nv, *qacc = L + nv * nv;
    for (int i = 0; i < nv; i++)
      for (int j = 0; j < nv; j++) A[i * nv + j] = d->qM[i * np + j] + (i == j ? o->timestep * m->dof_damping[i] : 0.0);
    chol_factor(L, nv, A, nv, nv);
    chol_solve(qacc, L, nv, d->efc_Ma, nv);
    o_advance(m, o, d, qacc);
    free(A);
  } else {
    o_advance(m, o, d, d->qacc);
  }
}

/* derivative.py:31-108 _qderiv_actuator_passive_vel: d(actuator force)/d(actuator velocity) */
static real act_vel_deriv(const OModel* m, const OOpt* o, const OData* d, int u) {
  real gain = m->actuator_gaintype[u] == 1 ? m->actuator_gainprm[10 * u + 2] : 0.0;
  real bias = m->actuator_biastype[u] == 1 ? m->actuator_biasprm[10 * u + 2] : 0.0;
  if (bias == 0.0 && gain == 0.0) return 0.0;
  if (m->actuator_forcelimited[u]) { /* clamped force has no velocity dependence */
    real f = d->actuator_force[u];
    if (f <= m->actuator_forcerange[2 * u] || f >= m->actuator_forcerange[2 * u + 1]) return 0.0;
  }
  real vel = bias;
  if (m->actuator_dyntype[u] != 0) {
    if (gain != 0.0) {
      int adr = m->actuator_actadr[u] + m->actuator_actnum[u] - 1;
      real act = d->act[adr];
      if (m->actuator_actearly[u])
        act = next_act(o->timestep, m->actuator_dyntype[u], m->actuator_dynprm + 10 * u, m->actuator_actrange + 2 * u, d->act[adr],
                       d->act_dot[adr], 1.0, m->actuator_actlimited[u]);
      vel += gain * act;
    }
  } else if (gain != 0.0) {
    vel += gain * d->ctrl[u];
  }
  return vel;
}
/* forward.implicit (forward.py:495-512) + derivative.deriv_smooth_vel dense (derivative.py:321-420):
 * (qM - dt * qDeriv) qacc' = efc_Ma, qDeriv = moment^T diag(dforce/dvel) moment - diag(damping) */
static void o_implicit(const OModel* m, const OOpt* o, OData* d) {
  int nv = m->nv, np = m->nv_pad;
  int all = DSBL_ACTUATION | DSBL_SPRING | DSBL_DAMPER;
  if ((o->disableflags & all) == all) { o_advance(m, o, d, d->qacc); return; }
  real* A = (real*)malloc(sizeof(real) * (2 * nv * nv + nv));
  real *L = A + nv * nv, *qacc = L + nv * nv;
  for (int i = 0; i < nv; i++)
    for (int j = 0; j < nv; j++) A[i * nv + j] = d->qM[i * np + j];
  if (m->nu > 0 && !(o->disableflags & DSBL_ACTUATION))
    for (int u = 0; u < m->nu; u++) {
      real vel = act_vel_deriv(m, o, d, u);
      if (vel == 0.0) continue;
      int adr = d->moment_rowadr[u], nnz = d->moment_rownnz[u];
      for (int a = 0; a < nnz; a++)
        for (int b = 0; b < nnz; b++)
          A[d->moment_colind[adr + a] * nv + d->moment_colind[adr + b]] -= o->timestep * d->actuator_moment[adr + a] * d->actuator_moment[adr + b] * vel;
    }
  if (!(o->disableflags & DSBL_DAMPER))
    for (int i = 0; i < nv; i++) A[i * nv + i] += o->timestep * m->dof_damping[i];
  chol_factor(L, nv, A, nv, nv);
  chol_solve(qacc, L, nv, d->efc_Ma, nv);
  o_advance(m, o, d, qacc);
  free(A);
}

/* ------------------------------------------------------------------------------------------ */
/* entry point                                                                                 */
/* ------------------------------------------------------------------------------------------ */
const char* mjo_field_name(int kind, int index) {
  static const char* k0[] = {
#define X(n) #n,
      MJB_MODEL_SIZES(X) 0};
  static const char* k1[] = {MJB_MODEL_INT_FIELDS(X) 0};
  static const char* k2[] = {MJB_MODEL_REAL_FIELDS(X) 0};
  static const char* k3[] = {MJB_OPT_INT_FIELDS(X) 0};
  static const char* k4[] = {MJB_OPT_REAL_FIELDS(X) 0};
  static const char* k5[] = {MJB_DATA_SIZES(X) 0};
#undef X
#define X(n, c) #n,
  static const char* k6[] = {MJB_DATA_INT_FIELDS(X) 0};
  static const char* k7[] = {MJB_DATA_REAL_FIELDS(X) 0};
#undef X
  const char** t[] = {k0, k1, k2, k3, k4, k5, k6, k7};
  if (kind < 0 || kind > 7) return 0;
  for (int i = 0; i <= index; i++)
    if (!t[kind][i]) return 0;
  return t[kind][index];
}

/* Runs `stages` for worlds [w0, w1) on host arrays (float64 / int32, same logical layout as
 * the device Data).  nthread > 1 uses OpenMP over worlds (cpu_baseline only). */
int mjo_run_stages(const int* msizes, const int* const* mint, const double* const* mreal, const int* oint,
                   const double* oreal, const int* dsizes, int* const* dint, double* const* dreal, uint32_t stages,
                   int w0, int w1, int nthread) {
  OModel m;
  OOpt o;
  OData d;
  int k;
  k = 0;
#define X(n) m.n = msizes[k++];
  MJB_MODEL_SIZES(X)
#undef X
  k = 0;
#define X(n) m.n = mint[k++];
  MJB_MODEL_INT_FIELDS(X)
#undef X
  k = 0;
#define X(n) m.n = mreal[k++];
  MJB_MODEL_REAL_FIELDS(X)
#undef X
  k = 0;
#define X(n) o.n = oint[k++];
  MJB_OPT_INT_FIELDS(X)
#undef X
  k = 0;
#define X(n) o.n = oreal[k++];
  MJB_OPT_REAL_FIELDS(X)
#undef X
  k = 0;
#define X(n) d.n = dsizes[k++];
  MJB_DATA_SIZES(X)
#undef X
  k = 0;
#define X(n, c) d.n = dint[k++];
  MJB_DATA_INT_FIELDS(X)
#undef X
  k = 0;
#define X(n, c) d.n = dreal[k++];
  MJB_DATA_REAL_FIELDS(X)
#undef X
  if (m.is_sparse) return -2;
  if ((stages & MJB_STAGE_COLLISION) && o.run_collision_detection) { d.nacon[0] = 0; d.ncollision[0] = 0; }
  int rc = 0;
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthread) if (nthread > 1)
  for (int w = w0; w < w1; w++) {
    OData v;
    world_view(&m, &d, w, &v);
    if (stages & MJB_STAGE_KINEMATICS) o_kinematics(&m, &v);
    if (stages & MJB_STAGE_COM_POS) o_com_pos(&m, &v);
    if (stages & MJB_STAGE_CRB) o_crb(&m, &v);
    if (stages & MJB_STAGE_FACTOR_M) o_factor_m(&m, &v);
    if ((stages & MJB_STAGE_COLLISION) && o.run_collision_detection)
      if (o_collision(&m, &o, &v, &d, w) < 0) rc = -3;
    if (stages & MJB_STAGE_MAKE_CONSTRAINT) o_make_constraint(&m, &o, &v, &d);
    if (stages & MJB_STAGE_TRANSMISSION) o_transmission(&m, &v);
    if (stages & MJB_STAGE_ZERO_SENSOR) { /* forward.py:978-984: sensordata / energy zeroing, energy_pos when enabled */
      v.energy[0] = v.energy[1] = 0;
      if (o.enableflags & ENBL_ENERGY) o_energy_pos(&m, &o, &v);
    }
    if (stages & MJB_STAGE_FWD_VELOCITY) o_fwd_velocity(&m, &o, &v);
    if ((stages & MJB_STAGE_ZERO_SENSOR) && (stages & MJB_STAGE_FWD_VELOCITY) && (o.enableflags & ENBL_ENERGY)) o_energy_vel(&m, &v); /* :989-991 */
    if (stages & MJB_STAGE_FWD_ACTUATION) o_fwd_actuation(&m, &o, &v);
    if (stages & MJB_STAGE_FWD_ACCELERATION) o_fwd_acceleration(&m, &v, 0);
    if (stages & MJB_STAGE_FWD_ACC_FACTOR) o_fwd_acceleration(&m, &v, 1);
    if (stages & MJB_STAGE_SOLVE) o_solve(&m, &o, &v, &d);
    if (stages & MJB_STAGE_EULER) o_euler(&m, &o, &v);
    if (stages & MJB_STAGE_IMPLICIT) o_implicit(&m, &o, &v);
  }
  return rc;
}

/* helpers exported for the known-answer tests (reference math_test.py / collision_primitive_core_test.py) */
void mjo_make_frame(const double* a, double* frame) { make_frame(frame, a); }
/* test hook: narrowphase candidates of one geom pair given directly (tests/golden/primitive_core.npz).  geom 1 and 2
 * are ordered by type as in collision_driver.py:351.  Returns the candidate count; out arrays hold 8 entries. */
int mjo_narrowphase_pair(int t1, int t2, const double* p1, const double* r1, const double* s1, const double* p2, const double* r2,
                         const double* s2, double margin, double* dist, double* pos, double* frame) {
  OModel m;
  OData d;
  memset(&m, 0, sizeof(m));
  memset(&d, 0, sizeof(d));
  int gtype[2] = {t1, t2};
  real size[6], xpos[6], xmat[18];
  memcpy(size, s1, 3 * sizeof(real)); memcpy(size + 3, s2, 3 * sizeof(real));
  memcpy(xpos, p1, 3 * sizeof(real)); memcpy(xpos + 3, p2, 3 * sizeof(real));
  memcpy(xmat, r1, 9 * sizeof(real)); memcpy(xmat + 9, r2, 9 * sizeof(real));
  m.geom_type = gtype; m.geom_size = size;
  d.geom_xpos = xpos; d.geom_xmat = xmat;
  OContactGeom out[8];
  memset(out, 0, sizeof(out));
  int n = narrowphase_pair(&m, &d, 0, 1, margin, 1e-6, 35, 0, out);
  for (int k = 0; k < n && k < 8; k++) {
    dist[k] = out[k].dist;
    memcpy(pos + 3 * k, out[k].pos, 3 * sizeof(real));
    memcpy(frame + 9 * k, out[k].frame, 9 * sizeof(real));
  }
  return n;
}
/* test hook: collision_gjk.py:2200 ccd on one geom pair (tests/golden/ccd_pairs.npz); returns ncollision */
static _Thread_local CcdPolytope g_dbg_pt;
/* test hook: the polytope left behind by the last mjo_ccd call of this thread */
int mjo_ccd_polytope(double* vert, int* face, double* norm2, int* nvert) {
  memcpy(vert, g_dbg_pt.vert, sizeof(g_dbg_pt.vert));
  memcpy(face, g_dbg_pt.face, sizeof(g_dbg_pt.face));
  memcpy(norm2, g_dbg_pt.face_norm2, sizeof(g_dbg_pt.face_norm2));
  *nvert = g_dbg_pt.nvert;
  return g_dbg_pt.nface;
}
int mjo_ccd_mesh(int t1, int t2, const double* p1, const double* r1, const double* s1, const double* v1, int nv1, const double* p2,
                 const double* r2, const double* s2, const double* v2, int nv2, double margin, double cutoff, double tolerance, int iterations,
                 double* dist, double* x1, double* x2) {
  CcdGeom g1, g2;
#define pt g_dbg_pt
  CcdGjk res;
  memcpy(g1.pos, p1, sizeof(g1.pos)); memcpy(g1.rot, r1, sizeof(g1.rot)); memcpy(g1.size, s1, sizeof(g1.size));
  memcpy(g2.pos, p2, sizeof(g2.pos)); memcpy(g2.rot, r2, sizeof(g2.rot)); memcpy(g2.size, s2, sizeof(g2.size));
  g1.margin = g2.margin = margin; g1.type = t1; g2.type = t2;
  g1.vert = v1; g1.nvert = nv1; g2.vert = v2; g2.nvert = nv2;
  g1.graph = g2.graph = 0;
  x1[0] = x1[1] = x1[2] = x2[0] = x2[1] = x2[2] = 0.0;
  int face_idx;
  return ccd_run(tolerance, cutoff, iterations, iterations, g1, g2, &pt, &res, dist, x1, x2, &face_idx);
#undef pt
}
int mjo_ccd(int t1, int t2, const double* p1, const double* r1, const double* s1, const double* p2, const double* r2, const double* s2,
            double margin, double cutoff, double tolerance, int iterations, double* dist, double* x1, double* x2) {
  return mjo_ccd_mesh(t1, t2, p1, r1, s1, 0, 0, p2, r2, s2, 0, 0, margin, cutoff, tolerance, iterations, dist, x1, x2);
}
/* test hook: solver.py:263 _eval_elliptic -> (cost, gradient, curvature) */
void mjo_eval_elliptic(double mu, const double* quad, const double* q1, const double* q2, double alpha, double* out) {
  pt3 r = eval_elliptic(mu, quad, q1, q2, alpha);
  out[0] = r.c; out[1] = r.g; out[2] = r.h;
}
void mjo_quat_integrate(double* q, const double* v, double dt) { quat_integrate(q, v, dt); }
void mjo_closest_segment_point(const double* a, const double* b, const double* pt, double* out) { closest_segment_point(out, a, b, pt); }
void mjo_mul_m(int nv, int nv_pad, const double* qM, const double* vec, double* res) {
  for (int i = 0; i < nv; i++) {
    double s = 0;
    for (int j = 0; j < nv; j++) s += qM[i * nv_pad + j] * vec[j];
    res[i] = s;
  }
}
void mjo_solve_m(int nv, const double* qLD, const double* y, double* x) { chol_solve(x, qLD, nv, y, nv); }

/* math.py:285-314 closest_segment_to_segment_points (pinned by the reference's math_test.py:27-95) */
void mjo_closest_segment_to_segment_points(const double* a0, const double* a1, const double* b0, const double* b1, double* best_a,
                                           double* best_b) {
  double da[3], db[3], am[3], bm[3], tr[3], la, lb;
  for (int i = 0; i < 3; i++) { da[i] = a1[i] - a0[i]; db[i] = b1[i] - b0[i]; }
  la = sqrt(da[0] * da[0] + da[1] * da[1] + da[2] * da[2]);
  lb = sqrt(db[0] * db[0] + db[1] * db[1] + db[2] * db[2]);
  if (la != 0.0) for (int i = 0; i < 3; i++) da[i] /= la; /* math.py:261-265 */
  if (lb != 0.0) for (int i = 0; i < 3; i++) db[i] /= lb;
  const double ha = 0.5 * la, hb = 0.5 * lb;
  for (int i = 0; i < 3; i++) { am[i] = a0[i] + da[i] * ha; bm[i] = b0[i] + db[i] * hb; tr[i] = am[i] - bm[i]; }
  const double ab = da[0] * db[0] + da[1] * db[1] + da[2] * db[2];
  const double at = da[0] * tr[0] + da[1] * tr[1] + da[2] * tr[2];
  const double bt = db[0] * tr[0] + db[1] * tr[1] + db[2] * tr[2];
  const double denom = 1.0 - ab * ab;
  const double ota = (-at + ab * bt) / (denom + 1e-6), otb = bt + ota * ab;
  const double ta = ota < -ha ? -ha : (ota > ha ? ha : ota), tb = otb < -hb ? -hb : (otb > hb ? hb : otb);
  double pa[3], pb[3], na[3], nb[3], d1 = 0.0, d2 = 0.0;
  for (int i = 0; i < 3; i++) { pa[i] = am[i] + da[i] * ta; pb[i] = bm[i] + db[i] * tb; }
  closest_segment_point(na, a0, a1, pb);
  closest_segment_point(nb, b0, b1, pa);
  for (int i = 0; i < 3; i++) { d1 += (pb[i] - na[i]) * (pb[i] - na[i]); d2 += (pa[i] - nb[i]) * (pa[i] - nb[i]); }
  if (d1 < d2) { memcpy(best_a, na, sizeof(na)); memcpy(best_b, pb, sizeof(pb)); }
  else { memcpy(best_a, pa, sizeof(pa)); memcpy(best_b, nb, sizeof(nb)); }
}
/* math.py:322-333 (pinned by math_test.py:97-126) */
int mjo_upper_tri_index(int n, int i, int j) { return (i * (2 * n - i - 3)) / 2 + j - 1; }
int mjo_upper_trid_index(int n, int i, int j) {
  if (j < i) { int t = i; i = j; j = t; }
  return (i * (2 * n - i - 1)) / 2 + j;
}
