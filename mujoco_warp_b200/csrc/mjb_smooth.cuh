// mjb_smooth.cuh -- position / velocity / actuation / acceleration stages, one warp per world.
//
// Replaces the reference's per-level kernel fan-out with float atomics (reference
// mujoco_warp/_src/smooth.py:358 kinematics, :602 com_pos, :889 crb, :1104 factor_m, :2015 com_vel,
// :1259 rne, :2605 transmission; forward.py:593-969) by warp-cooperative device functions:
// tree recursions are level-synchronous (__syncwarp between depth levels), subtree sums use the
// depth-first id ranges [b, subtreeend[b]) so that no atomics are needed and results are
// deterministic.
#pragma once
#ifndef MJB_SOL_COMPACT
#define MJB_SOL_COMPACT 1
#endif


#include "mjb_types.cuh"

// Lanes per world of the stage functions in this header.  32: one warp per world.  8 / 16: a warp carries 4 / 2 worlds, `lane` is
// the lane inside the world's group and every loop strides by MJB_LW.  The tree recursions keep 2-3 lanes busy per level whatever
// the group width, so narrower groups execute several worlds for the price of one; control flow in these stages depends on the
// model only, so the groups of a warp stay convergent and __syncwarp() keeps its full mask.  Set per translation unit.
// Measured on B200 (humanoid, 8192 worlds), 8 lanes per world and 7 warps (28 worlds) per CTA against one warp per world:
// smooth_position 69 -> 80 us, smooth_velocity 111 -> 95 us, euler 20 -> 17 us; 16 lanes: 115 us; 4 or 8 warps per CTA: 153 / 132 us.
// The stages are chains of dependent global-memory round trips (parent poses, subtree sums, ancestor walks): the narrower groups
// save instructions but leave fewer warps in flight to hide those latencies, so the gain depends on the stage.  Defaults
// (mjb_types.cuh): smooth_velocity and euler 8 lanes per world, smooth_position 32.  panda (8192 worlds): 29.0 -> 37.4 M env-steps/s.
#ifndef MJB_LW
#define MJB_LW 32
#endif
static_assert(MJB_LW == 32 || MJB_LW == 16 || MJB_LW == 8, "lanes per world: 8, 16 or 32");
// sum over the lanes of a world's group
__device__ __forceinline__ float lw_sum(float v) {
#pragma unroll
  for (int o = MJB_LW / 2; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}

// ---------------------------------------------------------------------------------------------
// smooth.kinematics (smooth.py:45-224)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void s_kinematics(const KParams& K, const unsigned char* mb, int w, int lane) {
  const float* qpos = DW(qpos);
  float* xpos_o = DW(xpos);
  float* xquat_o = DW(xquat);
  float* xanchor_o = DW(xanchor);
  float* xaxis_o = DW(xaxis);
  const int* tree = MI(body_tree);
  const int* tree_adr = MI(body_tree_adr);
  for (int lv = 1; lv < K.ms.nlevel; lv++) {
    const int a1 = tree_adr[lv + 1];
    for (int i = tree_adr[lv] + lane; i < a1; i += MJB_LW) {
      const int b = tree[i];
      const int pid = MI(body_parentid)[b];
      int jadr = MI(body_jntadr)[b];
      const int jnum = MI(body_jntnum)[b];
      if (jnum == 1 && MI(jnt_type)[jadr] == JNT_FREE) {  // smooth.py:83-95
        const int qa = MI(jnt_qposadr)[jadr];
        V3 p = ld3(qpos + qa);
        Q4 q = normalize(ld4(qpos + qa + 3));
        st3(xpos_o + 3 * b, p);
        st4(xquat_o + 4 * b, q);
        st3(xanchor_o + 3 * jadr, p);
        st3(xaxis_o + 3 * jadr, ld3(MR(jnt_axis) + 3 * jadr));
        continue;
      }
      V3 xpos;
      Q4 xquat;
      const int mocapid = MI(body_mocapid)[b];
      if (mocapid >= 0) {
        xpos = ld3(DW(mocap_pos) + 3 * mocapid);
        xquat = ld4(DW(mocap_quat) + 4 * mocapid);
      } else {
        xpos = ld3(MR(body_pos) + 3 * b);
        xquat = ld4(MR(body_quat) + 4 * b);
      }
      {
        Q4 pq = ld4(xquat_o + 4 * pid);
        xpos = rot_vec_quat(xpos, pq) + ld3(xpos_o + 3 * pid);
        xquat = mul_quat(pq, xquat);
      }
      for (int k = 0; k < jnum; k++, jadr++) {  // smooth.py:115-139
        const int qa = MI(jnt_qposadr)[jadr];
        const int jt = MI(jnt_type)[jadr];
        const V3 jaxis = ld3(MR(jnt_axis) + 3 * jadr);
        const V3 jpos = ld3(MR(jnt_pos) + 3 * jadr);
        const V3 xanchor = rot_vec_quat(jpos, xquat) + xpos;
        const V3 xaxis = rot_vec_quat(jaxis, xquat);
        if (jt == JNT_BALL) {
          Q4 ql = normalize(ld4(qpos + qa));
          xquat = mul_quat(xquat, ql);
          xpos = xanchor - rot_vec_quat(jpos, xquat);
        } else if (jt == JNT_SLIDE) {
          xpos = xpos + xaxis * (qpos[qa] - MR(qpos0)[qa]);
        } else if (jt == JNT_HINGE) {
          Q4 ql = axis_angle_to_quat(jaxis, qpos[qa] - MR(qpos0)[qa]);
          xquat = mul_quat(xquat, ql);
          xpos = xanchor - rot_vec_quat(jpos, xquat);
        }
        st3(xanchor_o + 3 * jadr, xanchor);
        st3(xaxis_o + 3 * jadr, xaxis);
      }
      st3(xpos_o + 3 * b, xpos);
      st4(xquat_o + 4 * b, normalize(xquat));
    }
    __syncwarp();
  }
  // body matrices and inertial frames (smooth.py:147-173)
  for (int b = lane; b < K.ms.nbody; b += MJB_LW) {
    const Q4 q = ld4(xquat_o + 4 * b);
    stm3(DW(xmat) + 9 * b, quat_to_mat(q));
    st3(DW(xipos) + 3 * b, ld3(xpos_o + 3 * b) + rot_vec_quat(ld3(MR(body_ipos) + 3 * b), q));
    stm3(DW(ximat) + 9 * b, quat_to_mat(mul_quat(q, ld4(MR(body_iquat) + 4 * b))));
  }
  // geoms (smooth.py:177-203); static world geoms keep their make_data pose
  for (int g = lane; g < K.ms.ngeom; g += MJB_LW) {
    if (MI(geom_static)[g]) continue;
    const int b = MI(geom_bodyid)[g];
    const Q4 q = ld4(xquat_o + 4 * b);
    st3(DW(geom_xpos) + 3 * g, ld3(xpos_o + 3 * b) + rot_vec_quat(ld3(MR(geom_pos) + 3 * g), q));
    stm3(DW(geom_xmat) + 9 * g, quat_to_mat(mul_quat(q, ld4(MR(geom_quat) + 4 * g))));
  }
  for (int s = lane; s < K.ms.nsite; s += MJB_LW) {  // smooth.py:207-224
    const int b = MI(site_bodyid)[s];
    const Q4 q = ld4(xquat_o + 4 * b);
    st3(DW(site_xpos) + 3 * s, ld3(xpos_o + 3 * b) + rot_vec_quat(ld3(MR(site_pos) + 3 * s), q));
    stm3(DW(site_xmat) + 9 * s, quat_to_mat(mul_quat(q, ld4(MR(site_quat) + 4 * s))));
  }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// smooth.com_pos (smooth.py:464-632)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void s_com_pos(const KParams& K, const unsigned char* mb, int w, int lane) {
  const float* xipos = DW(xipos);
  float* scom = DW(subtree_com);
  const int nb = K.ms.nbody;
  for (int b = lane; b < nb; b += MJB_LW) {  // subtree = id range [b, end)
    const int end = MI(body_subtreeend)[b];
    V3 acc = V3{0.f, 0.f, 0.f};
    for (int c = b; c < end; c++) acc = acc + ld3(xipos + 3 * c) * MR(body_mass)[c];
    const float mass = MR(body_subtreemass)[b];
    if (mass != 0.f) acc = acc * (1.0f / mass);
    st3(scom + 3 * b, acc);
  }
  __syncwarp();
  for (int b = lane; b < nb; b += MJB_LW) {  // smooth.py:511-553 (mju_inertCom)
    const M3 mat = ldm3(DW(ximat) + 9 * b);
    const V3 inert = ld3(MR(body_inertia) + 3 * b);
    const float mass = MR(body_mass)[b];
    const V3 dif = ld3(xipos + 3 * b) - ld3(scom + 3 * MI(body_rootid)[b]);
    const float* a = mat.m;
    float* res = DW(cinert) + 10 * b;
    // tmp = mat * diag(inert) * mat^T
    const float t00 = a[0] * a[0] * inert.x + a[1] * a[1] * inert.y + a[2] * a[2] * inert.z;
    const float t11 = a[3] * a[3] * inert.x + a[4] * a[4] * inert.y + a[5] * a[5] * inert.z;
    const float t22 = a[6] * a[6] * inert.x + a[7] * a[7] * inert.y + a[8] * a[8] * inert.z;
    const float t01 = a[0] * a[3] * inert.x + a[1] * a[4] * inert.y + a[2] * a[5] * inert.z;
    const float t02 = a[0] * a[6] * inert.x + a[1] * a[7] * inert.y + a[2] * a[8] * inert.z;
    const float t12 = a[3] * a[6] * inert.x + a[4] * a[7] * inert.y + a[5] * a[8] * inert.z;
    res[0] = t00 + mass * (dif.y * dif.y + dif.z * dif.z);
    res[1] = t11 + mass * (dif.x * dif.x + dif.z * dif.z);
    res[2] = t22 + mass * (dif.x * dif.x + dif.y * dif.y);
    res[3] = t01 - mass * dif.x * dif.y;
    res[4] = t02 - mass * dif.x * dif.z;
    res[5] = t12 - mass * dif.y * dif.z;
    res[6] = mass * dif.x;
    res[7] = mass * dif.y;
    res[8] = mass * dif.z;
    res[9] = mass;
  }
  for (int j = lane; j < K.ms.njnt; j += MJB_LW) {  // smooth.py:557-598
    const int b = MI(jnt_bodyid)[j];
    const int jt = MI(jnt_type)[j];
    float* res = DW(cdof) + 6 * MI(jnt_dofadr)[j];
    const V3 off = ld3(scom + 3 * MI(body_rootid)[b]) - ld3(DW(xanchor) + 3 * j);
    if (jt == JNT_FREE || jt == JNT_BALL) {
      const M3 xmat = ldm3(DW(xmat) + 9 * b);
      if (jt == JNT_FREE) {
        for (int i = 0; i < 18; i++) res[i] = 0.f;
        res[3] = 1.f;
        res[6 + 4] = 1.f;
        res[12 + 5] = 1.f;
        res += 18;
      }
#pragma unroll
      for (int a = 0; a < 3; a++) {
        const V3 ax = mat_col(xmat, a);
        st6(res + 6 * a, mk6(ax, cross(ax, off)));
      }
    } else if (jt == JNT_SLIDE) {
      st6(res, mk6(V3{0.f, 0.f, 0.f}, ld3(DW(xaxis) + 3 * j)));
    } else {
      const V3 ax = ld3(DW(xaxis) + 3 * j);
      st6(res, mk6(ax, cross(ax, off)));
    }
  }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// smooth.crb (smooth.py:807-912), dense qM (nv_pad x nv_pad, symmetric, zero padded)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void s_crb(const KParams& K, const unsigned char* mb, int w, int lane) {
  const int nb = K.ms.nbody, nv = K.ms.nv, np = K.ms.nv_pad;
  const float* cinert = DW(cinert);
  float* crb = DW(crb);
  for (int b = lane; b < nb; b += MJB_LW) {
    // reference accumulates children into parents except into the world body (smooth.py:820)
    const int end = b == 0 ? 1 : MI(body_subtreeend)[b];
    float acc[10];
#pragma unroll
    for (int i = 0; i < 10; i++) acc[i] = 0.f;
    for (int c = b; c < end; c++) {
#pragma unroll
      for (int i = 0; i < 10; i++) acc[i] += cinert[10 * c + i];
    }
#pragma unroll
    for (int i = 0; i < 10; i++) crb[10 * b + i] = acc[i];
  }
  float* qM = DW(qM);
  {
    float4* q4 = reinterpret_cast<float4*>(qM);
    const int n4 = (np * np) >> 2;
    for (int i = lane; i < n4; i += MJB_LW) q4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncwarp();
  const float* cdof = DW(cdof);
  for (int i = lane; i < nv; i += MJB_LW) {  // smooth.py:856-885
    const S6 ci = ld6(cdof + 6 * i);
    const S6 buf = inert_vec(crb + 10 * MI(dof_bodyid)[i], ci);
    qM[i * np + i] = MR(dof_armature)[i] + dot6(ci, buf);
    for (int j = MI(dof_parentid)[i]; j >= 0; j = MI(dof_parentid)[j]) {
      const float v = dot6(ld6(cdof + 6 * j), buf);
      qM[i * np + j] = v;
      qM[j * np + i] = v;
    }
  }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// Per-warp shared-memory scratch layout, shared by factor_m / fwd_acceleration / solve / euler:
//   H (nv x ld) | L (nv x ld) | 8 dof vectors (nv_pad each) | 5 row vectors (njmax_pad each)
// ld is a multiple of 4 with ld/4 odd so that LDS.128 row accesses by consecutive lanes are
// bank-conflict free.
// ---------------------------------------------------------------------------------------------
__host__ __device__ inline int chol_ld(int nv_pad) { return ((nv_pad >> 2) & 1) ? nv_pad : nv_pad + 4; }
// compact scratch of the stages that only factor / solve with one matrix: L (nv x ld) | x (nv_pad) | rd (nv_pad)
struct CholScr {
  float *L, *x, *rd;  // factor, right-hand side / solution, reciprocal diagonal of L
  int ld;
};
__host__ __device__ inline int chol_scratch_floats(int nv, int nv_pad) { return nv * chol_ld(nv_pad) + 2 * nv_pad; }
__device__ __forceinline__ CholScr chol_scr(float* scr, int nv, int nvp) {
  CholScr c;
  c.ld = chol_ld(nvp);
  c.L = scr;
  c.x = scr + nv * c.ld;
  c.rd = c.x + nvp;
  return c;
}
// MJB_SOL_JSMEM = 1 (nv <= 32): the world's constraint Jacobian rows are staged in shared memory for the whole solve and the Hessian
// moves to a per-world global buffer (KParams.hbuf) to make room -- H is touched once per Hessian update and once per factorisation,
// J four times per iteration; rows that do not fit (solver_jcap) leave the world on the global-memory path, and the 5 CTAs x 4 warps
// per SM are kept.  Measured on B200 (humanoid, 8192 worlds, parity-green): solve 200.3 us against 194.6 us with J read through
// L1 / L2 and H in shared memory -- the staging copy and the global-memory Hessian cost more than the shared-memory rows save.  Off.
#ifndef MJB_SOL_JSMEM
#define MJB_SOL_JSMEM 0
#endif
struct SolverScr {
  float *H, *L, *x, *grad, *search, *mv, *qacc, *Ma, *qfs, *rd, *Jaref, *jv, *D, *coef, *force;
  float* Jst;       // staging area for the rows of efc_J (jcap rows of nv_pad floats), or null
  const float* Jw;  // where this world's rows are read from: Jst or the global array (set by solve_begin's caller)
  int ld, jcap;
};
// shared-memory floats of one world without the staged rows
__host__ __device__ inline int solver_base_floats(int nv, int nv_pad, int njmax_pad) {
  return ((nv > 32 || MJB_SOL_JSMEM) ? 1 : 2) * nv * chol_ld(nv_pad) + 8 * nv_pad + 5 * njmax_pad;
}
// rows of J staged per world: what is left of the shared memory that lets 5 CTAs of 4 warps share an SM (227 kB - 1 kB per CTA)
__host__ __device__ inline int solver_jcap(int nv, int nv_pad, int njmax_pad) {
  if (!MJB_SOL_JSMEM || nv > 32) return 0;
  const int budget = (227 * 1024 - 5 * 1024) / (5 * 4 * 4);  // floats per world
  int cap = (budget - solver_base_floats(nv, nv_pad, njmax_pad)) / nv_pad;
  cap = cap < 0 ? 0 : cap & ~3;
  return cap < njmax_pad ? cap : njmax_pad;
}
__host__ __device__ inline int solver_scratch_floats(int nv, int nv_pad, int njmax_pad) {
  return solver_base_floats(nv, nv_pad, njmax_pad) + solver_jcap(nv, nv_pad, njmax_pad) * nv_pad;
}
__device__ __forceinline__ SolverScr solver_scr(float* scr, int nv, int nvp, int njp, float* hglobal = nullptr) {
  SolverScr s;
  s.ld = chol_ld(nvp);
  const bool hg = nv > 32 || MJB_SOL_JSMEM;  // H in global memory
  s.H = hg ? hglobal : scr;
  s.L = hg ? scr : scr + nv * s.ld;
  float* v = s.L + nv * s.ld;
  s.x = v; s.grad = v + nvp; s.search = v + 2 * nvp; s.mv = v + 3 * nvp; s.qacc = v + 4 * nvp; s.Ma = v + 5 * nvp; s.qfs = v + 6 * nvp; s.rd = v + 7 * nvp;
  float* r = v + 8 * nvp;
  s.Jaref = r; s.jv = r + njp; s.D = r + 2 * njp; s.coef = r + 3 * njp; s.force = r + 4 * njp;
  s.jcap = solver_jcap(nv, nvp, njp);
  s.Jst = s.jcap > 0 ? r + 5 * njp : nullptr;
  s.Jw = nullptr;
  return s;
}

// ---------------------------------------------------------------------------------------------
// dense Cholesky in shared memory: A (n x n, leading dimension ld, lower triangle used) is
// overwritten by L with A = L L^T (wp.tile_cholesky semantics, smooth.py:1085).  Left-looking,
// lane = row, 128-bit shared-memory loads for the row dot products.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float chol_row_dot(const float* ri, const float* rj, int j, float s) {
  const float4* a4 = reinterpret_cast<const float4*>(ri);
  const float4* b4 = reinterpret_cast<const float4*>(rj);
  const int j4 = j >> 2;
  for (int q = 0; q < j4; q++) {
    const float4 a = a4[q], b = b4[q];
    s -= a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
  }
  for (int k = j4 << 2; k < j; k++) s -= ri[k] * rj[k];
  return s;
}
static __device__ __noinline__ void warp_cholesky_smem(float* A, int ld, int n, int lane) {
  const int i0 = lane, i1 = lane + 32;
  for (int j = 0; j < n; j++) {
    const float* rj = A + j * ld;
    float s0 = 0.f, s1 = 0.f;
    if (i0 >= j && i0 < n) s0 = chol_row_dot(A + i0 * ld, rj, j, A[i0 * ld + j]);
    if (i1 >= j && i1 < n) s1 = chol_row_dot(A + i1 * ld, rj, j, A[i1 * ld + j]);
    const float sj = __shfl_sync(FULL, (j < 32) ? s0 : s1, j & 31);
    const float inv = rsqrtf(fmaxf(sj, MJ_MINVAL));
    if (i0 >= j && i0 < n) A[i0 * ld + j] = (i0 == j) ? sj * inv : s0 * inv;
    if (i1 >= j && i1 < n) A[i1 * ld + j] = (i1 == j) ? sj * inv : s1 * inv;
    __syncwarp();
  }
}

// x <- (L L^T)^-1 x with L in shared memory, x in shared memory (n <= 64).
static __device__ __noinline__ void warp_chol_solve_smem(const float* L, int ld, int n, float* x, int lane) {
  const int i0 = lane, i1 = lane + 32;
  float x0 = i0 < n ? x[i0] : 0.f, x1 = i1 < n ? x[i1] : 0.f;
  const float d0 = i0 < n ? 1.0f / L[i0 * ld + i0] : 1.f, d1 = i1 < n ? 1.0f / L[i1 * ld + i1] : 1.f;
  if (n <= 32) {
    for (int j = 0; j < n; j++) {  // forward: L y = b
      const float xj = __shfl_sync(FULL, x0 * d0, j);
      if (i0 == j) x0 = xj;
      if (i0 > j && i0 < n) x0 -= L[i0 * ld + j] * xj;
    }
    for (int j = n - 1; j >= 0; j--) {  // backward: L^T x = y
      const float xj = __shfl_sync(FULL, x0 * d0, j);
      if (i0 == j) x0 = xj;
      if (i0 < j) x0 -= L[j * ld + i0] * xj;
    }
  } else {
    for (int j = 0; j < n; j++) {
      const float xj = __shfl_sync(FULL, (j < 32) ? x0 * d0 : x1 * d1, j & 31);
      if (i0 == j) x0 = xj;
      if (i1 == j) x1 = xj;
      if (i0 > j && i0 < n) x0 -= L[i0 * ld + j] * xj;
      if (i1 > j && i1 < n) x1 -= L[i1 * ld + j] * xj;
    }
    for (int j = n - 1; j >= 0; j--) {
      const float xj = __shfl_sync(FULL, (j < 32) ? x0 * d0 : x1 * d1, j & 31);
      if (i0 == j) x0 = xj;
      if (i1 == j) x1 = xj;
      if (i0 < j) x0 -= L[j * ld + i0] * xj;
      if (i1 < j) x1 -= L[j * ld + i1] * xj;
    }
  }
  if (i0 < n) x[i0] = x0;
  if (i1 < n) x[i1] = x1;
  __syncwarp();
}

// Register-resident variants for n <= 32 (N = n rounded up to a multiple of 4, compile time):
// lane i owns row i of the matrix in registers; the padded rows/columns are the identity, so the
// fully unrolled loops need no bounds tests.  Right-looking: column j is scaled, broadcast lane by
// lane with shuffles and subtracted from the trailing rows (n^2/2 shuffle+FMA pairs, no shared-memory
// traffic besides one 128-bit row load and store per lane).
template <int N>
static __device__ __noinline__ void warp_cholesky_reg(float* A, int ld, int n, int lane) {
  float r[N];
  const bool in = lane < n;
  const float4* row4 = reinterpret_cast<const float4*>(A + (in ? lane : 0) * ld);
#pragma unroll
  for (int q = 0; q < N / 4; q++) {
    const float4 v = row4[q];
    r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
  }
#pragma unroll
  for (int k = 0; k < N; k++) {
    if (!in) r[k] = (k == lane) ? 1.0f : 0.0f;
    if (k >= N - 4 && k >= n) r[k] = (k == lane) ? 1.0f : 0.0f;
  }
#pragma unroll
  for (int j = 0; j < N; j++) {
    const float ajj = __shfl_sync(FULL, r[j], j);
    const float lij = r[j] * rsqrtf(fmaxf(ajj, MJ_MINVAL));
    r[j] = lij;
#pragma unroll
    for (int k = j + 1; k < N; k++) r[k] -= lij * __shfl_sync(FULL, lij, k);
  }
  if (in) {
    float4* out4 = reinterpret_cast<float4*>(A + lane * ld);
#pragma unroll
    for (int q = 0; q < N / 4; q++) out4[q] = make_float4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
  }
  __syncwarp();
}

template <int N>
static __device__ __noinline__ void warp_chol_solve_reg(const float* L, int ld, int n, float* x, int lane) {
  float r[N], c[N];  // r = row `lane` of L left of the diagonal, c = column `lane` of L below it
  const bool in = lane < n;
  const int i = in ? lane : 0;
  const float4* row4 = reinterpret_cast<const float4*>(L + i * ld);
#pragma unroll
  for (int q = 0; q < N / 4; q++) {
    const float4 v = row4[q];
    r[4 * q] = v.x; r[4 * q + 1] = v.y; r[4 * q + 2] = v.z; r[4 * q + 3] = v.w;
  }
#pragma unroll
  for (int j = 0; j < N; j++) {
    r[j] = (j < lane && in) ? r[j] : 0.f;
    c[j] = (j > lane && j < n) ? L[j * ld + i] : 0.f;
  }
  float xi = in ? x[lane] : 0.f;
  const float dinv = in ? 1.0f / L[i * ld + i] : 0.f;
  // forward L y = b: lane i accumulates b_i - sum_{k<i} L_ik y_k; y_k = that * dinv_k is broadcast
#pragma unroll
  for (int j = 0; j < N; j++) xi = fmaf(-r[j], __shfl_sync(FULL, xi * dinv, j), xi);
  xi *= dinv;
  // backward L^T x = y
#pragma unroll
  for (int j = N - 1; j >= 0; j--) xi = fmaf(-c[j], __shfl_sync(FULL, xi * dinv, j), xi);
  if (in) x[lane] = xi * dinv;
  __syncwarp();
}

__device__ __forceinline__ float dot4(const float4 a, const float4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }

// `env` (optional): env[i] = first column of row i that can be non-zero (the matrix's envelope: block-diagonal mass matrices of
// several kinematic trees, Hessians whose constraint rows touch few trees).  A Cholesky factor keeps the envelope, so the row
// updates start there and rows whose envelope begins right of a block column skip it: three 27-dof trees cost 3 x 27^3 instead of 81^3.
// MJB_CHOL_ENV = 0 (set by the solver's translation unit, whose dense Hessians have no fixed envelope) compiles the argument away.
#ifndef MJB_CHOL_ENV
#define MJB_CHOL_ENV 1
#endif
static __device__ __noinline__ void warp_cholesky_blk(float* A, int ld, int n, float* rdiag, int lane, const int* env_arg = nullptr) {
#if MJB_CHOL_ENV
  const int* env = env_arg;
#else
  constexpr const int* env = nullptr;
#endif
  const int nb = (n + 3) >> 2;
  #pragma unroll 1
  for (int J = 0; J < nb; J++) {
    const int c0 = 4 * J;
    // rows of the diagonal block; rows past the end alias the last row and are masked below
    const int r1 = min(c0 + 1, n - 1), r2 = min(c0 + 2, n - 1), r3 = min(c0 + 3, n - 1);
    const bool v1 = c0 + 1 < n, v2 = c0 + 2 < n, v3 = c0 + 3 < n;
    const float4* p0 = reinterpret_cast<const float4*>(A + c0 * ld);
    const float4* p1 = reinterpret_cast<const float4*>(A + r1 * ld);
    const float4* p2 = reinterpret_cast<const float4*>(A + r2 * ld);
    const float4* p3 = reinterpret_cast<const float4*>(A + r3 * ld);
    #pragma unroll 1
    for (int i = lane; i < n; i += MJB_LW) {
      if (i < c0) continue;
      const int e = env ? env[i] : 0;
      if (e > c0 + 3) continue;  // the row is zero in this block column
      float4* own = reinterpret_cast<float4*>(A + i * ld);
      float4 acc = own[J];
      #pragma unroll 1
      for (int q = e >> 2; q < J; q++) {
        const float4 o = own[q];
        acc.x -= dot4(o, p0[q]);
        acc.y -= dot4(o, p1[q]);
        acc.z -= dot4(o, p2[q]);
        acc.w -= dot4(o, p3[q]);
      }
      own[J] = acc;
    }
    __syncwarp();
    // 4x4 diagonal block -> its Cholesky factor (all lanes, from broadcast loads)
    const float4 s0 = p0[J], s1 = p1[J], s2 = p2[J], s3 = p3[J];
    const float d0 = fmaxf(s0.x, MJ_MINVAL), i0 = rsqrtf(d0);
    const float l10 = v1 ? s1.x * i0 : 0.f;
    const float d1 = v1 ? fmaxf(s1.y - l10 * l10, MJ_MINVAL) : 1.f, i1 = rsqrtf(d1);
    const float l20 = v2 ? s2.x * i0 : 0.f, l21 = v2 ? (s2.y - l20 * l10) * i1 : 0.f;
    const float d2 = v2 ? fmaxf(s2.z - l20 * l20 - l21 * l21, MJ_MINVAL) : 1.f, i2 = rsqrtf(d2);
    const float l30 = v3 ? s3.x * i0 : 0.f, l31 = v3 ? (s3.y - l30 * l10) * i1 : 0.f, l32 = v3 ? (s3.z - l30 * l20 - l31 * l21) * i2 : 0.f;
    const float d3 = v3 ? fmaxf(s3.w - l30 * l30 - l31 * l31 - l32 * l32, MJ_MINVAL) : 1.f, i3 = rsqrtf(d3);
    if (lane == 0) reinterpret_cast<float4*>(rdiag)[J] = make_float4(i0, i1, i2, i3);  // 1 / L_jj
    __syncwarp();
    #pragma unroll 1
    for (int i = lane; i < n; i += MJB_LW) {
      if (i < c0) continue;
      if (env && env[i] > c0 + 3) continue;
      float4* own = reinterpret_cast<float4*>(A + i * ld);
      const float4 a = own[J];
      const int m = i - c0;  // position inside the diagonal block (>= 4: below it)
      float4 l;
      l.x = (m == 0) ? d0 * i0 : a.x * i0;
      l.y = (m == 1) ? d1 * i1 : (a.y - l.x * l10) * i1;
      l.z = (m == 2) ? d2 * i2 : (a.z - l.x * l20 - l.y * l21) * i2;
      l.w = (m == 3) ? d3 * i3 : (a.w - l.x * l30 - l.y * l31 - l.z * l32) * i3;
      // zero the strictly upper part of the diagonal block and the padding columns
      if (m < 1 || !v1) l.y = 0.f;
      if (m < 2 || !v2) l.z = 0.f;
      if (m < 3 || !v3) l.w = 0.f;
      own[J] = l;
    }
    __syncwarp();
  }
}

// the 4x4 diagonal block J of L (rows/columns past n behave as identity)
struct DiagBlock {
  float r0, r1, r2, r3;  // reciprocal diagonal
  float l10, l20, l21, l30, l31, l32;
};
__device__ __forceinline__ DiagBlock load_diag_block(const float* L, int ld, int n, const float* rdiag, int J) {
  const int c0 = 4 * J;
  const bool v1 = c0 + 1 < n, v2 = c0 + 2 < n, v3 = c0 + 3 < n;
  const float4 rd = reinterpret_cast<const float4*>(rdiag)[J];
  const float4 s1 = reinterpret_cast<const float4*>(L + min(c0 + 1, n - 1) * ld)[J];
  const float4 s2 = reinterpret_cast<const float4*>(L + min(c0 + 2, n - 1) * ld)[J];
  const float4 s3 = reinterpret_cast<const float4*>(L + min(c0 + 3, n - 1) * ld)[J];
  DiagBlock b;
  b.r0 = rd.x; b.r1 = rd.y; b.r2 = rd.z; b.r3 = rd.w;
  b.l10 = v1 ? s1.x : 0.f;
  b.l20 = v2 ? s2.x : 0.f;
  b.l21 = v2 ? s2.y : 0.f;
  b.l30 = v3 ? s3.x : 0.f;
  b.l31 = v3 ? s3.y : 0.f;
  b.l32 = v3 ? s3.z : 0.f;
  return b;
}

// reciprocal diagonal of a factor that was loaded rather than computed here (padding entries are 1)
__device__ __forceinline__ void chol_rdiag(const float* L, int ld, int n, float* rdiag, int lane) {
  for (int i = lane; i < ((n + 3) & ~3); i += MJB_LW) rdiag[i] = i < n ? 1.0f / L[i * ld + i] : 1.0f;
  __syncwarp();
}

// x <- (L L^T)^-1 x with L and x in shared memory (x padded to a multiple of 4 entries).
// Blocked substitution: the 4x4 diagonal systems are solved redundantly by all lanes, the
// off-diagonal part is one strip (forward: 128-bit row load, backward: four column loads) per lane.
static __device__ __noinline__ void warp_chol_solve_blk(const float* L, int ld, int n, const float* rdiag, float* x, int lane, const int* env_arg = nullptr) {
#if MJB_CHOL_ENV
  const int* env = env_arg;
#else
  constexpr const int* env = nullptr;
#endif
  const int nb = (n + 3) >> 2;
  float4* x4 = reinterpret_cast<float4*>(x);
  #pragma unroll 1
  for (int i = n + lane; i < 4 * nb; i += MJB_LW) x[i] = 0.f;
  __syncwarp();
  #pragma unroll 1
  for (int J = 0; J < nb; J++) {  // forward: L y = b
    const DiagBlock b = load_diag_block(L, ld, n, rdiag, J);
    const float4 t = x4[J];
    float4 y;
    y.x = t.x * b.r0;
    y.y = (t.y - b.l10 * y.x) * b.r1;
    y.z = (t.z - b.l20 * y.x - b.l21 * y.y) * b.r2;
    y.w = (t.w - b.l30 * y.x - b.l31 * y.y - b.l32 * y.z) * b.r3;
    __syncwarp();
    if (lane == 0) x4[J] = y;
    #pragma unroll 1
    for (int i = 4 * J + 4 + lane; i < n; i += MJB_LW) {
      if (env && env[i] > 4 * J + 3) continue;
      x[i] -= dot4(reinterpret_cast<const float4*>(L + i * ld)[J], y);
    }
    __syncwarp();
  }
  #pragma unroll 1
  for (int J = nb - 1; J >= 0; J--) {  // backward: L^T x = y
    const DiagBlock b = load_diag_block(L, ld, n, rdiag, J);
    const float4 t = x4[J];
    float4 z;
    z.w = t.w * b.r3;
    z.z = (t.z - b.l32 * z.w) * b.r2;
    z.y = (t.y - b.l21 * z.z - b.l31 * z.w) * b.r1;
    z.x = (t.x - b.l10 * z.y - b.l20 * z.z - b.l30 * z.w) * b.r0;
    __syncwarp();
    if (lane == 0) x4[J] = z;
    const int c0 = 4 * J;
    const float* q0 = L + c0 * ld;
    const float* q1 = L + min(c0 + 1, n - 1) * ld;
    const float* q2 = L + min(c0 + 2, n - 1) * ld;
    const float* q3 = L + min(c0 + 3, n - 1) * ld;
    // rows past n have z == 0 there (identity padding), so the aliased loads contribute nothing
    const int i0 = env ? min(min(env[c0], env[min(c0 + 1, n - 1)]), min(env[min(c0 + 2, n - 1)], env[min(c0 + 3, n - 1)])) : 0;
    #pragma unroll 1
    for (int i = i0 + lane; i < c0; i += MJB_LW) x[i] -= q0[i] * z.x + q1[i] * z.y + q2[i] * z.z + q3[i] * z.w;
    __syncwarp();
  }
}

// REG = true: register-resident variant (fewest instructions, large straight-line code, 60+ registers):
// used by the solver kernel.  REG = false: blocked shared-memory variant with rolled loops (compact code,
// few registers): used by the smooth-dynamics kernels, where occupancy matters more.
template <bool REG>
__device__ __forceinline__ void warp_cholesky(float* A, int ld, int n, float* rdiag, int lane, const int* env = nullptr) {
  if (!REG) {
    warp_cholesky_blk(A, ld, n, rdiag, lane, env);
    return;
  }
  switch ((n + 3) >> 2) {
    case 1: warp_cholesky_reg<4>(A, ld, n, lane); break;
    case 2: warp_cholesky_reg<8>(A, ld, n, lane); break;
    case 3: warp_cholesky_reg<12>(A, ld, n, lane); break;
    case 4: warp_cholesky_reg<16>(A, ld, n, lane); break;
    case 5: warp_cholesky_reg<20>(A, ld, n, lane); break;
    case 6: warp_cholesky_reg<24>(A, ld, n, lane); break;
    case 7: warp_cholesky_reg<28>(A, ld, n, lane); break;
    case 8: warp_cholesky_reg<32>(A, ld, n, lane); break;
    default: warp_cholesky_smem(A, ld, n, lane); break;
  }
}
template <bool REG>
__device__ __forceinline__ void warp_chol_solve(const float* L, int ld, int n, const float* rdiag, float* x, int lane, const int* env = nullptr) {
  if (!REG) {
    warp_chol_solve_blk(L, ld, n, rdiag, x, lane, env);
    return;
  }
  switch ((n + 3) >> 2) {
    case 1: warp_chol_solve_reg<4>(L, ld, n, x, lane); break;
    case 2: warp_chol_solve_reg<8>(L, ld, n, x, lane); break;
    case 3: warp_chol_solve_reg<12>(L, ld, n, x, lane); break;
    case 4: warp_chol_solve_reg<16>(L, ld, n, x, lane); break;
    case 5: warp_chol_solve_reg<20>(L, ld, n, x, lane); break;
    case 6: warp_chol_solve_reg<24>(L, ld, n, x, lane); break;
    case 7: warp_chol_solve_reg<28>(L, ld, n, x, lane); break;
    case 8: warp_chol_solve_reg<32>(L, ld, n, x, lane); break;
    default: warp_chol_solve_smem(L, ld, n, x, lane); break;
  }
}

// copies the n leading rows of a global matrix (ld_src) into shared memory (ld); rows are copied
// whole (the Cholesky only reads the lower triangle).  Optional diagonal shift.
#if MJB_SOL_COMPACT
static __device__ __noinline__
#else
__device__ __forceinline__
#endif
void load_rows(float* A, int ld, const float* src, int ld_src, int n, int lane, const float* diag_add,
                                          float diag_scale) {
  // a row / column walk with running indices instead of a division per element (compact code)
  if ((ld_src & 3) == 0) {
    const int c4 = ld_src >> 2, ld4 = ld >> 2;
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* a4 = reinterpret_cast<float4*>(A);
    if (c4 <= 8) {  // rows of <= 32 floats: 4 rows per pass, lane = (row % 4, float4 column)
      const int q = lane & 7;
#pragma unroll 4
      for (int i = lane >> 3; i < n; i += MJB_LW / 8)
        if (q < c4) a4[i * ld4 + q] = s4[i * c4 + q];
    } else {
#pragma unroll 1
      for (int i = 0; i < n; i++)
        for (int q = lane; q < c4; q += MJB_LW) a4[i * ld4 + q] = s4[i * c4 + q];
    }
  } else {
#pragma unroll 1
    for (int i = 0; i < n; i++)
      for (int j = lane; j < ld_src; j += MJB_LW) A[i * ld + j] = src[i * ld_src + j];
  }
  __syncwarp();
  if (diag_add != nullptr) {
    for (int i = lane; i < n; i += MJB_LW) A[i * ld + i] += diag_scale * diag_add[i];
    __syncwarp();
  }
}

// writes the lower-triangular factor held in shared memory to d.qLD (nv x nv).  Only the lower triangle is written: the strictly
// upper part of d.qLD is zero from its allocation (make_data / put_data) and nothing ever writes there.
__device__ __forceinline__ void store_qLD(const KParams& K, int w, int lane, const float* L, int ld) {
  const int nv = K.ms.nv;
  float* qLD = DW(qLD);
  if (MJB_LW == 32 && nv <= 32) {
#pragma unroll 4
    for (int i = 0; i < nv; i++)
      if (lane <= i) qLD[i * nv + lane] = L[i * ld + lane];
  } else {
    for (int i = 0; i < nv; i++)
      for (int j = lane; j <= i; j += MJB_LW) qLD[i * nv + j] = L[i * ld + j];
  }
}

// envelope of the mass matrix and of everything with its sparsity (qM - dt qDeriv, qM + dt diag): rows start at their tree's first
// dof.  Only worth the extra loads when there is more than a warp's worth of dofs.
__device__ __forceinline__ const int* mass_env(const KParams& K, const unsigned char* mb) {
#if MJB_CHOL_ENV
  return K.ms.nv > 32 ? MI(dof_tree0) : nullptr;
#else
  return nullptr;
#endif
}

// smooth.factor_m dense (smooth.py:1091-1109): qLD = chol(qM).
__device__ __forceinline__ void s_factor_m(const KParams& K, const unsigned char* mb, int w, int lane, float* scr) {
  const CholScr s = chol_scr(scr, K.ms.nv, K.ms.nv_pad);
  load_rows(s.L, s.ld, DW(qM), K.ms.nv_pad, K.ms.nv, lane, nullptr, 0.f);
  warp_cholesky<false>(s.L, s.ld, K.ms.nv, s.rd, lane, mass_env(K, mb));
  store_qLD(K, w, lane, s.L, s.ld);
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// smooth.transmission (smooth.py:2042-2149): JOINT / JOINTINPARENT.  The sparse moment layout
// (rownnz/rowadr/colind) is static for joint transmissions and is written once by make_data; the
// reference re-derives it every step with a per-world atomic counter (smooth.py:2094).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void s_transmission(const KParams& K, const unsigned char* mb, int w, int lane) {
  const float* qpos = DW(qpos);
  for (int a = lane; a < K.ms.nu; a += MJB_LW) {
    const int j = MI(actuator_trnid)[2 * a];
    const int jt = MI(jnt_type)[j], qa = MI(jnt_qposadr)[j];
    const int trn = MI(actuator_trntype)[a];
    const float* gear = MR(actuator_gear) + 6 * a;
    float* mom = DW(actuator_moment) + DW(moment_rowadr)[a];
    if (jt == JNT_FREE) {
      DW(actuator_length)[a] = 0.f;
      V3 g3 = ld3(gear + 3);
      if (trn == 1) {
        Q4 q = normalize(ld4(qpos + qa + 3));
        g3 = rot_vec_quat(g3, Q4{q.w, -q.x, -q.y, -q.z});
      }
      mom[0] = gear[0]; mom[1] = gear[1]; mom[2] = gear[2];
      st3(mom + 3, g3);
    } else if (jt == JNT_BALL) {
      Q4 q = normalize(ld4(qpos + qa));
      V3 g3 = ld3(gear);
      if (trn == 1) g3 = rot_vec_quat(g3, Q4{q.w, -q.x, -q.y, -q.z});
      DW(actuator_length)[a] = dot(quat_to_vel(q), g3);
      st3(mom, g3);
    } else {
      DW(actuator_length)[a] = qpos[qa] * gear[0];
      mom[0] = gear[0];
    }
  }
  __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// forward.fwd_velocity (forward.py:593): actuator_velocity (:541), com_vel (smooth.py:1936-2038),
// passive spring/damper (passive.py:71-180, 536-565), rne (smooth.py:1113-1274)
// ---------------------------------------------------------------------------------------------
// `parts` selects the reference's sub-functions, which are public stage functions too: FV_ACTVEL actuator_velocity (forward.py:541),
// FV_PASSIVE passive.passive (passive.py:729), FV_COMVEL smooth.com_vel (smooth.py:2015: cvel, cdof_dot), FV_RNE smooth.rne
// (smooth.py:1259: cacc, cfrc_int, qfrc_bias).  com_vel and rne share one walk down the tree; run alone, each recomputes the
// velocities on the way and stores only its own outputs.
enum { FV_ACTVEL = 1, FV_PASSIVE = 2, FV_COMVEL = 4, FV_RNE = 8, FV_ALL = 15 };
__device__ __forceinline__ void s_fwd_velocity(const KParams& K, const unsigned char* mb, int w, int lane, int parts = FV_ALL) {
  const int nb = K.ms.nbody, nv = K.ms.nv;
  const float* qvel = DW(qvel);
  const float* qpos = DW(qpos);
  const float* cdof = DW(cdof);
  float* cdof_dot = DW(cdof_dot);
  float* cvel = DW(cvel);
  float* cacc = DW(cacc);
  float* cfrc = DW(cfrc_int);
  const int* tree = MI(body_tree);
  const int* tree_adr = MI(body_tree_adr);

  for (int a = lane; a < ((parts & FV_ACTVEL) ? K.ms.nu : 0); a += MJB_LW) {
    const int adr = DW(moment_rowadr)[a], nnz = DW(moment_rownnz)[a];
    float v = 0.f;
    for (int i = 0; i < nnz; i++) v += DW(actuator_moment)[adr + i] * qvel[DW(moment_colind)[adr + i]];
    DW(actuator_velocity)[a] = v;
  }
  // passive forces: lane per joint
  const int dsbl_spring = K.o.disableflags & DSBL_SPRING, dsbl_damper = K.o.disableflags & DSBL_DAMPER;
  for (int j = lane; j < ((parts & FV_PASSIVE) ? K.ms.njnt : 0); j += MJB_LW) {
    const int da = MI(jnt_dofadr)[j], qa = MI(jnt_qposadr)[j], jt = MI(jnt_type)[j];
    const float k = MR(jnt_stiffness)[j], c = MR(dof_damping)[da];
    const bool hs = k != 0.f && !dsbl_spring, hd = c != 0.f && !dsbl_damper;
    const int n = jt == JNT_FREE ? 6 : (jt == JNT_BALL ? 3 : 1);
    float sp[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (hs) {
      if (jt == JNT_FREE) {
        sp[0] = -k * (qpos[qa] - MR(qpos_spring)[qa]);
        sp[1] = -k * (qpos[qa + 1] - MR(qpos_spring)[qa + 1]);
        sp[2] = -k * (qpos[qa + 2] - MR(qpos_spring)[qa + 2]);
        V3 dif = quat_sub(normalize(ld4(qpos + qa + 3)), ld4(MR(qpos_spring) + qa + 3));
        sp[3] = -k * dif.x; sp[4] = -k * dif.y; sp[5] = -k * dif.z;
      } else if (jt == JNT_BALL) {
        V3 dif = quat_sub(normalize(ld4(qpos + qa)), ld4(MR(qpos_spring) + qa));
        sp[0] = -k * dif.x; sp[1] = -k * dif.y; sp[2] = -k * dif.z;
      } else {
        sp[0] = -k * (qpos[qa] - MR(qpos_spring)[qa]);
      }
    }
    for (int i = 0; i < n; i++) {
      const float dm = hd ? -c * qvel[da + i] : 0.f;
      DW(qfrc_spring)[da + i] = sp[i];
      DW(qfrc_damper)[da + i] = dm;
      DW(qfrc_gravcomp)[da + i] = 0.f;
      DW(qfrc_fluid)[da + i] = 0.f;
      DW(qfrc_passive)[da + i] = sp[i] + dm;
    }
  }
  // passive.py:246-272 _gravity_force + :536-565: gravity compensation through the CoM Jacobian (support.jac_dof), lane = dof
  if ((parts & FV_PASSIVE) && K.ms.ngravcomp && !(K.o.disableflags & DSBL_GRAVITY) && !(dsbl_spring && dsbl_damper)) {
    __syncwarp();
    const V3 g{K.o.gravity_x, K.o.gravity_y, K.o.gravity_z};
    for (int dof = lane; dof < K.ms.nv; dof += MJB_LW) {
      const int db = MI(dof_bodyid)[dof], dend = MI(body_subtreeend)[db];
      const S6 cd = ld6(DW(cdof) + 6 * dof);
      float q = 0.f;
      for (int b = db; b < dend; b++) {  // bodies the dof moves: its own body's subtree
        const float gc = MR(body_gravcomp)[b];
        if (gc == 0.f) continue;
        const V3 off = ld3(DW(xipos) + 3 * b) - ld3(DW(subtree_com) + 3 * MI(body_rootid)[b]);
        q += dot(bot(cd) + cross(top(cd), off), g * (-MR(body_mass)[b] * gc));
      }
      DW(qfrc_gravcomp)[dof] = q;
      if (!MI(jnt_actgravcomp)[MI(dof_jntid)[dof]]) DW(qfrc_passive)[dof] += q;
    }
  }
  if (!(parts & (FV_COMVEL | FV_RNE))) {
    __syncwarp();
    return;
  }
  // com_vel + cacc, level-synchronous (cacc needs cdof_dot of the same body only).  Run alone, com_vel keeps its running velocities
  // in cvel and rne its accelerations in cacc; the other quantity of the walk then travels through the scratch rows of cfrc_ext
  // (bodies x 6, rewritten to zero at the end) so that neither touches the other's outputs.
  const bool do_cv = (parts & FV_COMVEL) != 0, do_rne = (parts & FV_RNE) != 0;
  if (!do_cv) cvel = DW(cfrc_int);  // rne alone: velocities are staged in cfrc_int (defined by rne itself, after the walk)
  if (!do_rne) cacc = DW(cfrc_ext);  // com_vel alone: the accelerations of the walk are not an output
  if (lane < 6) {
    cvel[lane] = 0.f;
    float g = 0.f;
    if (!(K.o.disableflags & DSBL_GRAVITY)) g = lane == 3 ? -K.o.gravity_x : (lane == 4 ? -K.o.gravity_y : (lane == 5 ? -K.o.gravity_z : 0.f));
    cacc[lane] = g;
  }
  __syncwarp();
  for (int lv = 1; lv < K.ms.nlevel; lv++) {
    const int a1 = tree_adr[lv + 1];
    for (int i = tree_adr[lv] + lane; i < a1; i += MJB_LW) {
      const int b = tree[i], pid = MI(body_parentid)[b];
      S6 v = ld6(cvel + 6 * pid);
      S6 acc = ld6(cacc + 6 * pid);
      int dof = MI(body_dofadr)[b];
      const int j0 = MI(body_jntadr)[b], jn = MI(body_jntnum)[b];
      for (int j = j0; j < j0 + jn; j++) {
        const int jt = MI(jnt_type)[j];
        if (jt == JNT_FREE) {
          for (int k = 0; k < 3; k++) {
            axpy6(v, qvel[dof + k], ld6(cdof + 6 * (dof + k)));
            st6(cdof_dot + 6 * (dof + k), S6{{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}});
          }
          for (int k = 3; k < 6; k++) {
            const S6 cd = motion_cross(v, ld6(cdof + 6 * (dof + k)));
            st6(cdof_dot + 6 * (dof + k), cd);
            axpy6(acc, qvel[dof + k], cd);
          }
          for (int k = 3; k < 6; k++) axpy6(v, qvel[dof + k], ld6(cdof + 6 * (dof + k)));
          dof += 6;
        } else if (jt == JNT_BALL) {
          for (int k = 0; k < 3; k++) {
            const S6 cd = motion_cross(v, ld6(cdof + 6 * (dof + k)));
            st6(cdof_dot + 6 * (dof + k), cd);
            axpy6(acc, qvel[dof + k], cd);
          }
          for (int k = 0; k < 3; k++) axpy6(v, qvel[dof + k], ld6(cdof + 6 * (dof + k)));
          dof += 3;
        } else {
          const S6 c = ld6(cdof + 6 * dof);
          const S6 cd = motion_cross(v, c);
          st6(cdof_dot + 6 * dof, cd);
          axpy6(acc, qvel[dof], cd);
          axpy6(v, qvel[dof], c);
          dof += 1;
        }
      }
      st6(cvel + 6 * b, v);
      st6(cacc + 6 * b, acc);
      if (do_rne) {  // body force (smooth.py:1188-1209), own contribution; subtree sums follow
        const float* ci = DW(cinert) + 10 * b;
        S6 f = inert_vec(ci, acc);
        const S6 u = motion_cross_force(v, inert_vec(ci, v));
#pragma unroll
        for (int k = 0; k < 6; k++) f.v[k] += u.v[k];
        st6(DW(cfrc_ext) + 6 * b, f);  // staging: cfrc_ext is rewritten to zero below
      }
    }
    __syncwarp();
  }
  if (do_rne) {  // cfrc_int[b] = sum of the own forces over the subtree (smooth.py:1219-1240 without atomics)
    for (int b = lane; b < nb; b += MJB_LW) {
      const int end = MI(body_subtreeend)[b];
      S6 acc = S6{{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}};
      for (int c = (b == 0 ? 1 : b); c < end; c++) axpy6(acc, 1.0f, ld6(DW(cfrc_ext) + 6 * c));
      st6(cfrc + 6 * b, acc);
    }
    __syncwarp();
  }
  for (int b = lane; b < nb; b += MJB_LW) st6(DW(cfrc_ext) + 6 * b, S6{{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}});
  if (do_rne)
    for (int i = lane; i < nv; i += MJB_LW) DW(qfrc_bias)[i] = dot6(ld6(cdof + 6 * i), ld6(cfrc + 6 * MI(dof_bodyid)[i]));
  __syncwarp();
}

// forward.fwd_actuation (forward.py:617-927), no muscles / tendons
__device__ __forceinline__ void s_fwd_actuation(const KParams& K, const unsigned char* mb, int w, int lane) {
  const int nv = K.ms.nv, nu = K.ms.nu;
  if (!nu || (K.o.disableflags & DSBL_ACTUATION)) {
    for (int i = lane; i < K.ms.na; i += MJB_LW) DW(act_dot)[i] = 0.f;
    for (int i = lane; i < nv; i += MJB_LW) DW(qfrc_actuator)[i] = 0.f;
    __syncwarp();
    return;
  }
  for (int u = lane; u < nu; u += MJB_LW) {
    float ctrl = DW(ctrl)[u];
    if (MI(actuator_ctrllimited)[u] && !(K.o.disableflags & DSBL_CLAMPCTRL))
      ctrl = clampf(ctrl, MR(actuator_ctrlrange)[2 * u], MR(actuator_ctrlrange)[2 * u + 1]);
    float ctrl_act = ctrl;
    const int af = MI(actuator_actadr)[u];
    if (K.ms.na && af >= 0) {
      const int al = af + MI(actuator_actnum)[u] - 1, dyn = MI(actuator_dyntype)[u];
      const float* dynprm = MR(actuator_dynprm) + 10 * u;
      float act_dot = 0.f, act = 0.f;
      if (dyn == 1) act_dot = ctrl;
      else if (dyn == 2 || dyn == 3) { act = DW(act)[al]; act_dot = (ctrl - act) / fmaxf(dynprm[0], MJ_MINVAL); }
      DW(act_dot)[al] = act_dot;
      if (MI(actuator_actearly)[u]) {
        if (dyn == 1 || dyn == 0) act = DW(act)[al];
        ctrl_act = next_act(K.o.timestep, dyn, dynprm, MR(actuator_actrange) + 2 * u, act, act_dot, 1.0f, MI(actuator_actlimited)[u] != 0);
      } else {
        ctrl_act = DW(act)[al];
      }
    }
    const float len = DW(actuator_length)[u], vel = DW(actuator_velocity)[u];
    const float* gp = MR(actuator_gainprm) + 10 * u;
    const float* bp = MR(actuator_biasprm) + 10 * u;
    float gain = 0.f, bias = 0.f;
    const int gt = MI(actuator_gaintype)[u];
    if (gt == 0) gain = gp[0];
    else if (gt == 1) gain = gp[0] + gp[1] * len + gp[2] * vel;
    if (MI(actuator_biastype)[u] == 1) bias = bp[0] + bp[1] * len + bp[2] * vel;
    float force = gain * ctrl_act + bias;
    if (MI(actuator_forcelimited)[u]) force = clampf(force, MR(actuator_forcerange)[2 * u], MR(actuator_forcerange)[2 * u + 1]);
    DW(actuator_force)[u] = force;
  }
  __syncwarp();
  // qfrc_actuator = moment^T force (forward.py:783), deterministic and without atomics: lane u holds the
  // k-th entry (column, moment * force) of actuator u's moment row; the entries are broadcast one by one
  // and the lane that owns the dof adds them up
  for (int i = lane; i < ((nv + MJB_LW - 1) & ~(MJB_LW - 1)); i += MJB_LW) {
    float q = 0.f;
    for (int u0 = 0; u0 < nu; u0 += MJB_LW) {
      const int u = u0 + lane, cnt = min(MJB_LW, nu - u0);
      int adr = 0, nnz = 0;
      float force = 0.f;
      if (u < nu) {
        adr = DW(moment_rowadr)[u];
        nnz = DW(moment_rownnz)[u];
        force = DW(actuator_force)[u];
      }
      int maxnnz = nnz;
#pragma unroll
      for (int o = MJB_LW / 2; o > 0; o >>= 1) maxnnz = max(maxnnz, __shfl_xor_sync(FULL, maxnnz, o));
      for (int k = 0; k < maxnnz; k++) {
        const int col = k < nnz ? DW(moment_colind)[adr + k] : -1;
        const float val = k < nnz ? DW(actuator_moment)[adr + k] * force : 0.f;
        for (int uu = 0; uu < cnt; uu++) {
          const int c = __shfl_sync(FULL, col, uu, MJB_LW);
          const float v = __shfl_sync(FULL, val, uu, MJB_LW);
          if (c == i) q += v;
        }
      }
    }
    if (i >= nv) continue;
    const int j = MI(dof_jntid)[i];
    if (K.ms.ngravcomp && MI(jnt_actgravcomp)[j]) q += DW(qfrc_gravcomp)[i];
    if (MI(jnt_actfrclimited)[j]) q = clampf(q, MR(jnt_actfrcrange)[2 * j], MR(jnt_actfrcrange)[2 * j + 1]);
    DW(qfrc_actuator)[i] = q;
  }
  __syncwarp();
}

// forward.fwd_acceleration (forward.py:931-969) + support.apply_ft (support.py:175-216).
__device__ __forceinline__ void s_fwd_acceleration(const KParams& K, const unsigned char* mb, int w, int lane, float* scr, bool factorize) {
  const int nv = K.ms.nv;
  const CholScr s = chol_scr(scr, nv, K.ms.nv_pad);
  float* x = s.x;
  const float* xfrc = DW(xfrc_applied);
  // any applied wrench in this world?  (usually none: one coalesced pass instead of a subtree scan per dof)
  bool anyx = false;
  for (int k = lane; k < 6 * K.ms.nbody; k += MJB_LW) anyx |= xfrc[k] != 0.f;
  {
    const unsigned bal = __ballot_sync(FULL, anyx);
    const int g0 = (threadIdx.x & 31) & ~(MJB_LW - 1);
    anyx = ((bal >> g0) & (MJB_LW == 32 ? 0xffffffffu : ((1u << (MJB_LW & 31)) - 1u))) != 0u;
  }
  for (int i = lane; i < nv; i += MJB_LW) {
    float q = DW(qfrc_passive)[i] - DW(qfrc_bias)[i] + DW(qfrc_actuator)[i] + DW(qfrc_applied)[i];
    const S6 cd = ld6(DW(cdof) + 6 * i);
    const int db = MI(dof_bodyid)[i], end = anyx ? MI(body_subtreeend)[db] : db;
    for (int b = db; b < end; b++) {  // bodies in the subtree of the dof's body
      const S6 ft = ld6(xfrc + 6 * b);
      if (ft.v[0] == 0.f && ft.v[1] == 0.f && ft.v[2] == 0.f && ft.v[3] == 0.f && ft.v[4] == 0.f && ft.v[5] == 0.f) continue;
      const V3 off = ld3(DW(xipos) + 3 * b) - ld3(DW(subtree_com) + 3 * MI(body_rootid)[b]);
      q += dot(bot(cd), top(ft)) + dot(top(cd), bot(ft)) + dot(cross(top(cd), off), top(ft));
    }
    DW(qfrc_smooth)[i] = q;
    x[i] = q;
  }
  __syncwarp();
  if (factorize) {
    load_rows(s.L, s.ld, DW(qM), K.ms.nv_pad, nv, lane, nullptr, 0.f);
    warp_cholesky<false>(s.L, s.ld, nv, s.rd, lane, mass_env(K, mb));
    store_qLD(K, w, lane, s.L, s.ld);
  } else {
    load_rows(s.L, s.ld, DW(qLD), nv, nv, lane, nullptr, 0.f);
    chol_rdiag(s.L, s.ld, nv, s.rd, lane);
  }
  __syncwarp();
  warp_chol_solve<false>(s.L, s.ld, nv, s.rd, x, lane, mass_env(K, mb));
  for (int i = lane; i < nv; i += MJB_LW) DW(qacc_smooth)[i] = x[i];
  __syncwarp();
}

// derivative.py:31-108 _qderiv_actuator_passive_vel: d(actuator force) / d(actuator velocity)
__device__ __forceinline__ float act_vel_deriv(const KParams& K, const unsigned char* mb, int w, int u) {
  const float gain = MI(actuator_gaintype)[u] == 1 ? MR(actuator_gainprm)[10 * u + 2] : 0.f;
  const float bias = MI(actuator_biastype)[u] == 1 ? MR(actuator_biasprm)[10 * u + 2] : 0.f;
  if (bias == 0.f && gain == 0.f) return 0.f;
  if (MI(actuator_forcelimited)[u]) {  // a clamped force does not depend on the velocity
    const float f = DW(actuator_force)[u];
    if (f <= MR(actuator_forcerange)[2 * u] || f >= MR(actuator_forcerange)[2 * u + 1]) return 0.f;
  }
  float vel = bias;
  if (MI(actuator_dyntype)[u] != 0) {
    if (gain != 0.f) {
      const int adr = MI(actuator_actadr)[u] + MI(actuator_actnum)[u] - 1;
      float act = DW(act)[adr];
      if (MI(actuator_actearly)[u])
        act = next_act(K.o.timestep, MI(actuator_dyntype)[u], MR(actuator_dynprm) + 10 * u, MR(actuator_actrange) + 2 * u, act,
                       DW(act_dot)[adr], 1.0f, MI(actuator_actlimited)[u] != 0);
      vel += gain * act;
    }
  } else if (gain != 0.f) {
    vel += gain * DW(ctrl)[u];
  }
  return vel;
}

// forward.euler (forward.py:327-354) and forward.implicit (implicitfast, forward.py:495-512 with
// derivative.deriv_smooth_vel dense, derivative.py:321-420), both followed by _advance (:213-274).
//   euler:        (qM + dt diag(damping)) qacc' = efc_Ma
//   implicitfast: (qM - dt qDeriv) qacc' = efc_Ma,  qDeriv = moment^T diag(dforce/dvel) moment - diag(damping)
__device__ __forceinline__ void s_integrate(const KParams& K, const unsigned char* mb, int w, int lane, float* scr, bool implicit) {
  const int nv = K.ms.nv;
  const CholScr s = chol_scr(scr, nv, K.ms.nv_pad);
  const float dt = K.o.timestep;
  float* x = s.x;
  const int dsbl = K.o.disableflags;
  const bool damper = !(dsbl & DSBL_DAMPER);
  const int all3 = DSBL_ACTUATION | DSBL_SPRING | DSBL_DAMPER;
  const bool solve = implicit ? (dsbl & all3) != all3 : (damper && !(dsbl & DSBL_EULERDAMP));
  if (solve) {
    load_rows(s.L, s.ld, DW(qM), K.ms.nv_pad, nv, lane, damper ? MR(dof_damping) : nullptr, dt);
    if (implicit && K.ms.nu > 0 && !(dsbl & DSBL_ACTUATION)) {
      for (int u = 0; u < K.ms.nu; u++) {  // rows are short (one entry for joint transmissions)
        const float vel = act_vel_deriv(K, mb, w, u);
        if (MJB_LW == 32 && vel == 0.f) continue;  // world-dependent (ctrl): only a full-warp world may skip the __syncwarp below
        const int adr = DW(moment_rowadr)[u], nnz = DW(moment_rownnz)[u];
        for (int idx = lane; idx < nnz * nnz; idx += MJB_LW) {
          const int a = idx / nnz, b = idx - a * nnz;
          s.L[DW(moment_colind)[adr + a] * s.ld + DW(moment_colind)[adr + b]] -= dt * DW(actuator_moment)[adr + a] * DW(actuator_moment)[adr + b] * vel;
        }
        __syncwarp();
      }
    }
    warp_cholesky<false>(s.L, s.ld, nv, s.rd, lane, mass_env(K, mb));
    for (int i = lane; i < nv; i += MJB_LW) x[i] = DW(efc_Ma)[i];
    __syncwarp();
    warp_chol_solve<false>(s.L, s.ld, nv, s.rd, x, lane, mass_env(K, mb));
  } else {
    for (int i = lane; i < nv; i += MJB_LW) x[i] = DW(qacc)[i];
    __syncwarp();
  }
  for (int u = lane; u < K.ms.nu; u += MJB_LW) {
    const int a0 = MI(actuator_actadr)[u];
    if (a0 < 0) continue;
    for (int j = a0; j < a0 + MI(actuator_actnum)[u]; j++)
      DW(act)[j] = next_act(dt, MI(actuator_dyntype)[u], MR(actuator_dynprm) + 10 * u, MR(actuator_actrange) + 2 * u, DW(act)[j],
                            DW(act_dot)[j], 1.0f, MI(actuator_actlimited)[u] != 0);
  }
  float* qvel = DW(qvel);
  float* qpos = DW(qpos);
  for (int i = lane; i < nv; i += MJB_LW) {
    qvel[i] += x[i] * dt;
    DW(qacc_warmstart)[i] = DW(qacc)[i];
  }
  __syncwarp();
  for (int j = lane; j < K.ms.njnt; j += MJB_LW) {  // forward.py:52-112
    const int qa = MI(jnt_qposadr)[j], da = MI(jnt_dofadr)[j], jt = MI(jnt_type)[j];
    if (jt == JNT_FREE) {
      qpos[qa] += dt * qvel[da];
      qpos[qa + 1] += dt * qvel[da + 1];
      qpos[qa + 2] += dt * qvel[da + 2];
      st4(qpos + qa + 3, quat_integrate(ld4(qpos + qa + 3), ld3(qvel + da + 3), dt));
    } else if (jt == JNT_BALL) {
      st4(qpos + qa, quat_integrate(ld4(qpos + qa), ld3(qvel + da), dt));
    } else {
      qpos[qa] += dt * qvel[da];
    }
  }
  if (lane == 0) {
    DW(time)[0] += dt;
    int st = 0;
    if (DW(nefc)[0] > K.dsz.njmax) st |= MJB_STATUS_NEFC_OVERFLOW;  // forward.py:194
    if (st) DW(status)[0] |= st;
  }
  __syncwarp();
}

// sensor.energy_pos (sensor.py:2854): potential energy = -sum_b m_b g . xipos_b + joint springs (:2704-2812); energy[1] = 0.
static __device__ __noinline__ void s_energy_pos(const KParams& K, const unsigned char* mb, int w, int lane) {
  float e = 0.f;
  if (!(K.o.disableflags & DSBL_GRAVITY)) {
    const V3 g{K.o.gravity_x, K.o.gravity_y, K.o.gravity_z};
    for (int b = 1 + lane; b < K.ms.nbody; b += MJB_LW) e -= MR(body_mass)[b] * dot(g, ld3(DW(xipos) + 3 * b));
  }
  if (!(K.o.disableflags & DSBL_SPRING)) {
    const float* qpos = DW(qpos);
    for (int j = lane; j < K.ms.njnt; j += MJB_LW) {
      const float k = MR(jnt_stiffness)[j];
      if (k == 0.f) continue;
      const int qa = MI(jnt_qposadr)[j], t = MI(jnt_type)[j];
      const float* qs = MR(qpos_spring) + qa;
      if (t == JNT_FREE || t == JNT_BALL) {
        float s = 0.f;
        int qo = 0;
        if (t == JNT_FREE) {
          const V3 d0 = ld3(qpos + qa) - ld3(qs);
          s = dot(d0, d0);
          qo = 3;
        }
        const V3 dif = quat_sub(normalize(ld4(qpos + qa + qo)), ld4(qs + qo));
        e += 0.5f * k * (s + dot(dif, dif));
      } else {
        const float dq = qpos[qa] - qs[0];
        e += 0.5f * k * dq * dq;
      }
    }
  }
  e = lw_sum(e);
  if (lane == 0) { DW(energy)[0] = e; DW(energy)[1] = 0.f; }
}

// sensor.energy_vel (sensor.py:2922): kinetic energy = 0.5 qvel^T M qvel (dense qM)
static __device__ __noinline__ void s_energy_vel(const KParams& K, int w, int lane) {
  const int nv = K.ms.nv, np = K.ms.nv_pad;
  const float* qvel = DW(qvel);
  float e = 0.f;
  for (int i = lane; i < nv; i += MJB_LW) {
    const float* row = DW(qM) + (size_t)i * np;
    float s = 0.f;
    for (int j = 0; j < nv; j++) s += row[j] * qvel[j];
    e += qvel[i] * s;
  }
  e = lw_sum(e);
  if (lane == 0) DW(energy)[1] = 0.5f * e;
}
