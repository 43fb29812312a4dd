// mjb_types.cuh -- kernel parameter block, accessors and small vector math.
//
// Execution model: ONE WARP PER WORLD.  The reference launches one Warp kernel per stage (and per
// tree level) with one thread per (world, element) and float atomics (reference
// mujoco_warp/_src/forward.py:973-1018 drives ~100-150 launches per step).  Here a persistent
// grid of warps walks the worlds; each warp runs every selected stage for its world, lanes map to
// bodies / dofs / geoms / constraint rows, and all cross-lane traffic is __syncwarp + shuffles.
// Data fields keep the reference's (nworld, n, ...) row-major layout, so a warp touching
// consecutive elements of its world's row issues fully coalesced 128-byte transactions.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "mjb_fields.h"

// lanes per world of the smooth-dynamics stage groups (smooth_position, smooth_velocity, euler): see MJB_LW in mjb_smooth.cuh
#ifndef MJB_SMOOTH_LW
#define MJB_SMOOTH_LW 32
#endif
// per group (default: the common value); MJB_LW_WPB = warps per CTA of a group that carries several worlds per warp
#ifndef MJB_POS_LW
#define MJB_POS_LW MJB_SMOOTH_LW
#endif
#ifndef MJB_VEL_LW
#define MJB_VEL_LW 8  // measured (B200, humanoid 8192 worlds): smooth_velocity 111 -> 95 us with 8 lanes per world and 7 warps per CTA
#endif
#ifndef MJB_EUL_LW
#define MJB_EUL_LW 8  // euler 20 -> 17 us
#endif
#ifndef MJB_LW_WPB
#define MJB_LW_WPB 7  // warps per CTA of those groups (28 worlds per CTA); 4 warps: 153 us
#endif

#define MJ_MINVAL 1e-15f
#define MJ_MAXVAL 1e10f
#define MJ_MINIMP 1e-4f
#define MJ_MAXIMP 0.9999f
#define MJ_MINMU 1e-5f
#define MJ_PI 3.14159265358979323846f

enum { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum { GEOM_PLANE = 0, GEOM_HFIELD = 1, GEOM_SPHERE = 2, GEOM_CAPSULE = 3, GEOM_ELLIPSOID = 4, GEOM_CYLINDER = 5, GEOM_BOX = 6, GEOM_MESH = 7 };
enum {
  DSBL_CONSTRAINT = 1, DSBL_EQUALITY = 2, DSBL_FRICTIONLOSS = 4, DSBL_LIMIT = 8, DSBL_CONTACT = 16, DSBL_SPRING = 32,
  DSBL_DAMPER = 64, DSBL_GRAVITY = 128, DSBL_CLAMPCTRL = 256, DSBL_WARMSTART = 512, DSBL_ACTUATION = 2048,
  DSBL_REFSAFE = 4096, DSBL_EULERDAMP = 32768, DSBL_NATIVECCD = 131072
};
enum { CT_EQUALITY = 0, CT_FRICTION_DOF = 1, CT_LIMIT_JOINT = 3, CT_CONTACT_FRICTIONLESS = 5, CT_CONTACT_PYRAMIDAL = 6, CT_CONTACT_ELLIPTIC = 7 };
enum { ST_SATISFIED = 0, ST_QUADRATIC = 1, ST_LINEARNEG = 2, ST_LINEARPOS = 3, ST_CONE = 4 };
enum { SOLVER_CG = 1, SOLVER_NEWTON = 2 };
enum { BF_PLANE = 1, BF_SPHERE = 2, BF_AABB = 4, BF_OBB = 8 };

struct ModelSizes {
#define X(n) int n;
  MJB_MODEL_SIZES(X)
#undef X
};
struct ModelOffsets {  // byte offsets into the model blob
#define X(n) int n;
  MJB_MODEL_INT_FIELDS(X)
  MJB_MODEL_REAL_FIELDS(X)
#undef X
};
struct DataSizes {
#define X(n) int n;
  MJB_DATA_SIZES(X)
#undef X
};
struct DataPtrs {
#define X(n, c) int* n;
  MJB_DATA_INT_FIELDS(X)
#undef X
#define X(n, c) float* n;
  MJB_DATA_REAL_FIELDS(X)
#undef X
};
struct DataStrides {  // per-world element counts
#define X(n, c) int n;
  MJB_DATA_INT_FIELDS(X)
  MJB_DATA_REAL_FIELDS(X)
#undef X
};
struct OptView {
#define X(n) int n;
  MJB_OPT_INT_FIELDS(X)
#undef X
#define X(n) float n;
  MJB_OPT_REAL_FIELDS(X)
#undef X
};

// Kernel parameter block (lives in the constant bank; < 4 KB).
struct KParams {
  ModelSizes ms;
  ModelOffsets mo;
  DataSizes dsz;
  DataPtrs d;
  DataStrides st;
  OptView o;
  const unsigned char* blob;  // global copy of the model blob
  int blob_bytes;             // staged prefix of the blob (multiple of 16): everything, or the hot tables of a large model
  int cold_global;            // 1: the cold tables (pair lists, mesh vertices / graphs) stay in global memory behind the staged prefix
  int scratch_floats;         // per-warp shared-memory scratch
  uint32_t stages;
  float* hbuf;  // nv > 32: the solver's Hessians, nworld x nv x chol_ld floats in global memory (they would halve the warps per SM in shared memory); else null
  int* queue;  // dynamic world queue of this launch (zeroed by the host before the group kernels run), or null: static striding
  // optional extra operands for the small utility kernels
  float* x;
  const float* y;
};

// model accessors: `mb` is the base of the model blob (shared memory inside the main kernel)
#define MI(name) (reinterpret_cast<const int*>(mb + K.mo.name))
#define MR(name) (reinterpret_cast<const float*>(mb + K.mo.name))
// cold tables (large, streamed or rarely touched): read through `mb` when the whole blob is staged, else from the global copy
#define MIC(name) (reinterpret_cast<const int*>((K.cold_global ? K.blob : mb) + K.mo.name))
#define MRC(name) (reinterpret_cast<const float*>((K.cold_global ? K.blob : mb) + K.mo.name))
// per-world data pointer
#define DW(name) (K.d.name + (size_t)w * (size_t)K.st.name)

#define FULL 0xffffffffu

// ---------------------------------------------------------------------------------------------
// small vector math (reference math.py:24-334); quaternions are (w, x, y, z)
// ---------------------------------------------------------------------------------------------
struct V3 {
  float x, y, z;
};
struct Q4 {
  float w, x, y, z;
};
struct M3 {  // row major
  float m[9];
};
struct S6 {  // spatial vector (rot[3], lin[3])
  float v[6];
};

__device__ __forceinline__ V3 mk3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 ld3(const float* p) { return V3{p[0], p[1], p[2]}; }
__device__ __forceinline__ void st3(float* p, V3 a) { p[0] = a.x; p[1] = a.y; p[2] = a.z; }
__device__ __forceinline__ Q4 ld4(const float* p) { return Q4{p[0], p[1], p[2], p[3]}; }
__device__ __forceinline__ void st4(float* p, Q4 q) { p[0] = q.w; p[1] = q.x; p[2] = q.y; p[3] = q.z; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return V3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ float norm(V3 a) { return sqrtf(dot(a, a)); }
// wp.normalize semantics: a zero vector stays zero
__device__ __forceinline__ V3 normalize(V3 a) {
  float n = norm(a);
  return n > 0.f ? a * (1.0f / n) : V3{0.f, 0.f, 0.f};
}
__device__ __forceinline__ Q4 normalize(Q4 q) {
  float n = sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
  if (n > 0.f) {
    float s = 1.0f / n;
    return Q4{q.w * s, q.x * s, q.y * s, q.z * s};
  }
  return Q4{0.f, 0.f, 0.f, 0.f};
}
// math.py:24
__device__ __forceinline__ Q4 mul_quat(Q4 u, Q4 v) {
  return Q4{u.w * v.w - u.x * v.x - u.y * v.y - u.z * v.z, u.w * v.x + u.x * v.w + u.y * v.z - u.z * v.y,
            u.w * v.y - u.x * v.z + u.y * v.w + u.z * v.x, u.w * v.z + u.x * v.y - u.y * v.x + u.z * v.w};
}
// math.py:45
__device__ __forceinline__ V3 rot_vec_quat(V3 v, Q4 q) {
  V3 u = V3{q.x, q.y, q.z};
  float s = q.w;
  V3 r = 2.0f * (dot(u, v) * u) + (s * s - dot(u, u)) * v;
  return r + 2.0f * s * cross(u, v);
}
// math.py:53
__device__ __forceinline__ Q4 axis_angle_to_quat(V3 axis, float angle) {
  float s, c;
  sincosf(angle * 0.5f, &s, &c);
  return Q4{c, axis.x * s, axis.y * s, axis.z * s};
}
// math.py:60
__device__ __forceinline__ M3 quat_to_mat(Q4 q) {
  float q00 = q.w * q.w, q01 = q.w * q.x, q02 = q.w * q.y, q03 = q.w * q.z;
  float q11 = q.x * q.x, q12 = q.x * q.y, q13 = q.x * q.z, q22 = q.y * q.y, q23 = q.y * q.z, q33 = q.z * q.z;
  M3 r;
  r.m[0] = q00 + q11 - q22 - q33; r.m[1] = 2.0f * (q12 - q03); r.m[2] = 2.0f * (q13 + q02);
  r.m[3] = 2.0f * (q12 + q03); r.m[4] = q00 - q11 + q22 - q33; r.m[5] = 2.0f * (q23 - q01);
  r.m[6] = 2.0f * (q13 - q02); r.m[7] = 2.0f * (q23 + q01); r.m[8] = q00 - q11 - q22 + q33;
  return r;
}
__device__ __forceinline__ void stm3(float* p, const M3& a) {
#pragma unroll
  for (int i = 0; i < 9; i++) p[i] = a.m[i];
}
__device__ __forceinline__ M3 ldm3(const float* p) {
  M3 a;
#pragma unroll
  for (int i = 0; i < 9; i++) a.m[i] = p[i];
  return a;
}
__device__ __forceinline__ V3 mat_vec(const M3& a, V3 v) {
  return V3{a.m[0] * v.x + a.m[1] * v.y + a.m[2] * v.z, a.m[3] * v.x + a.m[4] * v.y + a.m[5] * v.z,
            a.m[6] * v.x + a.m[7] * v.y + a.m[8] * v.z};
}
__device__ __forceinline__ V3 mat_col(const M3& a, int c) { return V3{a.m[c], a.m[3 + c], a.m[6 + c]}; }

__device__ __forceinline__ S6 ld6(const float* p) {
  S6 s;
#pragma unroll
  for (int i = 0; i < 6; i++) s.v[i] = p[i];
  return s;
}
__device__ __forceinline__ void st6(float* p, const S6& s) {
#pragma unroll
  for (int i = 0; i < 6; i++) p[i] = s.v[i];
}
__device__ __forceinline__ V3 top(const S6& s) { return V3{s.v[0], s.v[1], s.v[2]}; }
__device__ __forceinline__ V3 bot(const S6& s) { return V3{s.v[3], s.v[4], s.v[5]}; }
__device__ __forceinline__ S6 mk6(V3 a, V3 b) { return S6{{a.x, a.y, a.z, b.x, b.y, b.z}}; }
__device__ __forceinline__ float dot6(const S6& a, const S6& b) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 6; i++) s += a.v[i] * b.v[i];
  return s;
}
__device__ __forceinline__ void axpy6(S6& y, float a, const S6& x) {
#pragma unroll
  for (int i = 0; i < 6; i++) y.v[i] += a * x.v[i];
}
// math.py:121 (i points at a 10-vector)
__device__ __forceinline__ S6 inert_vec(const float* i, const S6& s) {
  const float* v = s.v;
  S6 r;
  r.v[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  r.v[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  r.v[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  r.v[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  r.v[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  r.v[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
  return r;
}
// math.py:134
__device__ __forceinline__ S6 motion_cross(const S6& u, const S6& v) {
  return mk6(cross(top(u), top(v)), cross(bot(u), top(v)) + cross(top(u), bot(v)));
}
// math.py:148
__device__ __forceinline__ S6 motion_cross_force(const S6& v, const S6& f) {
  return mk6(cross(top(v), top(f)) + cross(bot(v), bot(f)), cross(top(v), bot(f)));
}
// math.py:162
__device__ __forceinline__ V3 quat_to_vel(Q4 q) {
  V3 axis = V3{q.x, q.y, q.z};
  float s = norm(axis);
  if (s == 0.f) return V3{0.f, 0.f, 0.f};
  float speed = 2.0f * atan2f(s, q.w);
  if (speed > MJ_PI) speed -= 2.0f * MJ_PI;
  return axis * (speed / s);
}
// math.py:178
__device__ __forceinline__ V3 quat_sub(Q4 qa, Q4 qb) { return quat_to_vel(mul_quat(Q4{qb.w, -qb.x, -qb.y, -qb.z}, qa)); }
// math.py:189
__device__ __forceinline__ Q4 quat_integrate(Q4 q, V3 v, float dt) {
  float n = norm(v);
  Q4 qr = axis_angle_to_quat(normalize(v), dt * n);
  return normalize(mul_quat(normalize(q), qr));
}
__device__ __forceinline__ float safe_div(float x, float y) { return x / (y != 0.f ? y : MJ_MINVAL); }
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
// support.next_act (support.py:38-64)
__device__ __forceinline__ float next_act(float dt, int dyntype, const float* dynprm, const float* actrange, float act, float act_dot,
                                          float scale, bool clamp) {
  if (dyntype == 3) {
    const float tau = fmaxf(MJ_MINVAL, dynprm[0]);
    act = act + scale * act_dot * tau * (1.0f - expf(-dt / tau));
  } else if (dyntype == 5) {
    return act;
  } else {
    act = act + scale * act_dot * dt;
  }
  if (clamp) act = clampf(act, actrange[0], actrange[1]);
  return act;
}

// math.py:203, 247: frame rows (a, b, c)
__device__ __forceinline__ M3 make_frame(V3 a) {
  a = normalize(a);
  V3 b = (-0.5f < a.y && a.y < 0.5f) ? V3{0.f, 1.f, 0.f} : V3{0.f, 0.f, 1.f};
  b = normalize(b - a * dot(a, b));
  if (norm(a) == 0.f) b = V3{0.f, 0.f, 0.f};
  V3 c = cross(a, b);
  M3 f;
  f.m[0] = a.x; f.m[1] = a.y; f.m[2] = a.z; f.m[3] = b.x; f.m[4] = b.y; f.m[5] = b.z; f.m[6] = c.x; f.m[7] = c.y; f.m[8] = c.z;
  return f;
}

// warp helpers
// Broadcast from lane 0: tells the compiler that a value that is equal on all lanes by construction
// (a load from a warp-uniform address, the result of a full reduction) is warp-uniform, so branches
// on it need no divergence handling.
__device__ __forceinline__ int uniform_i(int v) { return __shfl_sync(FULL, v, 0); }
__device__ __forceinline__ bool uniform_b(bool v) { return __shfl_sync(FULL, v ? 1 : 0, 0) != 0; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
// exclusive prefix sum over lanes
__device__ __forceinline__ int warp_excl_scan(int v, int lane) {
  int s = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(FULL, s, o);
    if (lane >= o) s += t;
  }
  return s - v;
}
