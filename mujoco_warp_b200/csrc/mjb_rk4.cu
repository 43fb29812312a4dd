// mjb_rk4.cu -- Runge-Kutta 4 integrator (reference forward.py:458 rungekutta4) around the step pipeline.
//
// The reference drives RK4 from the host: clone qpos / qvel / act, then per stage `_rk_perturb_state` (three launches),
// `forward` (~100 launches) and `_rk_accumulate` (two launches), then `_advance` (four launches), with six temporaries allocated
// per call (forward.py:465-475).  Here one small kernel per stage does save + accumulate + perturb (or, after the last stage,
// accumulate + restore + advance), the three extra `forward` passes go through mjb_run_stages, and the temporaries live in one
// scratch buffer of rk_stride() floats per world:
//     qpos_t0 (nq) | qvel_t0 (nv) | qvel_rk (nv) | qacc_rk (nv) | act_t0 (na) | act_dot_rk (na)
// Nothing is computed on the host, nothing synchronises, and the whole step is CUDA-graph capturable once the scratch exists
// (mjb_prepare, or the first call).
//
// mjb_rungekutta4_static is the same integrator for callers that must not allocate: its scratch is a file-scope device array.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>

#include "mjb_internal.h"

#define CUDA_OK(expr)                                                                                              \
  do {                                                                                                             \
    cudaError_t e__ = (expr);                                                                                      \
    if (e__ != cudaSuccess) return mjb_internal_fail(-10, (std::string(#expr) + ": " + cudaGetErrorString(e__)).c_str()); \
  } while (0)

// floats of scratch per world, rounded to 16 bytes
static inline int rk_stride(const ModelSizes& ms) { return (ms.nq + 3 * ms.nv + 2 * ms.na + 3) & ~3; }

struct RkArgs {
  float* scr;  // nworld x stride
  int stride;
  int stage;  // 0: save t0 + first accumulate + perturb; 1, 2: accumulate + perturb; 3: accumulate + restore + advance
  float a;    // tableau diagonal of the NEXT stage (perturbation scale), stages 0-2
  float b;    // tableau weight of THIS stage's derivative
};

// forward.py:52-112 _next_position: qpos_out = integrate(qpos_in, qvel * scale, timestep), lane = joint
__device__ __forceinline__ void rk_next_position(const KParams& K, const unsigned char* mb, int lane, const float* qpos_in, const float* qvel,
                                                 float scale, float dt, float* qpos_out) {
  for (int j = lane; j < K.ms.njnt; j += 32) {
    const int qa = MI(jnt_qposadr)[j], da = MI(jnt_dofadr)[j], jt = MI(jnt_type)[j];
    if (jt == JNT_FREE) {
      const V3 lin = ld3(qvel + da) * scale, ang = ld3(qvel + da + 3) * scale;
      st3(qpos_out + qa, ld3(qpos_in + qa) + dt * lin);
      st4(qpos_out + qa + 3, quat_integrate(ld4(qpos_in + qa + 3), ang, dt));
    } else if (jt == JNT_BALL) {
      st4(qpos_out + qa, quat_integrate(ld4(qpos_in + qa), ld3(qvel + da) * scale, dt));
    } else {
      qpos_out[qa] = qpos_in[qa] + dt * qvel[da] * scale;
    }
  }
}

// forward.py:130-167 _next_activation, lane = actuator
__device__ __forceinline__ void rk_next_activation(const KParams& K, const unsigned char* mb, int lane, const float* act_in, const float* act_dot,
                                                   float scale, bool limit, float dt, float* act_out) {
  for (int u = lane; u < K.ms.nu; u += 32) {
    const int a0 = MI(actuator_actadr)[u];
    if (a0 < 0) continue;
    for (int j = a0; j < a0 + MI(actuator_actnum)[u]; j++)
      act_out[j] = next_act(dt, MI(actuator_dyntype)[u], MR(actuator_dynprm) + 10 * u, MR(actuator_actrange) + 2 * u, act_in[j], act_dot[j], scale,
                            limit && MI(actuator_actlimited)[u] != 0);
  }
}

// One warp per world.  The model tables are read from the global blob (a few hundred bytes per world, L2 resident).
__global__ void __launch_bounds__(256) mjb_rk_kernel(const __grid_constant__ KParams K, const RkArgs A) {
  const unsigned char* mb = K.blob;
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const int nq = K.ms.nq, nv = K.ms.nv, na = K.ms.na;
  const float dt = K.o.timestep;
  for (int w = blockIdx.x * wpb + (threadIdx.x >> 5); w < K.dsz.nworld; w += gridDim.x * wpb) {
    float* s = A.scr + (size_t)w * A.stride;
    float *qpos_t0 = s, *qvel_t0 = s + nq, *qvel_rk = qvel_t0 + nv, *qacc_rk = qvel_rk + nv, *act_t0 = qacc_rk + nv, *act_dot_rk = act_t0 + na;
    float *qpos = DW(qpos), *qvel = DW(qvel), *qacc = DW(qacc), *act = DW(act), *act_dot = DW(act_dot);
    // forward.py:404-456 _rk_accumulate (stage 0 initialises: the reference zero-fills and accumulates B[0])
    if (A.stage == 0) {
      for (int i = lane; i < nq; i += 32) qpos_t0[i] = qpos[i];
      for (int i = lane; i < nv; i += 32) {
        qvel_t0[i] = qvel[i];
        qvel_rk[i] = A.b * qvel[i];
        qacc_rk[i] = A.b * qacc[i];
      }
      for (int i = lane; i < na; i += 32) {
        act_t0[i] = act[i];
        act_dot_rk[i] = A.b * act_dot[i];
      }
    } else {
      for (int i = lane; i < nv; i += 32) {
        qvel_rk[i] += A.b * qvel[i];
        qacc_rk[i] += A.b * qacc[i];
      }
      for (int i = lane; i < na; i += 32) act_dot_rk[i] += A.b * act_dot[i];
    }
    __syncwarp();
    if (A.stage < 3) {  // forward.py:355-400 _rk_perturb_state: position from the CURRENT stage velocity, then velocity, then activation
      rk_next_position(K, mb, lane, qpos_t0, qvel, A.a, dt, qpos);
      __syncwarp();
      for (int i = lane; i < nv; i += 32) qvel[i] = qvel_t0[i] + A.a * qacc[i] * dt;
      if (na) rk_next_activation(K, mb, lane, act_t0, act_dot, A.a, false, dt, act);
    } else {  // forward.py:482-491: restore the state, then _advance(m, d, qacc_rk, qvel_rk) (:213-274)
      if (na) {
        for (int i = lane; i < na; i += 32) act_dot[i] = act_dot_rk[i];
        rk_next_activation(K, mb, lane, act_t0, act_dot_rk, 1.0f, true, dt, act);
      }
      for (int i = lane; i < nv; i += 32) {
        qvel[i] = qvel_t0[i] + 1.0f * qacc_rk[i] * dt;
        DW(qacc_warmstart)[i] = qacc[i];  // forward.py:274 copies d.qacc: the LAST stage's acceleration
      }
      rk_next_position(K, mb, lane, qpos_t0, qvel_rk, 1.0f, dt, qpos);
      if (lane == 0) {
        DW(time)[0] += dt;
        if (DW(nefc)[0] > K.dsz.njmax) DW(status)[0] |= MJB_STATUS_NEFC_OVERFLOW;  // forward.py:194
      }
    }
    __syncwarp();
  }
}

static int rk_run(const MjbModel* m, MjbData* d, const MjbOption* opt, float* scr, cudaStream_t stream) {
  KParams K;
  int rc = mjb_internal_params(m, d, &K);
  if (rc) return rc;
  static_assert(sizeof(OptView) == sizeof(MjbOption), "option struct mismatch");
  memcpy(&K.o, opt, sizeof(OptView));
  const float A[3] = {0.5f, 0.5f, 1.0f};  // RK4 tableau (diagonal only), forward.py:461-462
  const float B[4] = {1.0f / 6.0f, 1.0f / 3.0f, 1.0f / 3.0f, 1.0f / 6.0f};
  RkArgs a;
  a.scr = scr;
  a.stride = rk_stride(K.ms);
  const int nworld = K.dsz.nworld;
  if (nworld <= 0) return 0;
  const int wpb = 8;
  int grid = (nworld + wpb - 1) / wpb;
  if (grid > 148 * 8) grid = 148 * 8;
  for (int stage = 0; stage < 4; stage++) {
    a.stage = stage;
    a.a = stage < 3 ? A[stage] : 0.f;
    a.b = B[stage];
    mjb_rk_kernel<<<grid, 32 * wpb, 0, stream>>>(K, a);
    mjb_internal_count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return mjb_internal_fail(-11, (std::string("mjb_rk_kernel launch: ") + cudaGetErrorString(e)).c_str());
    if (stage < 3) {
      rc = mjb_forward(m, d, opt, stream);
      if (rc) return rc;
    }
  }
  return 0;
}

extern "C" long long mjb_rk_scratch_floats(const MjbModel* m, const MjbData* d) {
  KParams K;
  if (mjb_internal_params(m, d, &K)) return -1;
  return (long long)rk_stride(K.ms) * K.dsz.nworld;
}

extern "C" int mjb_rungekutta4(const MjbModel* m, MjbData* d, const MjbOption* opt, void* s) {
  if (!m || !d || !opt) return mjb_internal_fail(-1, "null handle");
  const long long n = mjb_rk_scratch_floats(m, d);
  if (n < 0) return -1;
  float* scr = mjb_internal_rk_scratch(d, n > 0 ? (size_t)n : 1);  // allocates on the first call only (mjb_prepare does it ahead of a capture)
  if (!scr) return -10;
  return rk_run(m, d, opt, scr, static_cast<cudaStream_t>(s));
}

// ---- allocation-free variant -----------------------------------------------------------------------------------------------------
// Scratch of mjb_rungekutta4_static: resident on every GPU that loads the library until the process exits.
static __device__ __align__(256) float g_rk_static_scratch[MJB_RK_STATIC_MAX_FLOATS];

// Device address of g_rk_static_scratch on the CURRENT device (in host code the array's name is only a handle for the runtime,
// not an address a kernel can use), resolved once per device.
static int rk_static_address(float** out) {
  enum { MAX_DEVICES = 64 };
  static std::mutex mu;
  static float* ptr[MAX_DEVICES] = {nullptr};
  int dev = -1;
  CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES) return mjb_internal_fail(-2, "mjb_rungekutta4_static: device ordinal out of range");
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_rk_static_scratch);
    if (e != cudaSuccess || !p)
      return mjb_internal_fail(-10, (std::string("cudaGetSymbolAddress(g_rk_static_scratch): ") + cudaGetErrorString(e)).c_str());
    ptr[dev] = static_cast<float*>(p);
  }
  *out = ptr[dev];
  return 0;
}

extern "C" long long mjb_rk_static_max_floats(void) { return (long long)MJB_RK_STATIC_MAX_FLOATS; }

extern "C" int mjb_rungekutta4_static(const MjbModel* m, MjbData* d, const MjbOption* opt, void* s) {
  if (!m || !d || !opt) return mjb_internal_fail(-1, "null handle");
  const long long n = mjb_rk_scratch_floats(m, d);
  if (n < 0) return -1;
  if (n > (long long)MJB_RK_STATIC_MAX_FLOATS)
    return mjb_internal_fail(-6, ("mjb_rungekutta4_static: nworld x (nq + 3 nv + 2 na) = " + std::to_string(n) + " floats exceeds the static scratch of " +
                                  std::to_string((long long)MJB_RK_STATIC_MAX_FLOATS) + " floats (MJB_RK_STATIC_MAX_FLOATS); use mjb_rungekutta4")
                                     .c_str());
  float* scr = nullptr;
  const int rc = rk_static_address(&scr);
  if (rc) return rc;
  // nothing to clear: stage 0 overwrites every scratch word this call reads
  return rk_run(m, d, opt, scr, static_cast<cudaStream_t>(s));
}
