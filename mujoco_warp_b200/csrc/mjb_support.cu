// mjb_support.cu -- the callers either side of `step`: state packing, masked reset, contact forces.
//   mjb_get_state / mjb_set_state   reference support.py:572 / :711   (MuJoCo mj_getState / mj_setState layout)
//   mjb_reset_data                  reference io.py:1458-1640
//   mjb_contact_force               reference support.py:354 (contact_force_kernel :317, _decode_pyramid :239)
// All are bandwidth-bound copies: one launch each, asynchronous on the caller's stream, CUDA-graph capturable (the per-world
// masks stay on the device; nothing is read back).
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "mjb_internal.h"

#define CUDA_OK(expr)                                                                                              \
  do {                                                                                                             \
    cudaError_t e__ = (expr);                                                                                      \
    if (e__ != cudaSuccess) return mjb_internal_fail(-10, (std::string(#expr) + ": " + cudaGetErrorString(e__)).c_str()); \
  } while (0)

static int launched(void) {
  mjb_internal_count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return mjb_internal_fail(-11, (std::string("kernel launch: ") + cudaGetErrorString(e)).c_str());
  return 0;
}

// ---- state packing ---------------------------------------------------------------------------
// One segment per State bit (types.py:598-637): `n` floats per world at column `off` of the packed row.
enum { MJB_NSTATE_BITS = 13 };
struct StateSeg {
  float* real;  // float field (per-world stride `stride`), or
  int* integer; // eq_active (stored as int32; packed as 0.0 / 1.0)
  int n, stride, off;
};
struct StateLayout {
  StateSeg seg[11];
  int nseg, nstate;
};

static StateLayout state_layout(const KParams& K, int sig) {
  StateLayout L;
  L.nseg = 0;
  L.nstate = 0;
  auto add = [&](int bit, float* p, int* ip, int n, int stride) {
    if (!(sig & (1 << bit))) return;
    if (n > 0) L.seg[L.nseg++] = StateSeg{p, ip, n, stride, L.nstate};
    L.nstate += n;
  };
  add(0, K.d.time, nullptr, 1, K.st.time);
  add(1, K.d.qpos, nullptr, K.ms.nq, K.st.qpos);
  add(2, K.d.qvel, nullptr, K.ms.nv, K.st.qvel);
  add(3, K.d.act, nullptr, K.ms.na, K.st.act);
  add(4, K.d.qacc_warmstart, nullptr, K.ms.nv, K.st.qacc_warmstart);
  add(5, K.d.ctrl, nullptr, K.ms.nu, K.st.ctrl);
  add(6, K.d.qfrc_applied, nullptr, K.ms.nv, K.st.qfrc_applied);
  add(7, K.d.xfrc_applied, nullptr, 6 * K.ms.nbody, K.st.xfrc_applied);
  add(8, nullptr, K.d.eq_active, K.ms.neq, K.st.eq_active);
  add(9, K.d.mocap_pos, nullptr, 3 * K.ms.nmocap, K.st.mocap_pos);
  add(10, K.d.mocap_quat, nullptr, 4 * K.ms.nmocap, K.st.mocap_quat);
  return L;  // bits 11 (userdata) and 12 (plugin) carry no data here, as in the reference (types.py:638)
}

template <bool SET>
__global__ void __launch_bounds__(256) mjb_state_kernel(const StateLayout L, int nworld, float* __restrict__ state, int ld,
                                                        const uint8_t* __restrict__ active) {
  const long long total = (long long)nworld * L.nstate;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int w = (int)(t / L.nstate), j = (int)(t - (long long)w * L.nstate);
    if (active && !active[w]) continue;
    int s = 0;
#pragma unroll
    for (int k = 1; k < 11; k++) s += (k < L.nseg && j >= L.seg[k].off) ? 1 : 0;
    const StateSeg& g = L.seg[s];
    const int e = j - g.off;
    float* row = state + (size_t)w * ld + j;
    if (g.real) {
      float* p = g.real + (size_t)w * g.stride + e;
      if (SET) *p = *row;
      else *row = *p;
    } else {
      int* p = g.integer + (size_t)w * g.stride + e;
      if (SET) *p = *row != 0.f ? 1 : 0;
      else *row = *p ? 1.f : 0.f;
    }
  }
}

extern "C" int mjb_state_size(const MjbModel* m, const MjbData* d, int sig) {
  KParams K;
  if (int rc = mjb_internal_params(m, d, &K)) return rc;
  if (sig < 0 || sig >= (1 << MJB_NSTATE_BITS)) return mjb_internal_fail(-2, "invalid state signature >= 2^mjNSTATE");
  return state_layout(K, sig).nstate;
}

static int state_copy(bool set, const MjbModel* m, MjbData* d, float* state, int ld, int sig, const uint8_t* active, void* s) {
  KParams K;
  if (int rc = mjb_internal_params(m, d, &K)) return rc;
  if (sig < 0 || sig >= (1 << MJB_NSTATE_BITS)) return mjb_internal_fail(-2, "invalid state signature >= 2^mjNSTATE");
  const StateLayout L = state_layout(K, sig);
  if (L.nstate == 0 || K.dsz.nworld == 0) return 0;
  if (!state || ld < L.nstate) return mjb_internal_fail(-2, "state buffer is null or its row length is below the state size");
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  const long long total = (long long)K.dsz.nworld * L.nstate;
  int grid = (int)((total + 255) / 256);
  if (grid > 148 * 16) grid = 148 * 16;
  if (set) mjb_state_kernel<true><<<grid, 256, 0, stream>>>(L, K.dsz.nworld, state, ld, active);
  else mjb_state_kernel<false><<<grid, 256, 0, stream>>>(L, K.dsz.nworld, state, ld, active);
  return launched();
}
extern "C" int mjb_get_state(const MjbModel* m, MjbData* d, float* state, int ld, int sig, const uint8_t* active, void* s) {
  return state_copy(false, m, d, state, ld, sig, active, s);
}
extern "C" int mjb_set_state(const MjbModel* m, MjbData* d, const float* state, int ld, int sig, const uint8_t* active, void* s) {
  return state_copy(true, m, d, const_cast<float*>(state), ld, sig, active, s);
}

// ---- reset_data ------------------------------------------------------------------------------
// io.py:1601-1640 reset_contact: contacts of the reset worlds are cleared in place (the pool is not compacted; the next
// `collision` rebuilds it).  Runs BEFORE the per-world kernel, which may zero nacon, like the reference's launch order.
__global__ void __launch_bounds__(256) mjb_reset_contact_kernel(const KParams K, const uint8_t* __restrict__ reset) {
  const int nacon = min(K.d.nacon[0], K.dsz.naconmax);
  const int nmp = K.ms.nmaxpyramid;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < nacon; c += gridDim.x * blockDim.x) {
    const int w = K.d.contact_worldid[c];
    if (reset && w >= 0 && w < K.dsz.nworld && !reset[w]) continue;
    K.d.contact_dist[c] = 0.f;
    K.d.contact_includemargin[c] = 0.f;
    for (int i = 0; i < 3; i++) K.d.contact_pos[3 * c + i] = 0.f;
    for (int i = 0; i < 9; i++) K.d.contact_frame[9 * c + i] = 0.f;
    for (int i = 0; i < 5; i++) { K.d.contact_friction[5 * c + i] = 0.f; K.d.contact_solimp[5 * c + i] = 0.f; }
    for (int i = 0; i < 2; i++) { K.d.contact_solref[2 * c + i] = 0.f; K.d.contact_solreffriction[2 * c + i] = 0.f; K.d.contact_geom[2 * c + i] = 0; }
    for (int i = 0; i < nmp; i++) K.d.contact_efc_address[c * nmp + i] = -1;
    K.d.contact_dim[c] = 0;
    K.d.contact_worldid[c] = 0;
    K.d.contact_type[c] = 0;
    K.d.contact_geomcollisionid[c] = 0;
  }
}

// io.py:1470-1560 reset_xfrc_applied, reset_qM, reset_nworld, reset_mocap: one warp per world.
__global__ void __launch_bounds__(256) mjb_reset_world_kernel(const KParams K, const uint8_t* __restrict__ reset) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const unsigned char* mb = K.blob;
  for (int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < K.dsz.nworld; w += warps) {
    if (reset && !reset[w]) continue;
    if (lane == 0) {
      DW(solver_niter)[0] = 0;
      DW(ne)[0] = 0;
      DW(nf)[0] = 0;
      DW(nl)[0] = 0;
      DW(nefc)[0] = 0;
      DW(wcon_n)[0] = 0;  // library workspace: the world's contact index list is emptied with its contacts
      DW(status)[0] = 0;  // the overflow / NaN bits of a world end with its episode (they are sticky otherwise)
      DW(time)[0] = 0.f;
      DW(energy)[0] = 0.f;
      DW(energy)[1] = 0.f;
      if (w == 0) K.d.nacon[0] = 0;  // :1525 (sic: only a reset of world 0 zeroes the pooled counter)
    }
    const int nq = K.ms.nq, nv = K.ms.nv, nu = K.ms.nu, na = K.ms.na;
    for (int i = lane; i < nq; i += 32) DW(qpos)[i] = MR(qpos0)[i];
    for (int i = lane; i < nv; i += 32) {
      DW(qvel)[i] = 0.f;
      DW(qacc_warmstart)[i] = 0.f;
      DW(qfrc_applied)[i] = 0.f;
      DW(qacc)[i] = 0.f;
    }
    for (int i = lane; i < nu; i += 32) DW(ctrl)[i] = 0.f;
    for (int i = lane; i < min(nu, na); i += 32) {  // :1541-1545 (activations beyond nu are left alone, as in the reference)
      DW(act)[i] = 0.f;
      DW(act_dot)[i] = 0.f;
    }
    for (int i = lane; i < K.ms.neq; i += 32) DW(eq_active)[i] = MI(eq_active0)[i];
    for (int i = lane; i < 6 * K.ms.nbody; i += 32) DW(xfrc_applied)[i] = 0.f;
    for (int i = lane; i < K.st.qM; i += 32) DW(qM)[i] = 0.f;
    for (int b = lane; b < K.ms.nbody; b += 32) {
      const int mid = MI(body_mocapid)[b];
      if (mid < 0) continue;
      for (int i = 0; i < 3; i++) DW(mocap_pos)[3 * mid + i] = MR(body_pos)[3 * b + i];
      for (int i = 0; i < 4; i++) DW(mocap_quat)[4 * mid + i] = MR(body_quat)[4 * b + i];
    }
  }
}

extern "C" int mjb_reset_data(const MjbModel* m, MjbData* d, const uint8_t* reset, void* s) {
  KParams K;
  if (int rc = mjb_internal_params(m, d, &K)) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  if (K.dsz.nworld == 0) return 0;
  if (K.dsz.naconmax > 0) {
    int grid = (K.dsz.naconmax + 255) / 256;
    if (grid > 148 * 8) grid = 148 * 8;
    mjb_reset_contact_kernel<<<grid, 256, 0, stream>>>(K, reset);
    if (int rc = launched()) return rc;
  }
  int grid = (K.dsz.nworld + 7) / 8;
  if (grid > 148 * 8) grid = 148 * 8;
  mjb_reset_world_kernel<<<grid, 256, 0, stream>>>(K, reset);
  return launched();
}

// ---- contact_force ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128) mjb_contact_force_kernel(const KParams K, int cone, const int* __restrict__ contact_ids, int n,
                                                                int to_world_frame, float* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int c = contact_ids[t];
  if (c >= K.d.nacon[0] || c >= K.dsz.naconmax) return;  // support.py:331 (the output row is left untouched)
  float f[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c >= 0) {
    const int w = K.d.contact_worldid[c];
    const int condim = K.d.contact_dim[c], nmp = K.ms.nmaxpyramid, njmax = K.dsz.njmax;
    const int adr0 = K.d.contact_efc_address[c * nmp];
    const float* force = K.d.efc_force + (size_t)w * K.st.efc_force;
    if (adr0 >= 0) {
      if (cone == 0) {  // pyramidal: support.py:239-262
        if (condim == 1) {
          f[0] = force[adr0];
        } else {
          for (int i = 0; i < condim - 1; i++) {
            const int adr = adr0 + 2 * i;
            const float d1 = adr < njmax ? force[adr] : 0.f, d2 = adr + 1 < njmax ? force[adr + 1] : 0.f;
            f[0] += d1 + d2;
            f[i + 1] = (d1 - d2) * K.d.contact_friction[5 * c + i];
          }
        }
      } else {  // elliptic: one row per dimension (:296-299)
        for (int i = 0; i < condim && i < nmp; i++) {
          const int adr = K.d.contact_efc_address[c * nmp + i];
          if (adr >= 0 && adr < njmax) f[i] = force[adr];
        }
      }
    }
    if (to_world_frame) {  // row vector times the contact frame (:301-305)
      const float* fr = K.d.contact_frame + 9 * c;
      float g[6];
      for (int h = 0; h < 2; h++)
        for (int j = 0; j < 3; j++) g[3 * h + j] = f[3 * h] * fr[j] + f[3 * h + 1] * fr[3 + j] + f[3 * h + 2] * fr[6 + j];
      for (int i = 0; i < 6; i++) f[i] = g[i];
    }
  }
  for (int i = 0; i < 6; i++) out[6 * t + i] = f[i];
}

extern "C" int mjb_contact_force(const MjbModel* m, MjbData* d, int cone, const int* contact_ids, int n, int to_world_frame, float* force,
                                 void* s) {
  KParams K;
  if (int rc = mjb_internal_params(m, d, &K)) return rc;
  if (n <= 0) return 0;
  if (!contact_ids || !force) return mjb_internal_fail(-2, "contact_ids / force is null");
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  mjb_contact_force_kernel<<<(n + 127) / 128, 128, 0, stream>>>(K, cone, contact_ids, n, to_world_frame, force);
  return launched();
}
