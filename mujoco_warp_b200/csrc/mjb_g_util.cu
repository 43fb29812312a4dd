// mjb_g_util.cu -- instantiations of the `utility (solve_m, mul_m, ctrl_noise)` stage-group kernel (see mjb_groups.h)
#include "mjb_groups.h"
#include "mjb_smooth.cuh"

// smooth.solve_m (smooth.py:2848): x = (L L^T)^-1 y with L = d.qLD
__global__ void __launch_bounds__(256) mjb_solve_m_kernel(const __grid_constant__ KParams K) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  float* scr = reinterpret_cast<float*>(smem) + (size_t)warp * K.scratch_floats;
  const int nv = K.ms.nv;
  const CholScr c = chol_scr(scr, nv, K.ms.nv_pad);
  const int ld = c.ld;
  float* x = c.x;
  for (int w = blockIdx.x * wpb + warp; w < K.dsz.nworld; w += gridDim.x * wpb) {
    load_rows(scr, ld, DW(qLD), nv, nv, lane, nullptr, 0.f);
    chol_rdiag(scr, ld, nv, c.rd, lane);
    for (int i = lane; i < nv; i += 32) x[i] = K.y[(size_t)w * nv + i];
    __syncwarp();
    warp_chol_solve<false>(scr, ld, nv, c.rd, x, lane);
    for (int i = lane; i < nv; i += 32) K.x[(size_t)w * nv + i] = x[i];
    __syncwarp();
  }
}

// support.mul_m dense (support.py:105-129)
__global__ void mjb_mul_m_kernel(const __grid_constant__ KParams K) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int nv = K.ms.nv, np = K.ms.nv_pad;
  if (t >= K.dsz.nworld * nv) return;
  const int w = t / nv, i = t - w * nv;
  const float* row = DW(qM) + (size_t)i * np;
  const float* v = K.y + (size_t)w * nv;
  float s = 0.f;
  for (int j = 0; j < nv; j++) s += row[j] * v[j];
  K.x[(size_t)w * nv + i] = s;
}

// util_misc.halton (util_misc.py:60-73)
__device__ __forceinline__ float halton(int index, int base) {
  int n0 = index;
  const float b = (float)base;
  float f = 1.0f / b, hn = 0.f;
  while (n0 > 0) {
    const int n1 = n0 / base;
    const int r = n0 - n1 * base;
    hn += f * (float)r;
    f /= b;
    n0 = n1;
  }
  return hn;
}

// benchmark.ctrl_noise (benchmark.py:42-83)
__global__ void mjb_ctrl_noise_kernel(const __grid_constant__ KParams K, const float* ctrl_center, int step, float ctrlnoisestd,
                                      float ctrlnoiserate) {
  const unsigned char* mb = K.blob;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int nu = K.ms.nu;
  if (t >= K.dsz.nworld * nu) return;
  const int w = t / nu, a = t - w * nu;
  const float rate = expf(-K.o.timestep / ctrlnoiserate);
  const float scale = ctrlnoisestd * sqrtf(1.0f - rate * rate);
  float midpoint = 0.f, halfrange = 1.f;
  const float lo = MR(actuator_ctrlrange)[2 * a], hi = MR(actuator_ctrlrange)[2 * a + 1];
  const bool limited = MI(actuator_ctrllimited)[a] != 0;
  if (limited) {
    midpoint = 0.5f * (hi + lo);
    halfrange = 0.5f * (hi - lo);
  }
  if (ctrl_center != nullptr) midpoint = ctrl_center[a];
  float ctrl = rate * DW(ctrl)[a] + (1.0f - rate) * midpoint;
  ctrl += scale * halfrange * (2.0f * halton((step + 1) * (w + 1), a + 2) - 1.0f);
  if (limited) ctrl = clampf(ctrl, lo, hi);
  DW(ctrl)[a] = ctrl;
}

cudaError_t mjb_util_solve_m_smem(size_t smem) {
  return cudaFuncSetAttribute(mjb_solve_m_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}
void mjb_util_solve_m(const KParams& K, int grid, int threads, size_t smem, cudaStream_t stream) {
  mjb_solve_m_kernel<<<grid, threads, smem, stream>>>(K);
}
void mjb_util_mul_m(const KParams& K, int grid, cudaStream_t stream) { mjb_mul_m_kernel<<<grid, 256, 0, stream>>>(K); }
void mjb_util_ctrl_noise(const KParams& K, int grid, const float* ctrl_center, int step, float ctrlnoisestd, float ctrlnoiserate,
                         cudaStream_t stream) {
  mjb_ctrl_noise_kernel<<<grid, 256, 0, stream>>>(K, ctrl_center, step, ctrlnoisestd, ctrlnoiserate);
}

__global__ void mjb_clear_kernel(int* a, int* b, int* q, int nq) {
  const int t = threadIdx.x;
  if (t == 0 && a) *a = 0;
  if (t == 1 && b) *b = 0;
  if (q && t < nq) q[t] = 0;
}
void mjb_util_clear(int* a, int* b, int* q, int nq, cudaStream_t stream) { mjb_clear_kernel<<<1, 32, 0, stream>>>(a, b, q, nq); }
