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

// support.xfrc_accumulate (support.py:229 -> apply_ft :175-216): qfrc += J^T xfrc_applied.  Thread = (world, dof); a dof collects the
// wrenches of the bodies in its body's subtree (the bodies its motion moves), about the subtree's root centre of mass like cdof.
__global__ void mjb_xfrc_accumulate_kernel(const __grid_constant__ KParams K) {
  const unsigned char* mb = K.blob;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int nv = K.ms.nv;
  if (t >= K.dsz.nworld * nv) return;
  const int w = t / nv, i = t - w * nv;
  const float* xfrc = DW(xfrc_applied);
  const S6 cd = ld6(DW(cdof) + 6 * i);
  const int db = MI(dof_bodyid)[i], end = MI(body_subtreeend)[db];
  float q = 0.f;
  for (int b = db; b < end; b++) {
    const S6 ft = ld6(xfrc + 6 * b);
    const V3 off = ld3(DW(xipos) + 3 * b) - ld3(DW(subtree_com) + 3 * MI(body_rootid)[b]);
    q += dot(bot(cd), top(ft)) + dot(top(cd), bot(ft)) + dot(cross(top(cd), off), top(ft));
  }
  K.x[(size_t)w * nv + i] += q;
}

// derivative.deriv_smooth_vel dense (derivative.py:321-420): out = qM - dt qDeriv with qDeriv = moment^T diag(d force / d velocity) moment
// - diag(damping) (actuators unless DisableBit.ACTUATION, damping unless DisableBit.DAMPER).  Thread = (world, i, j); K.n = rows of out.
__global__ void mjb_deriv_smooth_vel_kernel(const __grid_constant__ KParams K, int ld_out) {
  const unsigned char* mb = K.blob;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int nv = K.ms.nv, np = K.ms.nv_pad;
  if (t >= K.dsz.nworld * nv * nv) return;
  const int w = t / (nv * nv), e = t - w * nv * nv, i = e / nv, j = e - i * nv;
  float q = 0.f;
  if (K.ms.nu > 0 && !(K.o.disableflags & DSBL_ACTUATION)) {
    for (int u = 0; u < K.ms.nu; u++) {
      const float vel = act_vel_deriv(K, mb, w, u);
      if (vel == 0.f) continue;
      const int adr = DW(moment_rowadr)[u], nnz = DW(moment_rownnz)[u];
      float mi = 0.f, mj = 0.f;
      for (int k = 0; k < nnz; k++) {
        const int c = DW(moment_colind)[adr + k];
        const float v = DW(actuator_moment)[adr + k];
        if (c == i) mi = v;
        if (c == j) mj = v;
      }
      q += mi * mj * vel;
    }
  }
  if (!(K.o.disableflags & DSBL_DAMPER) && i == j) q -= MR(dof_damping)[i];
  K.x[((size_t)w * ld_out + i) * ld_out + j] = DW(qM)[(size_t)i * np + j] - K.o.timestep * q;
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

void mjb_util_xfrc_accumulate(const KParams& K, int grid, cudaStream_t stream) { mjb_xfrc_accumulate_kernel<<<grid, 256, 0, stream>>>(K); }
void mjb_util_deriv_smooth_vel(const KParams& K, int grid, int ld_out, cudaStream_t stream) { mjb_deriv_smooth_vel_kernel<<<grid, 256, 0, stream>>>(K, ld_out); }
