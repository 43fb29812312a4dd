// mjb_groups.h -- the stage-group kernels live in their own translation units (mjb_g_*.cu) so that the library builds in parallel
// and an edit of one stage recompiles one unit.  Each accessor returns the kernel instantiation for (blob in shared memory?, variant).
#pragma once
#include <cuda_runtime.h>

#include "mjb_types.cuh"

typedef void (*GroupKernel)(const KParams);
GroupKernel mjb_gk_pos(bool in_smem);
GroupKernel mjb_gk_pos_narrow(bool in_smem);  // four worlds per warp (small models)
GroupKernel mjb_gk_col(bool in_smem, int variant);  // primitive-only variants: CollisionFeature bits 0..3
GroupKernel mjb_gk_col_convex(bool in_smem);        // all features (GJK / EPA, meshes)
GroupKernel mjb_gk_con(bool in_smem, int variant);  // variant 1: elliptic cones
GroupKernel mjb_gk_vel(bool in_smem);
GroupKernel mjb_gk_vel_wide(bool in_smem);  // one warp per world (nv > 32)
GroupKernel mjb_gk_sol(bool in_smem, int variant);  // variant 1: elliptic cones
GroupKernel mjb_gk_eul(bool in_smem);
GroupKernel mjb_gk_eul_wide(bool in_smem);
// small utility kernels (mjb_g_util.cu)
cudaError_t mjb_util_solve_m_smem(size_t smem);
void mjb_util_solve_m(const KParams& K, int grid, int threads, size_t smem, cudaStream_t stream);
void mjb_util_mul_m(const KParams& K, int grid, cudaStream_t stream);
void mjb_util_xfrc_accumulate(const KParams& K, int grid, cudaStream_t stream);
void mjb_util_deriv_smooth_vel(const KParams& K, int grid, int ld_out, cudaStream_t stream);
void mjb_util_ctrl_noise(const KParams& K, int grid, const float* ctrl_center, int step, float ctrlnoisestd, float ctrlnoiserate, cudaStream_t stream);
// zeroes *a, *b (either may be null) and q[0..nq) (q may be null): the per-call counters of mjb_run_stages
void mjb_util_clear(int* a, int* b, int* q, int nq, cudaStream_t stream);
