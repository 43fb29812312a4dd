// mjb_g_sol.cu -- instantiations of the `solve` stage-group kernel (see mjb_groups.h)
#define MJB_UNIT_GROUP 4
#include "mjb_groups.h"
#define MJB_CHOL_ENV 0  // the Hessian has no fixed envelope
#include "mjb_kernels.cuh"

GroupKernel mjb_gk_sol(bool in_smem, int variant) {
  if (variant) return in_smem ? mjb_group_kernel<G_SOL, true, 1> : mjb_group_kernel<G_SOL, false, 1>;
  return in_smem ? mjb_group_kernel<G_SOL, true, 0> : mjb_group_kernel<G_SOL, false, 0>;
}
