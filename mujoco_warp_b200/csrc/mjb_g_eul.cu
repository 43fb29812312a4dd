// mjb_g_eul.cu -- instantiations of the `euler / implicit` stage-group kernel (see mjb_groups.h)
#define MJB_UNIT_GROUP 5
#include "mjb_groups.h"
#define MJB_LW MJB_EUL_LW
#define MJB_CHOL_ENV 0  // this build serves nv <= 32 only: one tree block
#include "mjb_kernels.cuh"

GroupKernel mjb_gk_eul(bool in_smem) { return in_smem ? mjb_group_kernel<G_EUL, true, 0> : mjb_group_kernel<G_EUL, false, 0>; }
