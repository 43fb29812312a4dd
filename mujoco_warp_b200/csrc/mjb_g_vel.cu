// mjb_g_vel.cu -- instantiations of the `smooth_velocity` stage-group kernel (see mjb_groups.h): several worlds per warp
#define MJB_UNIT_GROUP 3
#include "mjb_groups.h"
#define MJB_LW MJB_VEL_LW
#define MJB_CHOL_ENV 0  // this build serves nv <= 32 only: one tree block
#include "mjb_kernels.cuh"

GroupKernel mjb_gk_vel(bool in_smem) { return in_smem ? mjb_group_kernel<G_VEL, true, 0> : mjb_group_kernel<G_VEL, false, 0>; }
