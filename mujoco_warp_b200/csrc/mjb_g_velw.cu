// mjb_g_velw.cu -- the one-warp-per-world build of the `smooth_velocity` stage-group kernel (variant 1: models with nv > 32)
#define MJB_UNIT_GROUP 3
#include "mjb_groups.h"
#define MJB_LW 32
#include "mjb_kernels.cuh"

GroupKernel mjb_gk_vel_wide(bool in_smem) { return in_smem ? mjb_group_kernel<G_VEL, true, 1> : mjb_group_kernel<G_VEL, false, 1>; }
