// mjb_g_eulw.cu -- the one-warp-per-world build of the `euler / implicit` stage-group kernel (variant 1: models with nv > 32)
#define MJB_UNIT_GROUP 5
#include "mjb_groups.h"
#define MJB_LW 32
#include "mjb_kernels.cuh"

GroupKernel mjb_gk_eul_wide(bool in_smem) { return in_smem ? mjb_group_kernel<G_EUL, true, 1> : mjb_group_kernel<G_EUL, false, 1>; }
