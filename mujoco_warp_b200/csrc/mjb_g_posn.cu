// mjb_g_posn.cu -- the four-worlds-per-warp build of the `smooth_position` stage-group kernel (variant 1: small models, full GPU)
#define MJB_UNIT_GROUP 0
#include "mjb_groups.h"
#define MJB_LW 8
#define MJB_CHOL_ENV 0  // this build serves nv <= 16 only: one tree block
#include "mjb_kernels.cuh"

GroupKernel mjb_gk_pos_narrow(bool in_smem) { return in_smem ? mjb_group_kernel<G_POS, true, 1> : mjb_group_kernel<G_POS, false, 1>; }
