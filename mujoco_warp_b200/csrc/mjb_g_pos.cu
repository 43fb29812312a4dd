// mjb_g_pos.cu -- instantiations of the `smooth_position` stage-group kernel (see mjb_groups.h)
#define MJB_UNIT_GROUP 0
#include "mjb_groups.h"
#define MJB_LW MJB_POS_LW
#include "mjb_kernels.cuh"

GroupKernel mjb_gk_pos(bool in_smem) { return in_smem ? mjb_group_kernel<G_POS, true, 0> : mjb_group_kernel<G_POS, false, 0>; }
