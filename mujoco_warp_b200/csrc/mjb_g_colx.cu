// mjb_g_colx.cu -- instantiations of the `collision (all features: GJK / EPA convex pairs, meshes)` stage-group kernel (see mjb_groups.h)
#define MJB_UNIT_GROUP 1
#include "mjb_groups.h"
#include "mjb_kernels.cuh"

GroupKernel mjb_gk_col_convex(bool in_smem) {
  return in_smem ? mjb_group_kernel<G_COL, true, COLF_COUNT - 1> : mjb_group_kernel<G_COL, false, COLF_COUNT - 1>;
}
