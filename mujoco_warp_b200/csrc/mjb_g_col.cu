// mjb_g_col.cu -- instantiations of the `collision (primitive pair types)` stage-group kernel (see mjb_groups.h)
#define MJB_UNIT_GROUP 1
#include "mjb_groups.h"
#include "mjb_kernels.cuh"

template <int V>
static GroupKernel pick(bool in_smem) {
  return in_smem ? mjb_group_kernel<G_COL, true, V> : mjb_group_kernel<G_COL, false, V>;
}
GroupKernel mjb_gk_col(bool in_smem, int variant) {
  switch (variant) {
    case 1: return pick<1>(in_smem);
    case 2: return pick<2>(in_smem);
    case 3: return pick<3>(in_smem);
    default: return pick<0>(in_smem);
  }
}
