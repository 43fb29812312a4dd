// mjb_g_con.cu -- instantiations of the `make_constraint` stage-group kernel (see mjb_groups.h)
#define MJB_UNIT_GROUP 2
#include "mjb_groups.h"
#include "mjb_kernels.cuh"

GroupKernel mjb_gk_con(bool in_smem, int variant) {
  if (variant) return in_smem ? mjb_group_kernel<G_CON, true, 1> : mjb_group_kernel<G_CON, false, 1>;
  return in_smem ? mjb_group_kernel<G_CON, true, 0> : mjb_group_kernel<G_CON, false, 0>;
}
