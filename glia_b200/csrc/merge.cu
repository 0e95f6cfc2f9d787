// K5 / K8 entry points (implemented in a later milestone of this round).
#include "handle.cuh"

extern "C" int glia_b200_merge_order_bc(glia_handle* h, const glia_rf* rf, uint32_t* order,
                                        double* sal, double* rows, int64_t* n_merges,
                                        int64_t* n_featurized) {
  (void)rf; (void)order; (void)sal; (void)rows; (void)n_merges; (void)n_featurized;
  if (!h) return GLIA_E_ARG;
  h->err = "merge_order_bc: not implemented yet";
  return GLIA_E_STATE;
}

extern "C" int glia_b200_bc_feat(glia_handle* h, const uint32_t* order, int64_t n,
                                 const double* sal_in, double init_sal, double sal_bias,
                                 double* rows) {
  (void)order; (void)n; (void)sal_in; (void)init_sal; (void)sal_bias; (void)rows;
  if (!h) return GLIA_E_ARG;
  h->err = "bc_feat: not implemented yet";
  return GLIA_E_STATE;
}
