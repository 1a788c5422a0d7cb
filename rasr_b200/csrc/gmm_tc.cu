// placeholder until the tcgen05 kernel lands (next commit): the model never "fits", AUTO selects direct
#include "gmm_internal.cuh"
int gmm_tc_pack(gmm_handle* h, const uint32_t*, const float*, const float*, const uint32_t*, const float*, const float*) {
    h->tensor.valid = false;
    return GMM_OK;
}
int gmm_tc_score(gmm_handle* h, uint32_t, const float*, float*, uint32_t*, size_t) {
    h->setError("tensor kernel not built");
    return GMM_ERR_UNSUPPORTED;
}
