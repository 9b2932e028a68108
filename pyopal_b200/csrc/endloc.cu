// pyopal_b200/csrc/endloc.cu -- end-location and alignment (full) modes.
#include "../../include/pyopal_cuda.h"
#include "kernels.cuh"
struct pyopal_cuda_db;
int pyopal_end_full_search(pyopal_cuda_db*, const unsigned char*, int, int, int, const int*, int, int, int,
                           int64_t, int64_t, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, unsigned char*,
                           const int64_t*, int32_t*) {
    return PYOPAL_CUDA_ERR_INTERNAL;
}
