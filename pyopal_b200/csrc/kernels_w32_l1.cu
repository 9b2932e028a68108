// Instantiations of search_kernel<K, 32, true, *> (see kernels.cuh).
#include "kernels.cuh"
namespace pyopal {
search_kernel_t get_search_kernel_w32_l1(int K, bool psmem) {
    if (!psmem) return (K == 32) ? search_kernel<32, 32, true, false> : nullptr;
    switch (K) {
        case 4: return search_kernel<4, 32, true, true>;
        case 8: return search_kernel<8, 32, true, true>;
        case 12: return search_kernel<12, 32, true, true>;
        case 16: return search_kernel<16, 32, true, true>;
        case 20: return search_kernel<20, 32, true, true>;
        case 24: return search_kernel<24, 32, true, true>;
        case 28: return search_kernel<28, 32, true, true>;
        case 32: return search_kernel<32, 32, true, true>;
        case 36: return search_kernel<36, 32, true, true>;
        case 40: return search_kernel<40, 32, true, true>;
        default: return nullptr;
    }
}
}  // namespace pyopal

namespace pyopal {
__global__ void build_profile_kernel(const KParams P, int8_t* narrow, int32_t* wide) {
    const int total = (P.A + 1) * P.pitch;
    for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < total; x += gridDim.x * blockDim.x) {
        int letter = x / P.pitch, trow = x - letter * P.pitch;
        int v = profile_entry(P, letter, trow);
        if (narrow) narrow[x] = (int8_t)v;
        if (wide) wide[x] = v;
    }
}
search_kernel_t get_search_kernel_wide(bool local) {
    return local ? search_kernel<32, 32, true, false, true> : search_kernel<32, 32, false, false, true>;
}
}  // namespace pyopal
