// Instantiations of search_kernel<K, 16, true, *> (see kernels.cuh).
#include "kernels.cuh"
namespace pyopal {
search_kernel_t get_search_kernel_w16_l1(int K, bool psmem) {
    if (!psmem) return (K == 32) ? search_kernel<32, 16, true, false> : nullptr;
    switch (K) {
        case 4: return search_kernel<4, 16, true, true>;
        case 8: return search_kernel<8, 16, true, true>;
        case 12: return search_kernel<12, 16, true, true>;
        case 16: return search_kernel<16, 16, true, true>;
        case 20: return search_kernel<20, 16, true, true>;
        case 24: return search_kernel<24, 16, true, true>;
        case 28: return search_kernel<28, 16, true, true>;
        case 32: return search_kernel<32, 16, true, true>;
        case 36: return search_kernel<36, 16, true, true>;
        case 40: return search_kernel<40, 16, true, true>;
        default: return nullptr;
    }
}
}  // namespace pyopal
