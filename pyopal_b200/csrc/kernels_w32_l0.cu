// Instantiations of search_kernel<K, 32, false, *> (see kernels.cuh).
#include "kernels.cuh"
namespace pyopal {
search_kernel_t get_search_kernel_w32_l0(int K, bool psmem) {
    if (!psmem) return (K == 32) ? search_kernel<32, 32, false, false> : nullptr;
    switch (K) {
        case 4: return search_kernel<4, 32, false, true>;
        case 8: return search_kernel<8, 32, false, true>;
        case 12: return search_kernel<12, 32, false, true>;
        case 16: return search_kernel<16, 32, false, true>;
        case 20: return search_kernel<20, 32, false, true>;
        case 24: return search_kernel<24, 32, false, true>;
        case 28: return search_kernel<28, 32, false, true>;
        case 32: return search_kernel<32, 32, false, true>;
        case 36: return search_kernel<36, 32, false, true>;
        case 40: return search_kernel<40, 32, false, true>;
        default: return nullptr;
    }
}
search_kernel_t get_search_kernel_bfix(bool wide) {
    return wide ? search_kernel<32, 32, false, false, true, true, true> : search_kernel<32, 32, false, false, false, true, true>;
}
}  // namespace pyopal
