// Instantiations of search_kernel<K, 16, true, *> (see kernels.cuh).
#include "kernels.cuh"
namespace pyopal {
search_kernel_t get_search_kernel_w16_l1(int K, bool psmem, bool tdadd) {
    if (!psmem) {
        if (K != 32) return nullptr;
        return tdadd ? search_kernel<32, 16, true, false, false, true> : search_kernel<32, 16, true, false, false, false>;
    }
    if (tdadd) switch (K) {
        case 4: return search_kernel<4, 16, true, true, false, true>;
        case 8: return search_kernel<8, 16, true, true, false, true>;
        case 12: return search_kernel<12, 16, true, true, false, true>;
        case 16: return search_kernel<16, 16, true, true, false, true>;
        case 20: return search_kernel<20, 16, true, true, false, true>;
        case 24: return search_kernel<24, 16, true, true, false, true>;
        case 28: return search_kernel<28, 16, true, true, false, true>;
        case 32: return search_kernel<32, 16, true, true, false, true>;
        case 36: return search_kernel<36, 16, true, true, false, true>;
        case 40: return search_kernel<40, 16, true, true, false, true>;
        default: return nullptr;
    }
    switch (K) {
        case 4: return search_kernel<4, 16, true, true, false, false>;
        case 8: return search_kernel<8, 16, true, true, false, false>;
        case 12: return search_kernel<12, 16, true, true, false, false>;
        case 16: return search_kernel<16, 16, true, true, false, false>;
        case 20: return search_kernel<20, 16, true, true, false, false>;
        case 24: return search_kernel<24, 16, true, true, false, false>;
        case 28: return search_kernel<28, 16, true, true, false, false>;
        case 32: return search_kernel<32, 16, true, true, false, false>;
        case 36: return search_kernel<36, 16, true, true, false, false>;
        case 40: return search_kernel<40, 16, true, true, false, false>;
        default: return nullptr;
    }
}
}  // namespace pyopal
