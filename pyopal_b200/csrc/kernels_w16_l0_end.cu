// Instantiations of search_kernel<K, 16, false, *, END = true> (see kernels.cuh): NW / HW / OV score + end cell.
#include "kernels.cuh"
namespace pyopal {
#define KEND(K, PS, TD) search_kernel<K, 16, false, PS, false, TD, false, true>
search_kernel_t get_search_kernel_w16_l0_end(int K, bool psmem, bool tdadd) {
    if (!psmem) {
        if (K != 32) return nullptr;
        return tdadd ? KEND(32, false, true) : KEND(32, false, false);
    }
    if (tdadd) switch (K) {
        case 4: return KEND(4, true, true);
        case 8: return KEND(8, true, true);
        case 12: return KEND(12, true, true);
        case 16: return KEND(16, true, true);
        case 20: return KEND(20, true, true);
        case 24: return KEND(24, true, true);
        case 28: return KEND(28, true, true);
        case 32: return KEND(32, true, true);
        case 36: return KEND(36, true, true);
        case 40: return KEND(40, true, true);
        default: return nullptr;
    }
    switch (K) {
        case 4: return KEND(4, true, false);
        case 8: return KEND(8, true, false);
        case 12: return KEND(12, true, false);
        case 16: return KEND(16, true, false);
        case 20: return KEND(20, true, false);
        case 24: return KEND(24, true, false);
        case 28: return KEND(28, true, false);
        case 32: return KEND(32, true, false);
        case 36: return KEND(36, true, false);
        case 40: return KEND(40, true, false);
        default: return nullptr;
    }
}
#undef KEND
}  // namespace pyopal
