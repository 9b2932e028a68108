// Instantiations of dual_kernel<K, TDADD> (see dual.cuh).
#include "dual.cuh"
namespace pyopal {
search_kernel_t get_dual_kernel(int K, bool tdadd) {
#define DK(K) (tdadd ? (search_kernel_t)dual_kernel<K, true> : (search_kernel_t)dual_kernel<K, false>)
    switch (K) {
        case 4: return DK(4);
        case 8: return DK(8);
        case 12: return DK(12);
        case 16: return DK(16);
        case 20: return DK(20);
        case 24: return DK(24);
        case 28: return DK(28);
        case 32: return DK(32);
        case 36: return DK(36);
        case 40: return DK(40);
        default: return nullptr;
    }
#undef DK
}
}  // namespace pyopal
