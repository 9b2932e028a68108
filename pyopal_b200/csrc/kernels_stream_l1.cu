// Instantiations of stream_kernel<K, LOCAL = true, TDADD> (see stream.cuh).
#include "stream.cuh"
namespace pyopal {
search_kernel_t get_stream_kernel_l1(int K, bool tdadd) {
#define SK(K, TD) stream_kernel<K, true, TD>
    if (tdadd) switch (K) {
        case 4: return SK(4, true);
        case 8: return SK(8, true);
        case 12: return SK(12, true);
        case 16: return SK(16, true);
        case 20: return SK(20, true);
        case 24: return SK(24, true);
        case 28: return SK(28, true);
        case 32: return SK(32, true);
        case 36: return SK(36, true);
        case 40: return SK(40, true);
        default: return nullptr;
    }
    switch (K) {
        case 4: return SK(4, false);
        case 8: return SK(8, false);
        case 12: return SK(12, false);
        case 16: return SK(16, false);
        case 20: return SK(20, false);
        case 24: return SK(24, false);
        case 28: return SK(28, false);
        case 32: return SK(32, false);
        case 36: return SK(36, false);
        case 40: return SK(40, false);
        default: return nullptr;
    }
#undef SK
}
}  // namespace pyopal
