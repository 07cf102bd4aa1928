// search_inst.cu -- the search kernels of one (window pitch, block size): compiled nine times with
// -DME_INST_PITCH=144|208|240 -DME_INST_B=8|16|32 (see _lib.py build()).  32- and 64-bit keys, one-row and
// multi-row tiles and the resident form are all instantiated here; me_api.cu picks by function pointer.
#include "sad_search.cuh"
#include "sad_persistent.cuh"
#include "sad_scan.cuh"
#include "kernel_choice.h"

#if !defined(ME_INST_PITCH) || !defined(ME_INST_B)
#error "compile with -DME_INST_PITCH=<144|208|240> -DME_INST_B=<8|16|32>"
#endif

namespace me {
namespace {

template <int B, int PITCH, bool K64, int RB>
KernelChoice choice()
{
    KernelChoice k;
    k.fn = (const void *)&sad_search_kernel<B, kNDY, kNQ, PITCH, K64, RB, kMaxThreads, 3, false>;
    k.fn_multirow = (const void *)&sad_search_kernel<B, kNDY, kNQ, PITCH, K64, RB, kMaxThreads, 3, true>;
    constexpr int CW = (B <= 16 && !K64) ? kConsumerWarps : kConsumerWarps32;
    k.fn_persistent = (const void *)&sad_search_persistent<B, kNDY, kNQ, PITCH, K64, RB, CW, kProducerWarps>;
    k.persist_threads = 32 * (kProducerWarps + CW);
    if constexpr (!K64) k.fn_scan = (const void *)&sad_scan_kernel<B, kNDY, kNQ, PITCH, RB, kMaxThreads, 3>;
    k.ndy = kNDY; k.nq = kNQ; k.maxt = kMaxThreads; k.pitch = PITCH; k.rank_bits = RB;
    return k;
}

}  // namespace

#define ME_CAT3(a, b, c) a##b##_##c
#define ME_NAME(P, B) ME_CAT3(kernel_choice_, P, B)

// 32-bit keys: (sad << RB) | rank with sad < 2^14 (8x8), 2^16 (16x16), 2^18 (32x32); 64-bit keys when the rank needs more.
KernelChoice ME_NAME(ME_INST_PITCH, ME_INST_B)(bool k64)
{
    constexpr int RB32 = ME_INST_B == 32 ? 14 : 16;
    return k64 ? choice<ME_INST_B, ME_INST_PITCH, true, 32>() : choice<ME_INST_B, ME_INST_PITCH, false, RB32>();
}

}  // namespace me
