// kernel_choice.h -- what the host side needs to know about one instantiation of the search kernels.
// The instantiations live in search_inst.cu, compiled once per (window pitch, block size) so that the nine
// translation units build in parallel; me_api.cu only sees these function pointers.
#pragma once

namespace me {

struct KernelChoice {
    const void *fn = nullptr;              // one CTA per tile, one block row per tile
    const void *fn_multirow = nullptr;     // the same kernel for tiles of several block rows
    const void *fn_persistent = nullptr;   // warp-specialised resident form
    const void *fn_scan = nullptr;         // the anchored walk as one cooperative kernel (32-bit keys, one-row tiles only)
    int ndy = 0, nq = 0, maxt = 0, pitch = 0, rank_bits = 0, persist_threads = 0;
};

constexpr int kNDY = 5, kNQ = 4, kMaxThreads = 256;
constexpr int kPitches[3] = {144, 208, 240};   // window row pitches built: odd multiples of 16 (bank-conflict free with NDY = 5)

#define ME_DECLARE_CHOICE(P, B) KernelChoice kernel_choice_##P##_##B(bool k64);
ME_DECLARE_CHOICE(144, 8) ME_DECLARE_CHOICE(144, 16) ME_DECLARE_CHOICE(144, 32)
ME_DECLARE_CHOICE(208, 8) ME_DECLARE_CHOICE(208, 16) ME_DECLARE_CHOICE(208, 32)
ME_DECLARE_CHOICE(240, 8) ME_DECLARE_CHOICE(240, 16) ME_DECLARE_CHOICE(240, 32)
#undef ME_DECLARE_CHOICE

}  // namespace me
