// sad_search.cuh -- device side of the displacement search (sm_100a only).
//
// One CTA = one tile = `tmb` horizontally adjacent macro-blocks of one block row of the previous
// frame and the (x_off + tmb*MB + 2R) x (B + 2R) search window around them in the current frame.
// A macro-block is the 16- or 32-pixel-wide column unit a thread patch works on: one 16x16 or
// 32x32 block, or two adjacent 8x8 blocks (which then share every loaded word).
//
//   staging   One TMA box brings the window into shared memory (x origin rounded down to 16 bytes:
//             on this hardware a tiled TMA load faults unless its innermost coordinate is 16-byte
//             aligned, measured with tools/probe/tma_probe.cu).  The CTA then derives FOUR copies
//             of it, the s-th shifted s bytes to the left (s = 0..3, one SHF per word, <1 % of the
//             SAD work).  Candidate dx = 4q + s - R then reads only ALIGNED 32-bit words of copy s,
//             so the inner loop needs no byte realignment on the pipe the SAD op uses.
//             Out-of-frame bytes are zero-filled by TMA and the candidates that touch them are
//             masked at argmin time (spec: in-frame domain only).
//   inner     Each thread owns an NDY x NQ patch of candidates (same s, consecutive q, consecutive
//             dy) and walks the window rows once; a current-frame word loaded into a register is
//             used for up to NDY*min(NQ,BW) VABSDIFF4.U8.ACC ops, a reference word for NDY*NQ.
//             The window pitch is a template constant, so every shared load is base + immediate.
//             Columns that do not fill an NQ-wide patch run as NDY x 1 patches afterwards.
//   argmin    key = (sad << RB) | rank, rank = (dy+R)*(2R+1) + (dx+R): the minimum key is the
//             first minimum in raster order whatever the reduction order (spec tie-break).
//   finalize  eight lanes per block recompute the four neighbour SADs from global memory, apply
//             the parabola fit and the gates and write the per-block record.
//   reduce    the last CTA of a pair (atomic ticket) radix-selects the per-axis medians.
//
// The spec this implements is written down in oracle/sad_oracle.c; reference call sites replaced:
// estimate_vel_vid.py:306-369 (calc_ECC_transform) and :522 (its caller).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/me_b200.h"

namespace me {

constexpr int kMaxKeys = 32;          // blocks per tile row
constexpr int kMaxTileKeys = 128;     // blocks per tile of the one-CTA-per-tile kernel (tmr rows of up to kMaxKeys)
constexpr int kMaxGroups = 96;        // column groups per block (wide + narrow)
constexpr int kRefPitch = 256;        // bytes per row of the reference strip in shared memory
constexpr unsigned kFull = 0xffffffffu;
constexpr int kCopySkew = 32;         // copy s starts s*32 bytes further into the banks (see run_item lane mapping)

// Compile-time geometry of a block size.
template <int B> struct Geo {
    static constexpr int MB = B < 16 ? 16 : B;      // macro-block width in pixels
    static constexpr int NSUB = MB / B;             // blocks per macro-block
    static constexpr int BW = MB / 4;               // words per macro-block row
    static constexpr int SW = B / 4;                // words per block row
};

struct SearchArgs {
    // frames in device memory: (slots, H, pitch) uint8
    const uint8_t *frames;
    long long frame_stride;
    int frame_pitch;
    int W, H;
    const int32_t *pairs;       // device, (prev, cur) x P
    // block grid
    int nbx, nby, nb;
    int R, U;                   // U = 2R + 1
    int Ry, Uy, dy0;            // vertical range Ry <= R: Uy = 2Ry + 1 rows of the candidate square are searched, from dy index dy0 = R - Ry
    int tmb, tiles_x, tiles_per_pair;   // macro-blocks per tile row
    int tmr;                    // block rows per tile (> 1 only for single-slab plans with little work per row: small ranges)
    unsigned magic_nj[2];       // fastdiv by the macro-blocks of a tile row, [last tile in x]
    int nmbx;                   // macro-blocks per frame row = ceil(nbx / NSUB)
    // shared-memory window: 4 shifted copies, each alloc_rows x PITCH bytes
    int copy_bytes;
    int box_rows, n_row_boxes;
    int x_off;                  // (X0 - R) - (16-byte aligned TMA x origin), 0..15
    int n_bufs;                 // 4 when x_off == 0 (TMA lands in copy 0), else 5 (raw + 4 copies)
    // candidate slabs: the (2R+1)^2 candidate square is cut into nsx x nsy rectangles, one CTA each, so
    // that a window never exceeds one TMA box / the shared-memory budget.  Slab (sx,sy) starts at
    // (u, dyi) = (sx*su, sy*sd); the last slab of each axis takes the remainder (it is the largest).
    int nsx, nsy, su, sd;
    // candidate tiling inside a slab: column groups [0, n_wide) are NQ wide, then n_narrow single columns.
    // Index 0 = regular slab, 1 = last slab in x.
    int n_wide[2], n_narrow[2];
    unsigned short group[2][kMaxGroups];   // (s << 12) | q0
    // fastdiv magics for the common case (a slab whose dy groups are all in the domain), [last slab in y][last slab in x]
    int full_nd[2];
    unsigned magic_nd[2], magic_wide[2][2], magic_narrow[2][2];
    unsigned magic_nw[2], magic_nn[2];   // fastdiv by n_wide / n_narrow, [last slab in x]
    // cross-slab argmin (only when nsx*nsy > 1): all-ones on entry, restored by the finalizing CTA
    unsigned long long *blk_key;   // [P][nb]
    unsigned int *tile_ticket;     // [P][tiles_per_pair] zero on entry, reset by the finalizing CTA
    // gates
    int min_range, min_curv, subpixel, min_valid;
    int use_tma;
    int repeat;                 // debug (ME_REPEAT): run the candidate phase this many times; results are unchanged
    // outputs
    int16_t *blk_mv;            // [P][nb][2]
    uint32_t *blk_sad;          // [P][nb]
    float *blk_sub;             // [P][nb][2]
    uint8_t *blk_flags;         // [P][nb]
    me_pair_result *pair_out;   // [P]
    unsigned int *pair_ticket;  // [P] scratch: zero on entry, reset to zero by the last CTA
};

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t vsad4(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;   // SASS: VABSDIFF4.U8.ACC
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_inval(uint64_t *bar)
{
    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t done, addr = smem_u32(bar);
    do {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    } while (!done);
}
// 3-D tiled TMA load (x, y, frame) -> shared; completes on the mbarrier.  SASS: UTMALDG.3D
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *tm, int x, int y, int z, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
                 " [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar)) : "memory");
}

// n / d for n*d < 2^32 by one multiply-high (FMA pipe, not the ALU the SAD op needs); magic 0 means d == 1.
__device__ __forceinline__ unsigned fastdiv_magic(unsigned d) { return d <= 1 ? 0u : (unsigned)(0x100000000ull / d + 1ull); }
__device__ __forceinline__ int fastdiv(int n, unsigned magic) { return magic ? (int)__umulhi((unsigned)n, magic) : n; }

template <bool K64> struct KeyOf { using T = unsigned int; };
template <> struct KeyOf<true> { using T = unsigned long long; };

__device__ __forceinline__ unsigned int warp_min_key(unsigned int v) { return __reduce_min_sync(kFull, v); }
__device__ __forceinline__ unsigned long long warp_min_key(unsigned long long v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        unsigned long long t = __shfl_xor_sync(kFull, v, o);
        v = t < v ? t : v;
    }
    return v;
}

// ---------------------------------------------------------------- inner loop
// acc[sub][d][q] += SAD of block `sub` of the macro-block for dy index dyi0 + d, column q0 + q.
//   cur : word pointer to copy s, window row dyi0, word j*BW + q0
//   ref : word pointer to the reference strip, row 0, word j*BW
// LW = words per shared load of the current-frame row (4 needs j*BW + q0 to be a multiple of 4).
template <int B, int ROWS, int NDY, int NQP, int PITCH, int LW>
__device__ __forceinline__ void sad_rows(const uint32_t *__restrict__ cur, const uint32_t *__restrict__ ref,
                                         uint32_t (&acc)[Geo<B>::NSUB][NDY][NQP])
{
    constexpr int BW = Geo<B>::BW, SW = Geo<B>::SW;
    constexpr int PW = PITCH / 4, RPW = kRefPitch / 4;
    constexpr int NCW = NQP + BW - 1;               // current-frame words needed per window row
    constexpr int NLD = (NCW + LW - 1) / LW;
    uint32_t refw[NDY][BW];                         // rotating window of the last NDY reference rows
#pragma unroll
    for (int yy = 0; yy < ROWS + NDY - 1; ++yy) {
        uint32_t cw[NLD * LW];
        const uint32_t *crow = cur + yy * PW;
#pragma unroll
        for (int l = 0; l < NLD; ++l) {
            if constexpr (LW == 4) {
                const uint4 v = *reinterpret_cast<const uint4 *>(crow + 4 * l);
                cw[4 * l] = v.x; cw[4 * l + 1] = v.y; cw[4 * l + 2] = v.z; cw[4 * l + 3] = v.w;
            } else {
                cw[l] = crow[l];
            }
        }
        if (yy < ROWS) {
            const uint32_t *rrow = ref + yy * RPW;
#pragma unroll
            for (int l = 0; l < BW / 4; ++l) {
                const uint4 v = *reinterpret_cast<const uint4 *>(rrow + 4 * l);
                refw[yy % NDY][4 * l] = v.x; refw[yy % NDY][4 * l + 1] = v.y;
                refw[yy % NDY][4 * l + 2] = v.z; refw[yy % NDY][4 * l + 3] = v.w;
            }
        }
#pragma unroll
        for (int d = 0; d < NDY; ++d) {
            const int i = yy - d;                   // reference row paired with this window row for dy index d
            if (i >= 0 && i < ROWS) {
#pragma unroll
                for (int w = 0; w < BW; ++w)
#pragma unroll
                    for (int q = 0; q < NQP; ++q)
                        acc[w / SW][d][q] = vsad4(refw[i % NDY][w], cw[q + w], acc[w / SW][d][q]);
            }
        }
    }
}

// A 32-row block runs as four 8-row slices through the same (rolled) body of 1280 SAD ops, the size of the
// 16x16 patch: the fully unrolled 36-row form is 80 KB of straight-line code and starved the warps of
// instructions (r01e ncu: 1.9 no-instruction stalls per issue, still 1.0 with two 16-row halves); the price
// is 4 extra window rows of loads per slice.
template <int B, int NDY, int NQP, int PITCH, int LW, bool SLICE16 = false>
__device__ __forceinline__ void sad_patch(const uint32_t *__restrict__ cur, const uint32_t *__restrict__ ref,
                                          uint32_t (&acc)[Geo<B>::NSUB][NDY][NQP])
{
    if constexpr (B == 16 && SLICE16) {
        // multi-row tiles only (small ranges, short CTA phases): there the warps of an SM are spread over the whole
        // kernel and the 20 KB unrolled body misses the instruction cache (ncu: 2.7 no-instruction stalls per issue at
        // +-8); two rolled 8-row slices gave +13 % at +-8, +2.4 % at +-16.  4-row slices gained nothing more.
#pragma unroll 1
        for (int part = 0; part < 2; ++part)
            sad_rows<B, 8, NDY, NQP, PITCH, LW>(cur + part * 8 * (PITCH / 4), ref + part * 8 * (kRefPitch / 4), acc);
    } else if constexpr (B == 32) {
#pragma unroll 1
        for (int part = 0; part < 4; ++part)
            sad_rows<B, 8, NDY, NQP, PITCH, LW>(cur + part * 8 * (PITCH / 4), ref + part * 8 * (kRefPitch / 4), acc);
    } else {                                        // one-row tiles: 16 rows as two slices measured slower (0.854 vs 0.867)
        sad_rows<B, B, NDY, NQP, PITCH, LW>(cur, ref, acc);
    }
}

// ---------------------------------------------------------------- shifted window copies
// Derive the four byte-shifted copies of a staged window: copy s, byte k of a row = raw byte (x_off + s + k).
// Threads first, first + step, ... of the calling group take one 16-byte chunk each per round (the whole CTA in
// sad_search_kernel, the producer warps in sad_search_persistent).
template <int PITCH>
__device__ __forceinline__ void build_shifted_copies(const SearchArgs &a, const uint8_t *s_raw, uint8_t *s_win, int first, int step)
{
    const int chunks = a.n_row_boxes * a.box_rows * (PITCH >> 4);   // 16-byte chunks per copy
    const uint32_t *raw = reinterpret_cast<const uint32_t *>(s_raw);
    if (a.x_off == 0) {                                             // copy 0 is the TMA destination itself
        for (int i = first; i < chunks; i += step) {
            const uint4 v = *reinterpret_cast<const uint4 *>(raw + 4 * i);
            const uint32_t w4 = raw[4 * i + 4];                     // next chunk's first word (slack at the very end)
#pragma unroll
            for (int s = 1; s < 4; ++s) {
                uint4 o;
                o.x = __funnelshift_r(v.x, v.y, 8 * s);
                o.y = __funnelshift_r(v.y, v.z, 8 * s);
                o.z = __funnelshift_r(v.z, v.w, 8 * s);
                o.w = __funnelshift_r(v.w, w4, 8 * s);
                *reinterpret_cast<uint4 *>(s_win + (size_t)s * (a.copy_bytes + kCopySkew) + 16 * (size_t)i) = o;
            }
        }
    } else if ((a.x_off & 3) == 0) {
        // window origin a whole number of words off the aligned one (even ranges: +-8, +-24, +-40): copy s is the raw
        // row from word x_off/4 on, funnel-shifted by s bytes -- no per-word selects
        const uint32_t *rawo = raw + (a.x_off >> 2);
        for (int i = first; i < chunks; i += step) {
            const uint32_t e0 = rawo[4 * i], e1 = rawo[4 * i + 1], e2 = rawo[4 * i + 2], e3 = rawo[4 * i + 3], e4 = rawo[4 * i + 4];
            *reinterpret_cast<uint4 *>(s_win + 16 * (size_t)i) = make_uint4(e0, e1, e2, e3);
#pragma unroll
            for (int s = 1; s < 4; ++s) {
                uint4 o;
                o.x = __funnelshift_r(e0, e1, 8 * s);
                o.y = __funnelshift_r(e1, e2, 8 * s);
                o.z = __funnelshift_r(e2, e3, 8 * s);
                o.w = __funnelshift_r(e3, e4, 8 * s);
                *reinterpret_cast<uint4 *>(s_win + (size_t)s * (a.copy_bytes + kCopySkew) + 16 * (size_t)i) = o;
            }
        }
    } else {
        for (int i = first; i < chunks; i += step) {
            uint32_t w[9];
            const uint4 v0 = *reinterpret_cast<const uint4 *>(raw + 4 * i);
            const uint4 v1 = *reinterpret_cast<const uint4 *>(raw + 4 * i + 4);   // may run into the slack: unused bytes only
            w[0] = v0.x; w[1] = v0.y; w[2] = v0.z; w[3] = v0.w; w[4] = v1.x; w[5] = v1.y; w[6] = v1.z; w[7] = v1.w;
            w[8] = raw[4 * i + 8];
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int sh = a.x_off + s, wo = sh >> 2;
                const unsigned bits = (unsigned)(sh & 3) * 8;
                uint32_t e[5];                  // wo is 0..4: select statically to keep w[] in registers
#pragma unroll
                for (int k = 0; k < 5; ++k)
                    e[k] = wo == 0 ? w[k] : wo == 1 ? w[k + 1] : wo == 2 ? w[k + 2] : wo == 3 ? w[k + 3] : w[k + 4];
                uint4 o;
                o.x = __funnelshift_r(e[0], e[1], bits);
                o.y = __funnelshift_r(e[1], e[2], bits);
                o.z = __funnelshift_r(e[2], e[3], bits);
                o.w = __funnelshift_r(e[3], e[4], bits);
                *reinterpret_cast<uint4 *>(s_win + (size_t)s * (a.copy_bytes + kCopySkew) + 16 * (size_t)i) = o;
            }
        }
    }
}

// ---------------------------------------------------------------- finalize helpers
// 4 bytes from an arbitrary byte address, using aligned word loads only.  The second word is
// touched only when it holds a needed byte, so an in-frame request never leaves its row pitch.
__device__ __forceinline__ uint32_t ld_bytes4(const uint8_t *p)
{
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t *q = reinterpret_cast<const uint32_t *>(a & ~uintptr_t(3));
    const unsigned sh = (unsigned)(a & 3) * 8;
    uint32_t lo = __ldg(q);
    uint32_t hi = sh ? __ldg(q + 1) : 0u;
    return __funnelshift_r(lo, hi, sh);
}
__device__ __forceinline__ uint32_t hmin_bytes(uint32_t v)
{
    return min(min(v & 255u, (v >> 8) & 255u), min((v >> 16) & 255u, v >> 24));
}
__device__ __forceinline__ uint32_t hmax_bytes(uint32_t v)
{
    return max(max(v & 255u, (v >> 8) & 255u), max((v >> 16) & 255u, v >> 24));
}
__device__ __forceinline__ uint32_t fkey(float f)      // order-preserving float -> uint
{
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float funkey(uint32_t k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// Reduction over the block's kFinLanes lanes, gates, parabola, the 17-byte record (lane 0 writes).
__device__ __forceinline__ void finalize_record(const SearchArgs &a, int pair, int bx, int by, uint32_t sad, int dx, int dy,
                                                bool xin, bool yin, uint32_t sxm, uint32_t sxp, uint32_t sym, uint32_t syp,
                                                uint32_t vmn, uint32_t vmx, int lane, unsigned mask)
{
    sxm = __reduce_add_sync(mask, sxm);
    sxp = __reduce_add_sync(mask, sxp);
    sym = __reduce_add_sync(mask, sym);
    syp = __reduce_add_sync(mask, syp);
    const uint32_t mn = __reduce_min_sync(mask, hmin_bytes(vmn));
    const uint32_t mx = __reduce_max_sync(mask, hmax_bytes(vmx));
    if (lane == 0) {
        unsigned flags = 0;
        if ((int)(mx - mn) >= a.min_range) flags |= ME_FLAG_TEXTURED;
        float fx = (float)dx, fy = (float)dy;
        const int c0 = (int)sad, mc = a.min_curv < 1 ? 1 : a.min_curv;
        if (xin) {
            flags |= ME_FLAG_X_INTERIOR;
            const int den = (int)sxm - 2 * c0 + (int)sxp;
            if (den >= mc) flags |= ME_FLAG_X_PEAKED;
            if (den > 0 && a.subpixel)
                fx = __fadd_rn(fx, __fmul_rn(0.5f, __fdiv_rn((float)((int)sxm - (int)sxp), (float)den)));
        }
        if (yin) {
            flags |= ME_FLAG_Y_INTERIOR;
            const int den = (int)sym - 2 * c0 + (int)syp;
            if (den >= mc) flags |= ME_FLAG_Y_PEAKED;
            if (den > 0 && a.subpixel)
                fy = __fadd_rn(fy, __fmul_rn(0.5f, __fdiv_rn((float)((int)sym - (int)syp), (float)den)));
        }
        const size_t b = (size_t)pair * a.nb + (size_t)by * a.nbx + bx;
        reinterpret_cast<short2 *>(a.blk_mv)[b] = make_short2((short)dx, (short)dy);
        a.blk_sad[b] = sad;
        reinterpret_cast<float2 *>(a.blk_sub)[b] = make_float2(fx, fy);
        a.blk_flags[b] = (uint8_t)flags;
        // no fence here: every caller synchronises its CTA (or warp) after the finalize loop and ONE thread then fences before it
        // takes the pair's ticket -- fences are cumulative, so that one covers these stores (a fence per block was ~1 k cycles in
        // front of every ticket: r02 scan trace)
    }
}


// Eight lanes (one quarter of a warp, `mask` names them, `lane` = 0..7 inside the group): neighbour SADs,
// gates, parabola, per-block record.  Four blocks are finalized per warp pass.
constexpr int kFinLanes = 8;
template <int B>
__device__ __forceinline__ void finalize_block(const SearchArgs &a, int pair, const uint8_t *prevf,
                                               const uint8_t *curf, int bx, int by, uint32_t sad, uint32_t rank,
                                               int lane, unsigned mask)
{
    constexpr int SW = B / 4;
    const int R = a.R, U = a.U, x0 = bx * B, y0 = by * B;
    const int dyi = (int)(rank / (uint32_t)U), ui = (int)rank - dyi * U;
    const int dx = ui - R, dy = dyi - R;
    const int dxlo = max(-R, -x0), dxhi = min(R, a.W - B - x0);
    const int dylo = max(-a.Ry, -y0), dyhi = min(a.Ry, a.H - B - y0);
    const bool xin = dx - 1 >= dxlo && dx + 1 <= dxhi;
    const bool yin = dy - 1 >= dylo && dy + 1 <= dyhi;
    const size_t pitch = (size_t)a.frame_pitch;
    uint32_t sxm = 0, sxp = 0, sym = 0, syp = 0, vmn = 0xffffffffu, vmx = 0u;
    for (int idx = lane; idx < B * SW; idx += kFinLanes) {
        const int row = idx / SW, wd = idx - row * SW;
        const uint32_t rw = __ldg(reinterpret_cast<const uint32_t *>(prevf + (size_t)(y0 + row) * pitch + x0) + wd);
        vmn = __vminu4(vmn, rw);
        vmx = __vmaxu4(vmx, rw);
        const uint8_t *c = curf + (size_t)(y0 + dy + row) * pitch + (x0 + dx + 4 * wd);
        if (xin) {
            sxm = vsad4(rw, ld_bytes4(c - 1), sxm);
            sxp = vsad4(rw, ld_bytes4(c + 1), sxp);
        }
        if (yin) {
            sym = vsad4(rw, ld_bytes4(c - pitch), sym);
            syp = vsad4(rw, ld_bytes4(c + pitch), syp);
        }
    }
    finalize_record(a, pair, bx, by, sad, dx, dy, xin, yin, sxm, sxp, sym, syp, vmn, vmx, lane, mask);
}

// Same step for a CTA that still holds the block's whole candidate window in shared memory (single-slab tiles):
// the neighbour SADs come from the shifted copies with aligned word loads -- no global loads, no realignment,
// no 64-bit address arithmetic (r01f ncu at +-8: the global-memory form was ~15 % of all issued instructions
// and the source of the long-scoreboard stalls).  jr / jb: block row in the tile / block in the row.
template <int B, int PITCH, int FL = 8>
__device__ __forceinline__ void finalize_block_smem(const SearchArgs &a, int pair, const uint8_t *s_win, const uint8_t *s_ref,
                                                    int jr, int jb, int bx, int by, uint32_t sad, uint32_t rank,
                                                    int lane, unsigned mask)
{
    constexpr int SW = B / 4, PW = PITCH / 4, RPW = kRefPitch / 4;
    const int R = a.R, U = a.U, x0 = bx * B, y0 = by * B;
    const int dyi = (int)(rank / (uint32_t)U), ui = (int)rank - dyi * U;
    const int dx = ui - R, dy = dyi - R;
    const int dxlo = max(-R, -x0), dxhi = min(R, a.W - B - x0);
    const int dylo = max(-a.Ry, -y0), dyhi = min(a.Ry, a.H - B - y0);
    const bool xin = dx - 1 >= dxlo && dx + 1 <= dxhi;
    const bool yin = dy - 1 >= dylo && dy + 1 <= dyhi;
    const size_t cstride = (size_t)a.copy_bytes + kCopySkew;
    // word pointers of the candidate (u, dyi) = copy u & 3, row dyi - dy0 + jr*B (the window starts at dy index dy0), word jb*SW + (u >> 2)
    auto cand = [&](int u, int d) {
        return reinterpret_cast<const uint32_t *>(s_win + (size_t)(u & 3) * cstride) + (d - a.dy0 + jr * B) * PW + jb * SW + (u >> 2);
    };
    const uint32_t *ref = reinterpret_cast<const uint32_t *>(s_ref) + jr * B * RPW + jb * SW;
    const uint32_t *cxm = cand(xin ? ui - 1 : ui, dyi), *cxp = cand(xin ? ui + 1 : ui, dyi);
    const uint32_t *cym = cand(ui, yin ? dyi - 1 : dyi), *cyp = cand(ui, yin ? dyi + 1 : dyi);
    uint32_t sxm = 0, sxp = 0, sym = 0, syp = 0, vmn = 0xffffffffu, vmx = 0u;
#pragma unroll 2
    for (int idx = lane; idx < B * SW; idx += FL) {
        const int row = idx / SW, wd = idx - row * SW;
        const uint32_t rw = ref[row * RPW + wd];
        vmn = __vminu4(vmn, rw);
        vmx = __vmaxu4(vmx, rw);
        const int o = row * PW + wd;
        sxm = vsad4(rw, cxm[o], sxm);
        sxp = vsad4(rw, cxp[o], sxp);
        sym = vsad4(rw, cym[o], sym);
        syp = vsad4(rw, cyp[o], syp);
    }
    finalize_record(a, pair, bx, by, sad, dx, dy, xin, yin, sxm, sxp, sym, syp, vmn, vmx, lane, mask);
}


// ---------------------------------------------------------------- per-pair median (last CTA)
// The CTA that takes the pair's last ticket reduces it while the rest of the grid works on other
// pairs -- except for the last pair of a launch, where this code IS the tail of the launch (r01: 2.8 ms
// of a 4.7 ms 4K launch in its first form: 11 passes with one dependent load chain per thread).
// Now: an exact 3-level radix select (11 + 11 + 10 bits of the order-preserving float key), both axes
// in the same pass, eight independent block records in flight per thread.  The level-0 histogram
// also yields the counts, so the whole reduction is three streaming passes over the pair's records
// (plus one when an even count needs the next larger key).
constexpr uint32_t kNoKey = 0u;        // fkey() of a finite float is never 0
constexpr int kSelBins = 2048;         // 11 bits per level; 2 axes x 2048 x 4 B = 16 KB of the CTA's shared memory
constexpr int kSelUnroll = 8;

__device__ __forceinline__ void pair_keys(const uint8_t *fl, const float2 *sub, int b, uint32_t &kx, uint32_t &ky, unsigned &tex, float2 &v)
{
    const unsigned f = __ldcg(fl + b);
    tex = f & ME_FLAG_TEXTURED;
    v = __ldcg(sub + b);
    kx = (f & (ME_FLAG_TEXTURED | ME_FLAG_X_PEAKED)) == (ME_FLAG_TEXTURED | ME_FLAG_X_PEAKED) ? fkey(v.x) : kNoKey;
    ky = (f & (ME_FLAG_TEXTURED | ME_FLAG_Y_PEAKED)) == (ME_FLAG_TEXTURED | ME_FLAG_Y_PEAKED) ? fkey(v.y) : kNoKey;
}

// Few blocks per pair (the 640x64 recordings have 160): the three radix passes, each clearing and scanning 2 x 2048 bins,
// are a fixed ~10 us at the tail of every launch of the anchored scan, which waits for exactly that.  Here the order
// statistics come from counting: with lt(x) = #(valid keys < x), the m-th smallest key is max{x : lt(x) <= m} -- one compare
// per key pair, no tie-break, and the selection is an atomicMax.  A (axis, block) key is counted by four adjacent lanes, a
// quarter of the key list each (16-byte broadcast reads, four independent counters), joined by two shuffles.  The CTA that
// runs this is the only busy one on its SM, one warp per scheduler: every dependent instruction costs its full latency, so
// the instruction COUNT is what matters (r02 trace at 128 threads: first form, one thread per block and 14 instructions per
// key, 14.7 k cycles per pair; lt and le counted separately and joined in shared memory, 10.4 k cycles in the counting alone).
constexpr int kSmallNb = 512;
static __device__ void reduce_pair_small(const SearchArgs &a, int pair, unsigned *scratch, unsigned long long *sub_phase = nullptr)
{
    long long rc0 = sub_phase ? clock64() : 0;
    int rci = 0;
    auto rmark = [&]() {                         // trace only: cycles of [0] clear + key loads, [1] sums, [2] counting + selection, [3] record
        if (sub_phase && threadIdx.x == 0) { const long long t = clock64(); atomicAdd(sub_phase + rci++, (unsigned long long)(t - rc0)); rc0 = t; }
    };
    uint32_t *kx = scratch;                      // [2][kSmallNb] keys, x then y; invalid = kNoKey = 0, below every real key
    unsigned *bc = scratch + 2 * kSmallNb;       // [0],[1] lower / upper middle key x; [2],[3] y; [4] textured; [5],[6] counts; [8..11] field sums
    const uint8_t *fl = a.blk_flags + (size_t)pair * a.nb;
    const float2 *sub = reinterpret_cast<const float2 *>(a.blk_sub) + (size_t)pair * a.nb;
    const int T = blockDim.x, tid = threadIdx.x, nb = a.nb;
    const int nb16 = (nb + 15) & ~15;            // every quarter of the key list is whole 16-byte words: pad with kNoKey
    if (tid < 12) bc[tid] = 0;
    for (int i = nb + tid; i < nb16; i += T) kx[i] = kx[kSmallNb + i] = kNoKey;
    __syncthreads();
    unsigned ntex = 0, nx = 0, ny = 0;
    long long fsx = 0, fsy = 0;
    for (int b0 = tid; b0 < nb; b0 += 2 * T) {   // two records in flight per thread (each is two dependent-free L2 reads)
        uint32_t x[2], y[2];
        unsigned tex[2];
        float2 v[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            x[u] = y[u] = kNoKey; tex[u] = 0; v[u] = make_float2(0.f, 0.f);
            if (b0 + u * T < nb) pair_keys(fl, sub, b0 + u * T, x[u], y[u], tex[u], v[u]);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            if (b0 + u * T >= nb) continue;
            kx[b0 + u * T] = x[u]; kx[kSmallNb + b0 + u * T] = y[u];
            ntex += tex[u];
            if (x[u] != kNoKey) { ++nx; fsx += (long long)__float2ll_rn(__fmul_rn(v[u].x, 1048576.0f)); }
            if (y[u] != kNoKey) { ++ny; fsy += (long long)__float2ll_rn(__fmul_rn(v[u].y, 1048576.0f)); }
        }
    }
    rmark();
    // warp sums: the 64-bit field sums as a low 16-bit and a high part (|v| <= 128.5 and at most four records per lane keep
    // both inside 32 bits over a warp), one REDUX each instead of five 64-bit shuffle rounds
    ntex = __reduce_add_sync(kFull, ntex); nx = __reduce_add_sync(kFull, nx); ny = __reduce_add_sync(kFull, ny);
    const unsigned lox = __reduce_add_sync(kFull, (unsigned)(fsx & 0xffff)), loy = __reduce_add_sync(kFull, (unsigned)(fsy & 0xffff));
    const int hix = __reduce_add_sync(kFull, (int)(fsx >> 16)), hiy = __reduce_add_sync(kFull, (int)(fsy >> 16));
    if ((tid & 31) == 0) {
        const long long wx = (long long)hix * 65536 + (long long)lox, wy = (long long)hiy * 65536 + (long long)loy;
        if (ntex) atomicAdd(&bc[4], ntex);
        if (nx) atomicAdd(&bc[5], nx);
        if (ny) atomicAdd(&bc[6], ny);
        if (wx) atomicAdd(reinterpret_cast<unsigned long long *>(bc + 8), (unsigned long long)wx);
        if (wy) atomicAdd(reinterpret_cast<unsigned long long *>(bc + 10), (unsigned long long)wy);
    }
    __syncthreads();
    rmark();
    const unsigned cx = bc[5], cy = bc[6];
    const int per = nb16 / 16;                                                   // 16-byte words per quarter of a key list
    uint32_t best[4] = {0u, 0u, 0u, 0u};                                         // running max of the keys with lt <= m: [2 * axis + (upper middle)]
    // compare + predicated add, two instructions (the compiler's form of `c += kk < x` is three: compare, add, select)
    auto lt1 = [](unsigned &c, uint32_t kk, uint32_t x) {
        asm("{\n\t.reg .pred p;\n\tsetp.lt.u32 p, %1, %2;\n\t@p add.u32 %0, %0, 1;\n\t}" : "+r"(c) : "r"(kk), "r"(x));
    };
    constexpr int kRegWords = 10;                                                // 160 keys per axis: the recordings
    if (per <= kRegWords) {
        // A lane's role is fixed -- axis (tid >> 2) & 1, quarter tid & 3 -- so its quarter of the keys sits in registers for the whole
        // loop, and a group of eight lanes (both axes, four quarters) takes two blocks per round: a round is one key read per
        // block, the compares, and two shuffles; no shared-memory traffic in between (r02 trace: rounds that re-read the keys from
        // shared memory cost ~790 cycles each at one warp per scheduler, twice what their instruction count explains).
        const int axis = (tid >> 2) & 1, sq = tid & 3, grp = tid >> 3, ngrp = T >> 3;
        const uint32_t *k = kx + axis * kSmallNb;
        const uint4 *k4 = reinterpret_cast<const uint4 *>(k) + sq * per;
        uint4 kr[kRegWords];
#pragma unroll
        for (int u = 0; u < kRegWords; ++u) kr[u] = u < per ? k4[u] : make_uint4(~0u, ~0u, ~0u, ~0u);   // padding is never below a key
        const unsigned c = axis ? cy : cx, inval = (unsigned)nb16 - c, m1 = (c - 1) / 2, m2 = c / 2;
        uint32_t bl = 0u, bh = 0u;
        for (int base = 0; base < nb; base += 2 * ngrp) {                        // uniform trip count: the shuffles take whole warps
            const int ba = base + grp, bb = base + ngrp + grp;
            const uint32_t xa = ba < nb ? k[ba] : kNoKey, xb = bb < nb ? k[bb] : kNoKey;
            unsigned a0 = 0, a1 = 0, a2 = 0, a3 = 0, b0 = 0, b1 = 0, b2 = 0, b3 = 0;
#pragma unroll
            for (int u = 0; u < kRegWords; ++u) {
                lt1(a0, kr[u].x, xa); lt1(a1, kr[u].y, xa); lt1(a2, kr[u].z, xa); lt1(a3, kr[u].w, xa);
                lt1(b0, kr[u].x, xb); lt1(b1, kr[u].y, xb); lt1(b2, kr[u].z, xb); lt1(b3, kr[u].w, xb);
            }
            unsigned lta = (a0 + a1) + (a2 + a3), ltb = (b0 + b1) + (b2 + b3);
            lta += __shfl_xor_sync(kFull, lta, 1); ltb += __shfl_xor_sync(kFull, ltb, 1);
            lta += __shfl_xor_sync(kFull, lta, 2); ltb += __shfl_xor_sync(kFull, ltb, 2);
            if (sq == 0) {
                lta -= inval; ltb -= inval;                                      // every invalid (and padding-0) key counted as smaller
                if (xa != kNoKey) { if (lta <= m1) bl = max(bl, xa); if (lta <= m2) bh = max(bh, xa); }
                if (xb != kNoKey) { if (ltb <= m1) bl = max(bl, xb); if (ltb <= m2) bh = max(bh, xb); }
            }
        }
        best[0] = axis ? 0u : bl; best[1] = axis ? 0u : bh; best[2] = axis ? bl : 0u; best[3] = axis ? bh : 0u;
    } else {
        const int total = 8 * nb;                                                // tasks = (axis, block, quarter)
        for (int base = 0; base < total; base += T) {                            // uniform trip count: the shuffles below take whole warps
            const int t = base + tid, idx = t >> 2, sq = t & 3;
            const bool live = t < total;
            const int axis = live && idx >= nb, blk = live ? idx - axis * nb : 0;
            const uint32_t *k = kx + axis * kSmallNb;
            const uint32_t x = k[blk];
            const uint4 *k4 = reinterpret_cast<const uint4 *>(k) + sq * per;
            unsigned c0 = 0, c1 = 0, c2 = 0, c3 = 0;
            auto count = [&](const uint4 kk) { lt1(c0, kk.x, x); lt1(c1, kk.y, x); lt1(c2, kk.z, x); lt1(c3, kk.w, x); };
            int w = 0;
            for (; w + 8 <= per; w += 8) {                                       // eight loads in flight
                uint4 v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = k4[w + u];
#pragma unroll
                for (int u = 0; u < 8; ++u) count(v[u]);
            }
            for (; w < per; ++w) count(k4[w]);
            unsigned lt = (c0 + c1) + (c2 + c3);
            lt += __shfl_xor_sync(kFull, lt, 1);
            lt += __shfl_xor_sync(kFull, lt, 2);
            if (live && sq == 0 && x != kNoKey) {
                const unsigned c = axis ? cy : cx;
                lt -= (unsigned)nb16 - c;                                        // every invalid (and padding) key counted as smaller
                const uint32_t lo = lt <= (c - 1) / 2 ? x : 0u, hi = lt <= c / 2 ? x : 0u;
                if (axis) { best[2] = max(best[2], lo); best[3] = max(best[3], hi); }
                else      { best[0] = max(best[0], lo); best[1] = max(best[1], hi); }
            }
        }
    }
    // half of all keys pass lt <= m: one shared-memory atomic per warp and target, not one per key (r02 trace: 320 same-address
    // atomicMax made this loop 8.5 k cycles)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t m = __reduce_max_sync(kFull, best[i]);
        if ((tid & 31) == 0 && m) atomicMax(&bc[i], m);
    }
    __syncthreads();
    rmark();
    // the record: x fields by thread 0, y fields by the first thread of the next warp (two double divisions side by side)
    me_pair_result *out = a.pair_out + pair;
    const int ty = T >= 64 ? 32 : 0;
    if (tid == 0) {
        out->pdx = cx ? __fmul_rn(0.5f, __fadd_rn(funkey(bc[0]), funkey(bc[1]))) : 0.0f;
        out->flag = (int)cx >= (a.min_valid > 1 ? a.min_valid : 1);
        out->nvalid_x = (int)cx;
        out->n_textured = (int)bc[4];
        const long long sx = *reinterpret_cast<const long long *>(bc + 8);
        out->mean_dx = cx ? (float)__ddiv_rn((double)sx, __dmul_rn((double)cx, 1048576.0)) : 0.0f;
        a.pair_ticket[pair] = 0;                    // leave the scratch clean for the next launch
    }
    if (tid == ty) {
        out->pdy = cy ? __fmul_rn(0.5f, __fadd_rn(funkey(bc[2]), funkey(bc[3]))) : 0.0f;
        out->nvalid_y = (int)cy;
        const long long sy = *reinterpret_cast<const long long *>(bc + 10);
        out->mean_dy = cy ? (float)__ddiv_rn((double)sy, __dmul_rn((double)cy, 1048576.0)) : 0.0f;
    }
    rmark();
}

// scratch: the CTA's shared memory from offset 0 (at least 2*kSelBins*4 + 128 bytes)
static __device__ void reduce_pair(const SearchArgs &a, int pair, unsigned *scratch, unsigned long long *sub_phase = nullptr)
{
    if (a.nb <= kSmallNb) { reduce_pair_small(a, pair, scratch, sub_phase); return; }     // uniform across the CTA
    unsigned *hist = scratch;                    // [2][kSelBins]
    unsigned *bc = scratch + 2 * kSelBins;       // per axis: [0] chosen bin, [1] rank inside it, [2] its count, [3] total; [8],[9] next keys; [10] textured; [12..15] field sums
    const uint8_t *fl = a.blk_flags + (size_t)pair * a.nb;
    const float2 *sub = reinterpret_cast<const float2 *>(a.blk_sub) + (size_t)pair * a.nb;
    const int T = blockDim.x, tid = threadIdx.x;
    uint32_t prefix[2] = {0, 0}, mask[2] = {0, 0};
    unsigned k[2] = {0, 0}, cnt[2] = {0, 0}, neq[2] = {0, 0}, idx[2] = {0, 0};
    bool even[2] = {false, false};
    for (int level = 0; level < 3; ++level) {
        const int shift = level == 0 ? 21 : level == 1 ? 10 : 0;
        const uint32_t bmask = level == 2 ? 1023u : 2047u;
        for (int i = tid; i < 2 * kSelBins; i += T) hist[i] = 0;
        if (level == 0 && tid < 6) bc[10 + tid] = 0;          // [10] textured count, [12..15] the two 64-bit field sums
        __syncthreads();
        unsigned ntex = 0;
        long long fsx = 0, fsy = 0;                  // spec "mean": exact 2^-20 fixed-point sums of the valid displacements
        for (int b0 = tid; b0 < a.nb; b0 += kSelUnroll * T) {
            uint32_t kx[kSelUnroll], ky[kSelUnroll];
#pragma unroll
            for (int u = 0; u < kSelUnroll; ++u) {
                kx[u] = ky[u] = kNoKey;
                unsigned tex = 0;
                float2 v = make_float2(0.f, 0.f);
                if (b0 + u * T < a.nb) pair_keys(fl, sub, b0 + u * T, kx[u], ky[u], tex, v);
                ntex += tex;
                if (level == 0) {
                    if (kx[u] != kNoKey) fsx += (long long)__float2ll_rn(__fmul_rn(v.x, 1048576.0f));
                    if (ky[u] != kNoKey) fsy += (long long)__float2ll_rn(__fmul_rn(v.y, 1048576.0f));
                }
            }
#pragma unroll
            for (int u = 0; u < kSelUnroll; ++u) {
                if (kx[u] != kNoKey && (kx[u] & mask[0]) == prefix[0]) atomicAdd(&hist[(kx[u] >> shift) & bmask], 1u);
                if (ky[u] != kNoKey && (ky[u] & mask[1]) == prefix[1]) atomicAdd(&hist[kSelBins + ((ky[u] >> shift) & bmask)], 1u);
            }
        }
        if (level == 0) {                            // the first pass also counts the textured blocks and sums the field
            ntex = __reduce_add_sync(kFull, ntex);
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                fsx += __shfl_xor_sync(kFull, fsx, o);
                fsy += __shfl_xor_sync(kFull, fsy, o);
            }
            if ((tid & 31) == 0) {
                if (ntex) atomicAdd(&bc[10], ntex);
                if (fsx) atomicAdd(reinterpret_cast<unsigned long long *>(bc + 12), (unsigned long long)fsx);
                if (fsy) atomicAdd(reinterpret_cast<unsigned long long *>(bc + 14), (unsigned long long)fsy);
            }
        }
        __syncthreads();
        if (tid < 64) {                              // warp 0: x axis, warp 1: y axis
            const int axis = tid >> 5, lane = tid & 31;
            const unsigned *h = hist + axis * kSelBins;
            constexpr int per = kSelBins / 32;
            unsigned mine = 0;
            for (int i = 0; i < per; ++i) mine += h[lane * per + i];
            unsigned incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned t = __shfl_up_sync(kFull, incl, o);
                if (lane >= o) incl += t;
            }
            const unsigned total = __shfl_sync(kFull, incl, 31);
            unsigned kk = k[axis];
            if (level == 0) kk = total ? (total - 1) / 2 : 0;       // lower middle of this axis
            const unsigned excl = incl - mine;
            if (lane == 0) { bc[axis * 4 + 3] = total; if (total == 0) { bc[axis * 4] = 0; bc[axis * 4 + 1] = 0; bc[axis * 4 + 2] = 0; } }
            if (total && kk >= excl && kk < incl) {                 // exactly one lane
                unsigned cum = excl;
                int sel = lane * per;
                for (int i = 0; i < per; ++i) {
                    const unsigned c = h[lane * per + i];
                    if (kk < cum + c) { sel = lane * per + i; break; }
                    cum += c;
                }
                bc[axis * 4] = (unsigned)sel;
                bc[axis * 4 + 1] = kk - cum;
                bc[axis * 4 + 2] = h[sel];
            }
        }
        __syncthreads();
#pragma unroll
        for (int axis = 0; axis < 2; ++axis) {
            if (level == 0) { cnt[axis] = bc[axis * 4 + 3]; even[axis] = cnt[axis] && (cnt[axis] & 1u) == 0; }
            prefix[axis] |= bc[axis * 4] << shift;
            mask[axis] |= bmask << shift;
            k[axis] = bc[axis * 4 + 1];
            if (level == 2) { neq[axis] = bc[axis * 4 + 2]; idx[axis] = bc[axis * 4 + 1]; }
        }
        __syncthreads();
    }
    // upper middle: the same key unless the lower middle is the last of its run and the count is even
    uint32_t v2[2] = {prefix[0], prefix[1]};
    const bool need0 = even[0] && idx[0] + 1 >= neq[0], need1 = even[1] && idx[1] + 1 >= neq[1];
    if (need0 || need1) {                            // uniform across the CTA
        if (tid == 0) { bc[8] = 0xffffffffu; bc[9] = 0xffffffffu; }
        __syncthreads();
        uint32_t bx = 0xffffffffu, by = 0xffffffffu;
        for (int b0 = tid; b0 < a.nb; b0 += kSelUnroll * T) {
            uint32_t kx[kSelUnroll], ky[kSelUnroll];
#pragma unroll
            for (int u = 0; u < kSelUnroll; ++u) {
                kx[u] = ky[u] = kNoKey;
                unsigned tex;
                float2 v;
                if (b0 + u * T < a.nb) pair_keys(fl, sub, b0 + u * T, kx[u], ky[u], tex, v);
            }
#pragma unroll
            for (int u = 0; u < kSelUnroll; ++u) {
                if (kx[u] != kNoKey && kx[u] > prefix[0] && kx[u] < bx) bx = kx[u];
                if (ky[u] != kNoKey && ky[u] > prefix[1] && ky[u] < by) by = ky[u];
            }
        }
        bx = __reduce_min_sync(kFull, bx);
        by = __reduce_min_sync(kFull, by);
        if ((tid & 31) == 0) { atomicMin(&bc[8], bx); atomicMin(&bc[9], by); }
        __syncthreads();
        if (need0) v2[0] = bc[8];
        if (need1) v2[1] = bc[9];
    }
    if (tid == 0) {
        me_pair_result r;
        r.pdx = cnt[0] ? __fmul_rn(0.5f, __fadd_rn(funkey(prefix[0]), funkey(v2[0]))) : 0.0f;
        r.pdy = cnt[1] ? __fmul_rn(0.5f, __fadd_rn(funkey(prefix[1]), funkey(v2[1]))) : 0.0f;
        r.flag = (int)cnt[0] >= (a.min_valid > 1 ? a.min_valid : 1);
        r.nvalid_x = (int)cnt[0];
        r.nvalid_y = (int)cnt[1];
        r.n_textured = (int)bc[10];
        const long long sx = *reinterpret_cast<const long long *>(bc + 12), sy = *reinterpret_cast<const long long *>(bc + 14);
        r.mean_dx = cnt[0] ? (float)__ddiv_rn((double)sx, __dmul_rn((double)cnt[0], 1048576.0)) : 0.0f;
        r.mean_dy = cnt[1] ? (float)__ddiv_rn((double)sy, __dmul_rn((double)cnt[1], 1048576.0)) : 0.0f;
        a.pair_out[pair] = r;
        a.pair_ticket[pair] = 0;                    // leave the scratch clean for the next launch
    }
}

// ---------------------------------------------------------------- one work item -> block keys
// Runs one NDY x NQP candidate patch of macro-block j and folds its best candidate(s) into s_key.
template <int B, int NDY, int NQP, int PITCH, int LW, bool K64, int RB, bool SLICE16 = false>
__device__ __forceinline__ void run_item(const SearchArgs &a, bool active, int j, int dyi0, int s, int q0,
                                         const uint8_t *s_win, const uint8_t *s_ref,
                                         typename KeyOf<K64>::T *s_key, bool tile_interior, int dyi_lo, int dyi_hi,
                                         int X0, int lane, int ubase, int dbase, int slab_w, int jy = 0, int key_row = 0)
{
    using KeyT = typename KeyOf<K64>::T;
    using G = Geo<B>;
    constexpr KeyT kNone = ~KeyT(0);
    // Thread-local key = sad * KM + (index in patch): order-preserving for any KM > max index.  KM is
    // deliberately NOT a power of two: the compiler then emits IMAD (FMA pipe, idle here) instead of
    // LEA/SHF, which would compete with VABSDIFF4 for the ALU pipe.
    constexpr uint32_t KM = 37;
    static_assert(NDY * NQP <= 37, "patch too large for the packed local key");
    KeyT best[G::NSUB];
#pragma unroll
    for (int sb = 0; sb < G::NSUB; ++sb) best[sb] = kNone;
    if (active) {
        uint32_t acc[G::NSUB][NDY][NQP];
#pragma unroll
        for (int sb = 0; sb < G::NSUB; ++sb)
#pragma unroll
            for (int d = 0; d < NDY; ++d)
#pragma unroll
                for (int q = 0; q < NQP; ++q) acc[sb][d][q] = 0;
        const uint32_t *cur = reinterpret_cast<const uint32_t *>(s_win + (size_t)s * (a.copy_bytes + kCopySkew)) +
                              (dyi0 + jy * B) * (PITCH / 4) + j * G::BW + q0;
        const uint32_t *ref = reinterpret_cast<const uint32_t *>(s_ref) + jy * B * (kRefPitch / 4) + j * G::BW;
        sad_patch<B, NDY, NQP, PITCH, LW, SLICE16>(cur, ref, acc);

        // everything below is in SLAB-LOCAL index space: u = dx + R - ubase, dyi = dy + R - dbase
        const int u0 = 4 * q0 + s;
        int u_lo[G::NSUB], u_hi[G::NSUB];
        bool whole = dyi0 >= dyi_lo && dyi0 + NDY - 1 <= dyi_hi;
#pragma unroll
        for (int sb = 0; sb < G::NSUB; ++sb) {
            u_lo[sb] = 0; u_hi[sb] = slab_w - 1;
            if (!tile_interior) {
                const int x0 = X0 + j * G::MB + sb * B;
                u_lo[sb] = max(0, a.R - x0 - ubase); u_hi[sb] = min(slab_w - 1, min(2 * a.R, a.R + a.W - B - x0) - ubase);
                whole = whole && u0 >= u_lo[sb] && u0 + 4 * (NQP - 1) <= u_hi[sb];
            }
        }
        uint32_t lk[G::NSUB];                                           // sad * KM + index in patch, d-major
        // A real (warp-uniform) branch: left to itself the compiler if-converts the two paths and then
        // every patch pays the per-candidate domain test on the SAD op's pipe.
        if (__all_sync(__activemask(), whole)) {
#pragma unroll
            for (int sb = 0; sb < G::NSUB; ++sb) {
                lk[sb] = 0xffffffffu;
#pragma unroll
                for (int d = 0; d < NDY; ++d)
#pragma unroll
                    for (int q = 0; q < NQP; ++q) lk[sb] = min(lk[sb], acc[sb][d][q] * KM + (uint32_t)(d * NQP + q));
            }
        } else {
#pragma unroll
            for (int sb = 0; sb < G::NSUB; ++sb) {
                lk[sb] = 0xffffffffu;
#pragma unroll
                for (int d = 0; d < NDY; ++d)
#pragma unroll
                    for (int q = 0; q < NQP; ++q) {
                        const int dyi = dyi0 + d, u = u0 + 4 * q;
                        const bool ok = dyi >= dyi_lo && dyi <= dyi_hi && u >= u_lo[sb] && u <= u_hi[sb];
                        const uint32_t v = acc[sb][d][q] * KM + (uint32_t)(d * NQP + q);
                        if (ok) lk[sb] = min(lk[sb], v);
                    }
            }
        }
#pragma unroll
        for (int sb = 0; sb < G::NSUB; ++sb) {
            if (lk[sb] != 0xffffffffu) {
                const uint32_t sad = lk[sb] / KM;
                const int idx = (int)(lk[sb] - sad * KM), d = idx / NQP, q = idx - d * NQP;
                const uint32_t rank = (uint32_t)((dbase + dyi0 + d) * a.U + ubase + u0 + 4 * q);
                best[sb] = ((KeyT)sad << RB) | (KeyT)rank;
            }
        }
    }
    // block-level argmin: one shared atomic per warp when the warp sits in one macro-block
    const int kj = key_row + j;                                         // macro-block slot in the tile's key table
    const int j0 = __shfl_sync(kFull, kj, 0);
    const bool uniform = __all_sync(kFull, kj == j0);
#pragma unroll
    for (int sb = 0; sb < G::NSUB; ++sb) {
        if (uniform) {
            const KeyT m = warp_min_key(best[sb]);
            if (lane == 0 && m != kNone) atomicMin(&s_key[j0 * G::NSUB + sb], m);
        } else if (best[sb] != kNone) {
            atomicMin(&s_key[kj * G::NSUB + sb], best[sb]);
        }
    }
}

// ---------------------------------------------------------------- the search kernel
// MR = tiles of several block rows (a.tmr > 1).  A separate instantiation: carrying the row arithmetic in the
// one-row kernel cost it 3 % (r01e: 0.846 -> 0.819 at 1080p/16/+-32).
constexpr bool kSliceMR = true;               // multi-row tiles (small ranges): 16-row patches as two rolled 8-row slices
// search_tile: everything one CTA does for one unit (tile x candidate slab) -- the body of sad_search_kernel, and of every
// iteration of sad_scan_kernel (sad_scan.cuh), which walks units in a loop.  `smem` is the CTA's dynamic shared memory.
// SCAN: the build inside sad_scan_kernel, where one unit's latency is what counts: the pair's frame slots come from the caller
// (the barrier word) instead of the pair list, a tile with few blocks is finalized by a whole warp per block, and under
// ME_SCAN_TRACE thread 0 adds the SM-clock cycles of each phase to phase[0..5] (load, shifted copies, candidates, finalize,
// ticket, pair reduction) when `phase` is not null.
struct TileTrace {
    unsigned long long *phase;       // the traced CTA only: cycle sums per phase
    unsigned long long *unit_max;    // every CTA: atomicMax of (records written - t_step) in ns, the step's slowest unit
    unsigned long long *reduce;      // every CTA that reduces a pair: [0] cycles, [1] how often, [2..5] reduce_pair_small's parts
    unsigned long long t_step;       // globaltimer when this CTA began the step
};
template <int B, int NDY, int NQ, int PITCH, bool K64, int RB, bool MR, bool SCAN = false>
__device__ __forceinline__ void search_tile(const CUtensorMap &tm_win, const CUtensorMap &tm_ref, const SearchArgs &a, const int unit,
                                            uint8_t *smem, const TileTrace *trace = nullptr, int scan_prev = 0, int scan_cur = 0)
{
    constexpr bool TRC = SCAN;
    unsigned long long *phase = TRC && trace ? trace->phase : nullptr;
    long long tc0 = 0;
    int tci = 0;
    auto mark = [&]() {
        if constexpr (TRC) {
            if (phase && threadIdx.x == 0) {
                const long long t = clock64();
                phase[tci++] += (unsigned long long)(t - tc0);
                tc0 = t;
            }
        }
    };
    if constexpr (TRC) { if (phase && threadIdx.x == 0) tc0 = clock64(); }
    using KeyT = typename KeyOf<K64>::T;
    using G = Geo<B>;
    constexpr KeyT kNone = ~KeyT(0);
    constexpr int LWW = (NQ % 4 == 0) ? 4 : 1;       // wide patches start on 16-byte boundaries when NQ % 4 == 0

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int nslab = a.nsx * a.nsy;
    const int tile_id = unit / nslab, slab = unit - tile_id * nslab;               // slabs of a tile are adjacent units
    const int pair = tile_id / a.tiles_per_pair;
    const int tr = tile_id - pair * a.tiles_per_pair;
    const int ty = tr / a.tiles_x, tx = tr - ty * a.tiles_x;
    const int sy = slab / a.nsx, sx = slab - sy * a.nsx;
    const int ubase = sx * a.su, dbase = a.dy0 + sy * a.sd;
    const int lastx = sx == a.nsx - 1;
    const int slab_w = lastx ? a.U - ubase : a.su, slab_h = (sy == a.nsy - 1) ? a.dy0 + a.Uy - dbase : a.sd;
    const int prev_slot = SCAN ? scan_prev : __ldg(a.pairs + 2 * pair), cur_slot = SCAN ? scan_cur : __ldg(a.pairs + 2 * pair + 1);
    const int tmr = MR ? a.tmr : 1;
    const int X0 = tx * a.tmb * G::MB, Y0 = ty * tmr * B;
    const int nj = min(a.tmb, a.nmbx - tx * a.tmb);                    // macro-blocks in a row of this tile
    const int nblk = min(nj * G::NSUB, a.nbx - tx * a.tmb * G::NSUB);  // blocks in a row of this tile
    const int nrows = MR ? min(tmr, a.nby - ty * tmr) : 1;             // block rows in this tile

    uint8_t *s_win = smem;                                            // 4 shifted copies
    uint8_t *s_raw = a.x_off ? smem + 4 * (size_t)a.copy_bytes + 128 : smem;   // TMA destination (128-byte aligned)
    uint8_t *s_ref = smem + (size_t)a.n_bufs * a.copy_bytes + 128 + 256;       // copy skews + 256 B of slack for tile over-reads
    const int ref_bytes = kRefPitch * B * tmr;
    KeyT *s_key = reinterpret_cast<KeyT *>(s_ref + ref_bytes);         // [tmr][tmb * NSUB] <= kMaxTileKeys
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_ref + ref_bytes + kMaxTileKeys * 8);
    int *s_last = reinterpret_cast<int *>(s_bar + 1);

    // ---- this slab's share of the block row's domain, in slab-local index space.  dy groups that lie wholly
    //      outside it are not enumerated at all (most of the work on the 64-pixel-tall recordings), and a slab
    //      with none loads nothing.
    //      A tile of several block rows enumerates the union of its rows' ranges (the last row reaches furthest
    //      up, the first furthest down); each item then tests against its own row.
    const int twoR = 2 * a.R;
    const int Ylast = Y0 + (nrows - 1) * B;
    const int dyi_lo = max(0, a.R - Ylast - dbase), dyi_hi = min(slab_h - 1, min(twoR, a.R + a.H - B - Y0) - dbase);
    const int xlast = X0 + (nblk - 1) * B;
    const bool x_interior = a.R - X0 - ubase <= 0 && min(twoR, a.R + a.W - B - xlast) - ubase >= slab_w - 1;
    const bool tile_interior = x_interior && a.R - Y0 - dbase <= 0 && min(twoR, a.R + a.H - B - Ylast) - dbase >= slab_h - 1;
    // dy groups are laid from the first valid row (dyi_lo) in steps of NDY; the last one starts NDY rows before the last
    // valid row instead of hanging over it.  Its first rows then repeat candidates of the group before (same keys, harmless
    // for the argmin), and every group lies inside the valid range: no masked rows, test-free argmin path -- also for the
    // block rows at the top and bottom of the frame (all rows of the 64-pixel-tall recordings).
    const int dy_cap = max(dyi_hi - NDY + 1, 0);
    const int nd = dyi_hi >= dyi_lo ? (dyi_hi - dyi_lo) / NDY + 1 : 0; // 0: this slab holds no candidate of this block row
    for (int i = tid; i < kMaxTileKeys; i += (int)blockDim.x) s_key[i] = kNone;

    const int wx0 = X0 - a.R - a.x_off + ubase, wy0 = Y0 - a.R + dbase;   // wx0 is a multiple of 16 (su is)
    if (nd == 0) {
        __syncthreads();                                               // s_key initialised; nothing to load or search
    } else if (a.use_tma) {
        if (tid == 0) mbar_init(s_bar, 1);
        __syncthreads();
        if (tid == 0) {
            // the scan kernel reuses this shared memory unit after unit: the generic-proxy stores of the unit before (shifted copies,
            // reduction scratch; ordered before this point by the barrier) come before the async-proxy writes of the loads below
            if constexpr (SCAN) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            const uint32_t bytes = (uint32_t)a.n_row_boxes * a.box_rows * PITCH + (uint32_t)ref_bytes;
            mbar_expect_tx(s_bar, bytes);
            for (int rb = 0; rb < a.n_row_boxes; ++rb)
                tma_load_3d(s_raw + (size_t)rb * a.box_rows * PITCH, &tm_win, wx0, wy0 + rb * a.box_rows,
                            cur_slot, s_bar);
            tma_load_3d(s_ref, &tm_ref, X0, Y0, prev_slot, s_bar);
        }
        mbar_wait(s_bar, 0);
    } else {
        // Debug loader (ME_NO_TMA=1): same bytes with plain loads, zero outside the frame.
        const uint8_t *curf = a.frames + (long long)cur_slot * a.frame_stride;
        const uint8_t *prevf = a.frames + (long long)prev_slot * a.frame_stride;
        const int rows = a.n_row_boxes * a.box_rows;
        for (int i = tid; i < rows * PITCH; i += blockDim.x) {
            const int r = i / PITCH, c = i - r * PITCH;
            const int gx = wx0 + c, gy = wy0 + r;
            uint8_t v = 0;
            if (gx >= 0 && gx < a.W && gy >= 0 && gy < a.H) v = curf[(size_t)gy * a.frame_pitch + gx];
            s_raw[i] = v;
        }
        for (int i = tid; i < ref_bytes; i += blockDim.x) {
            const int r = i / kRefPitch, c = i - r * kRefPitch;
            const int gx = X0 + c, gy = Y0 + r;
            s_ref[i] = (gx < a.W && gy < a.H) ? prevf[(size_t)gy * a.frame_pitch + gx] : 0;   // r runs over all tmr block rows
        }
        __syncthreads();
    }

    mark();                                                            // [0] pair list + TMA
    // ---- derive the byte-shifted copies
    if (nd > 0) {
        build_shifted_copies<PITCH>(a, s_raw, s_win, tid, (int)blockDim.x);
        __syncthreads();
        if (a.use_tma && tid == 0) mbar_inval(s_bar);                  // the scan kernel initialises it again for its next unit
    }

    mark();                                                            // [1] shifted copies
    // ---- candidate patches.  The magic numbers of the usual case come precomputed from the host: a
    //      64-bit division per warp and CTA showed up as 10 ALU instructions per work item in r01c.
    const int lasty = sy == a.nsy - 1;
    const bool usual = nd == a.full_nd[lasty];
    const unsigned short *grp = a.group[lastx];
    const int n_wide = a.n_wide[lastx], n_narrow = a.n_narrow[lastx];
    const int njt = nj * nrows;                                        // macro-blocks in the tile, row-major
    const unsigned magic_row = a.magic_nj[tx == a.tiles_x - 1];
    const int key_stride = a.tmb;                                      // macro-block slots per row of s_key
    // row of macro-block jj, and that row's own share of the dy domain
    auto row_domain = [&](int jj, int dyi0, bool &act, int &jx, int &jy, int &lo, int &hi, bool &inter) {
        jy = fastdiv(jj, magic_row); jx = jj - jy * nj;
        const int y0r = Y0 + jy * B;
        lo = max(0, a.R - y0r - dbase); hi = min(slab_h - 1, min(twoR, a.R + a.H - B - y0r) - dbase);
        inter = x_interior && lo == 0 && hi == slab_h - 1;
        act = act && dyi0 + NDY - 1 >= lo && dyi0 <= hi;                // a dy group wholly outside this row's range: nothing to do
    };
    for (int rep = 0; rep < a.repeat; ++rep) {
    if (nd > 0) {   // wide patches: NDY x NQ
        const int per_j = nd * n_wide;
        const int total = njt * per_j;
        const unsigned magic_pj = usual ? a.magic_wide[lasty][lastx] : fastdiv_magic((unsigned)per_j);
        const int iters = (total + (int)blockDim.x - 1) / (int)blockDim.x;
        for (int it = 0; it < iters; ++it) {
            const int w_raw = it * (int)blockDim.x + tid;
            bool active = w_raw < total;
            const int w = active ? w_raw : total - 1;
            const int j = fastdiv(w, magic_pj), r = w - j * per_j;
            const int dq = fastdiv(r, a.magic_nw[lastx]), cg = r - dq * n_wide;                       // column group varies fastest across lanes
            const int dyi0 = min(dyi_lo + dq * NDY, dy_cap);
            const unsigned g = grp[cg];
            if constexpr (MR) {
                int jx, jy, lo, hi; bool inter;
                row_domain(j, dyi0, active, jx, jy, lo, hi, inter);
                run_item<B, NDY, NQ, PITCH, LWW, K64, RB, kSliceMR>(a, active, jx, dyi0, (int)(g >> 12), (int)(g & 0xfffu), s_win, s_ref,
                                                          s_key, inter, lo, hi, X0, lane, ubase, dbase, slab_w, jy, jy * key_stride);
            } else {
                run_item<B, NDY, NQ, PITCH, LWW, K64, RB>(a, active, j, dyi0, (int)(g >> 12), (int)(g & 0xfffu), s_win, s_ref,
                                                          s_key, tile_interior, dyi_lo, dyi_hi, X0, lane, ubase, dbase, slab_w);
            }
        }
    }
    if (nd > 0 && n_narrow > 0) {   // leftover columns: NDY x 1
        const int per_j = nd * n_narrow;
        const int total = njt * per_j;
        const unsigned magic_pj = usual ? a.magic_narrow[lasty][lastx] : fastdiv_magic((unsigned)per_j);
        const int iters = (total + (int)blockDim.x - 1) / (int)blockDim.x;
        for (int it = 0; it < iters; ++it) {
            // handed out from the LAST warp down: the first warps carry the extra wide pass of a ragged tile
            const int w_raw = it * (int)blockDim.x + ((int)blockDim.x - 1 - tid);
            bool active = w_raw < total;
            const int w = active ? w_raw : total - 1;
            const int j = fastdiv(w, magic_pj), r = w - j * per_j;
            const int dq = fastdiv(r, a.magic_nn[lastx]), cg = r - dq * n_narrow;
            const int dyi0 = min(dyi_lo + dq * NDY, dy_cap);
            const unsigned g = grp[n_wide + cg];
            if constexpr (MR) {
                int jx, jy, lo, hi; bool inter;
                row_domain(j, dyi0, active, jx, jy, lo, hi, inter);
                run_item<B, NDY, 1, PITCH, 1, K64, RB, kSliceMR>(a, active, jx, dyi0, (int)(g >> 12), (int)(g & 0xfffu), s_win, s_ref,
                                                       s_key, inter, lo, hi, X0, lane, ubase, dbase, slab_w, jy, jy * key_stride);
            } else {
                run_item<B, NDY, 1, PITCH, 1, K64, RB>(a, active, j, dyi0, (int)(g >> 12), (int)(g & 0xfffu), s_win, s_ref,
                                                       s_key, tile_interior, dyi_lo, dyi_hi, X0, lane, ubase, dbase, slab_w);
            }
        }
    }
    }
    __syncthreads();
    mark();                                                            // [2] candidates

    // ---- several slabs per tile: fold this CTA's block minima into the global keys; the CTA that
    //      arrives last at the tile's ticket owns the finalize step.
    if (nslab > 1) {
        const size_t kbase = (size_t)pair * a.nb + (size_t)ty * a.nbx + (size_t)tx * a.tmb * G::NSUB;
        KeyT *gkey = reinterpret_cast<KeyT *>(a.blk_key);               // 8 bytes reserved per block, KeyT used
        if (tid < nblk && s_key[tid] != kNone) atomicMin(&gkey[kbase + tid], s_key[tid]);
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            const unsigned old = atomicAdd(&a.tile_ticket[tile_id], 1u);
            *s_last = (old == (unsigned)nslab - 1u);
        }
        __syncthreads();
        if (!*s_last) return;
        __threadfence();
        if (tid < nblk) {
            s_key[tid] = __ldcg(&gkey[kbase + tid]);
            gkey[kbase + tid] = kNone;                                  // leave the scratch as we found it
        }
        if (tid == 0) a.tile_ticket[tile_id] = 0;
        __syncthreads();
    }

    // ---- per-block finalize (one warp per block)
    {
        const uint8_t *curf = a.frames + (long long)cur_slot * a.frame_stride;
        const uint8_t *prevf = a.frames + (long long)prev_slot * a.frame_stride;
        constexpr KeyT rmask = (KeyT(1) << RB) - 1;
        constexpr int per_warp = 32 / kFinLanes;
        const int grp = lane / kFinLanes;
        if (SCAN && nslab == 1 && nblk * nrows <= nwarps) {            // few blocks: a whole warp per block, a quarter of the passes
            if (warp < nblk * nrows) {
                const int jr = MR ? warp / nblk : 0, jb = warp - jr * nblk;
                const KeyT key = s_key[jr * a.tmb * G::NSUB + jb];
                finalize_block_smem<B, PITCH, 32>(a, pair, s_win, s_ref, jr, jb, tx * a.tmb * G::NSUB + jb, ty * tmr + jr,
                                                  (uint32_t)(key >> RB), (uint32_t)(key & rmask), lane, 0xffffffffu);
            }
        } else
        for (int jt = warp * per_warp + grp; jt < nblk * nrows; jt += nwarps * per_warp) {
            const int jr = MR ? jt / nblk : 0, jb = jt - jr * nblk;    // block row in the tile, block in the row
            const KeyT key = s_key[jr * a.tmb * G::NSUB + jb];
            const unsigned fmask = ((1u << kFinLanes) - 1u) << (grp * kFinLanes);
            if (nslab == 1)        // the whole candidate square of the block is still in this CTA's shared memory
                finalize_block_smem<B, PITCH>(a, pair, s_win, s_ref, jr, jb, tx * a.tmb * G::NSUB + jb, ty * tmr + jr,
                                              (uint32_t)(key >> RB), (uint32_t)(key & rmask), lane % kFinLanes, fmask);
            else
                finalize_block<B>(a, pair, prevf, curf, tx * a.tmb * G::NSUB + jb, ty * tmr + jr, (uint32_t)(key >> RB),
                                  (uint32_t)(key & rmask), lane % kFinLanes, fmask);
        }
    }
    __syncthreads();
    mark();                                                            // [3] finalize
    if constexpr (TRC) {
        if (trace && trace->unit_max && threadIdx.x == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            atomicMax(trace->unit_max, t - trace->t_step);
        }
    }

    // ---- ticket: the last CTA of this pair reduces it.  Thread 0 fences on both sides of its ticket (release: the CTA's
    //      records, ordered before it by the barrier above; acquire: the other CTAs' records, read through L2 by all threads
    //      after the barrier below) -- fences are cumulative, so the other threads need none (128 fences were ~2 k cycles)
    if (tid == 0) {
        __threadfence();
        const unsigned old = atomicAdd(&a.pair_ticket[pair], 1u);
        const bool last = old == (unsigned)a.tiles_per_pair - 1u;
        if (last) __threadfence();
        *s_last = last;
    }
    __syncthreads();
    mark();                                                            // [4] ticket
    if (*s_last) {
        long long tr0 = 0;
        if constexpr (TRC) { if (trace && tid == 0) tr0 = clock64(); }
        reduce_pair(a, pair, reinterpret_cast<unsigned *>(smem), TRC && trace ? trace->reduce + 2 : nullptr);
        if constexpr (TRC) {
            __syncthreads();
            if (trace && tid == 0) { atomicAdd(trace->reduce, (unsigned long long)(clock64() - tr0)); atomicAdd(trace->reduce + 1, 1ull); }
        }
    }
}

template <int B, int NDY, int NQ, int PITCH, bool K64, int RB, int MAXT, int MINB, bool MR>
__global__ void __launch_bounds__(MAXT, MINB)
sad_search_kernel(const __grid_constant__ CUtensorMap tm_win, const __grid_constant__ CUtensorMap tm_ref,
                  const SearchArgs a)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    search_tile<B, NDY, NQ, PITCH, K64, RB, MR>(tm_win, tm_ref, a, (int)blockIdx.x, smem);
}

} // namespace me
