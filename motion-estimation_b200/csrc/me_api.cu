// me_api.cu -- host side of libme_b200.so: the C ABI declared in include/me_b200.h.
//
// Owns the device frame ring, the pinned staging buffers, the two streams (copy / compute), the TMA
// descriptors and the launch configuration.  There is no CPU fallback: every entry point either
// runs the CUDA path or returns an error.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <cmath>
#include <mutex>
#include <set>
#include <string>
#include <utility>
#include <vector>

#include "../../include/me_b200.h"
#include "sad_search.cuh"
#include "sad_persistent.cuh"
#include "sad_scan.cuh"
#include "kernel_choice.h"
#include "misc_kernels.cuh"
#include "clahe.cuh"
#include "nlm.cuh"

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CU_TRY(expr)                                                                              \
    do {                                                                                          \
        cudaError_t e_ = (expr);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail(ME_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// (x, y, frame) uint8 view of a frame stack; a box is box_w x box_h bytes of one frame.
int make_tensor_map(CUtensorMap *tm, const uint8_t *base, int W, int H, int n_frames, size_t pitch,
                    int box_w, int box_h)
{
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return fail(ME_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)n_frames};
    cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * (cuuint64_t)H};
    cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1u};
    cuuint32_t estr[3] = {1u, 1u, 1u};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(ME_ECUDA, "cuTensorMapEncodeTiled failed (CUresult %d; W=%d H=%d pitch=%zu box=%dx%d)", (int)r, W,
                    H, pitch, box_w, box_h);
    return ME_OK;
}

using me::KernelChoice;
using me::kNDY;
using me::kNQ;
using me::kMaxThreads;
using me::kPitches;

// One instantiation per (block size, window pitch, key width), each pitch x block in its own translation unit
// (search_inst.cu); see DESIGN.md "inner loop".
KernelChoice pick_kernel(int B, int pitch, bool k64)
{
#define ME_PICK(P)                                                                                   \
    switch (B) {                                                                                     \
    case 8:  return me::kernel_choice_##P##_8(k64);                                                  \
    case 16: return me::kernel_choice_##P##_16(k64);                                                 \
    default: return me::kernel_choice_##P##_32(k64);                                                 \
    }
    switch (pitch) {
    case 144: ME_PICK(144)
    case 208: ME_PICK(208)
    default:  ME_PICK(240)
    }
#undef ME_PICK
}

// A kernel's dynamic shared-memory limit belongs to the function and the device, not to a context: raise it once to the
// opt-in maximum so that contexts (and the static-scratch path) with different window sizes that share an
// instantiation cannot lower it under each other's launches.
constexpr int kSmemOptInMax = 227 * 1024;
int raise_smem_limit(const void *fn, bool prefer_smem = true)
{
    static std::mutex mu;
    static std::set<std::pair<const void *, int>> done;
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    if (done.count({fn, dev})) return ME_OK;
    cudaFuncAttributes fa;
    CU_TRY(cudaFuncGetAttributes(&fa, fn));          // static shared memory counts against the same 227 KB
    CU_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemOptInMax - (int)fa.sharedSizeBytes));
    if (prefer_smem) CU_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    done.insert({fn, dev});
    return ME_OK;
}


int env_int(const char *name, int dflt)
{
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

} // namespace

struct me_ctx {
    int device = 0;
    int H = 0, W = 0;
    me_params p{};
    int max_frames = 0, max_pairs = 0;
    size_t pitch = 0, frame_stride = 0;
    uint8_t *d_frames = nullptr;
    // staging
    uint8_t *h_stage[2] = {nullptr, nullptr};
    size_t stage_bytes = 0;
    cudaEvent_t stage_free[2] = {nullptr, nullptr};
    int stage_next = 0;
    uint8_t *d_bgr = nullptr;
    size_t d_bgr_bytes = 0;
    uint8_t *d_luts = nullptr;          // CLAHE lookup tables, allocated on first use
    size_t d_luts_bytes = 0;
    uint8_t *d_nlm = nullptr;           // denoised frames + accumulators + tables of me_denoise, allocated on first use
    size_t d_nlm_bytes = 0;
    int *d_nlm_meta = nullptr;          // weight tables + per-frame table index of me_denoise
    size_t d_nlm_meta_bytes = 0;
    std::vector<int> nlm_meta;          // what they were built for: (n_frames, windows, h, temporal windows); same again = reuse
    int nlm_meta_shift = 0;
    bool nlm_lut_small = false;
    // streams / events
    cudaStream_t s_copy = nullptr, s_compute = nullptr;
    cudaEvent_t ev_upload = nullptr, ev_search = nullptr, ev_k0 = nullptr, ev_k1 = nullptr;
    bool have_upload = false, have_search = false;
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;     // me_timer_start / me_timer_stop
    cudaEvent_t ev_c0 = nullptr, ev_c1 = nullptr;     // me_copy_timer_start / me_copy_timer_stop
    cudaEvent_t ev_rec = nullptr;                     // last me_copy_pair_records_device: the next search must not overwrite the records before it
    bool have_rec = false;
    // the last few searches and the ring slots they read: an upload waits only for those it overwrites
    struct SearchRecord { int lo = 0, hi = -1; cudaEvent_t done = nullptr; bool used = false; };
    SearchRecord recent[4];
    int recent_next = 0;
    // pairs + results
    // Pair lists and pair records live in a ring of kSearchRing slots (slot s at base + s * max_pairs entries), so that a search
    // can be queued while the one before it still runs and its records have not been fetched: me_search never waits for the
    // previous search, only for the one kSearchRing launches ago.
    static constexpr int kSearchRing = 4;
    int32_t *h_pairs = nullptr, *d_pairs = nullptr;
    me_pair_result *h_pair_out = nullptr;
    int slot = 0;                                  // ring slot (= ticket) of the most recent search
    cudaEvent_t ev_slot[kSearchRing] = {nullptr, nullptr, nullptr, nullptr};
    bool slot_used[kSearchRing] = {false, false, false, false};
    int slot_pairs[kSearchRing] = {0, 0, 0, 0};
    cudaStream_t s_fetch = nullptr;                // record fetches: neither behind queued searches nor behind frame uploads
    int last_pairs = 0;
    // launch configuration
    me::SearchArgs args{};
    CUtensorMap tm_win{}, tm_ref{};
    KernelChoice kc{};
    int threads = 0, smem = 0, sm_count = 0, win_pitch = 0;
    int plan_pairs = 1;                           // pairs per launch the tile shape is planned for (small frames: enough CTAs)
    me_ctx *big = nullptr;                        // second plan for launches of many pairs (plan fields + tensor maps only)
    bool force_persistent = false;                // ME_PERSISTENT=2: also for launches of a few units (tests)
    int pstage_bytes = 0, smem_persistent = 0;    // resident kernel: one stage of window copies, total dynamic smem (0 = not usable)
    int *d_ready = nullptr, *d_error = nullptr;  // resident kernel: queue of completed pairs, wait-timeout flag
    size_t n_tile_tickets = 0;                    // entries of args.tile_ticket (slabbed plans only)
    uint8_t *d_scan = nullptr;                    // me_anchored_scan on the device: [records max_frames | state 16 ints | barrier: arrivals, -, 64-bit word | trace]
    int64_t launches = 0;
    int64_t cand_px = 0, cand_px_valid = 0;
};

namespace {

// Fill ctx->args / launch shape for the context's geometry.  Returns ME_OK or an error.
int configure(me_ctx *c)
{
    const int B = c->p.block, R = c->p.range, U = 2 * R + 1, W = c->W, H = c->H;
    const int MB = B < 16 ? 16 : B, NSUB = MB / B, NDY = kNDY, NQ = kNQ;
    me::SearchArgs &a = c->args;
    // vertical range: the caller's, and never more than H - B -- a larger |dy| leaves the frame for every block (the 64-pixel-tall
    // recordings: +-48 of a +-64 square), so dropping those rows of the candidate square changes no result
    int Ry = (c->p.range_y > 0 && c->p.range_y < R) ? c->p.range_y : R;
    if (Ry > H - B && H - B >= 1) Ry = H - B;
    const int Uy = 2 * Ry + 1;
    a.W = W; a.H = H; a.R = R; a.U = U;
    a.Ry = Ry; a.Uy = Uy; a.dy0 = R - Ry;
    a.nbx = W / B; a.nby = H / B; a.nb = a.nbx * a.nby;
    a.nmbx = ceil_div(a.nbx, NSUB);
    a.frame_pitch = (int)c->pitch;
    a.frame_stride = (long long)c->frame_stride;
    a.min_range = c->p.min_range; a.min_curv = c->p.min_curv; a.subpixel = c->p.subpixel; a.min_valid = c->p.min_valid;

    // key width: (sad << RB) | rank must fit; 32-bit keys use RB = 16 (B <= 16) or 14 (B = 32)
    if (R > 128) return fail(ME_EUNSUPPORTED, "range %d: ranges up to 128 are built", R);
    const bool k64 = (long long)U * U > (B == 32 ? (1ll << 14) : (1ll << 16));   // R = 128 needs a 17-bit rank

    // TMA needs a 16-byte aligned x origin; tiles and slabs start at multiples of 16, so the slack is fixed.
    a.x_off = (16 - R % 16) % 16;
    a.n_bufs = a.x_off ? 5 : 4;

    // Launch plan.  The candidate square is cut into nsx x nsy slabs (one CTA each) until a slab's
    // window fits the shared-memory cap; within that, the widest tile and the fewest slabs win.  The
    // window pitch is a template constant (every shared load becomes base + immediate) from a short
    // list of odd multiples of 16 bytes: with NDY = 5 the 8 lanes of a quarter-warp, which read rows
    // NDY apart, then land in 8 different 16-byte bank groups (r01 ncu: a 256-byte pitch cost 5.4
    // wavefronts per LDS.128 and made the kernel shared-memory bound).
    // 3 CTAs per SM for 8/16-pixel blocks; 32-pixel blocks do better with two larger tiles per SM (r01 sweep: 0.74 vs 0.50)
    const int smem_cap = env_int("ME_SMEM_CAP", B == 32 ? 110 * 1024 : 72 * 1024);
    const int forced_tmb = env_int("ME_TBX", 0), forced_pitch = env_int("ME_PITCH", 0);
    const int forced_nsx = env_int("ME_NSX", 0), forced_nsy = env_int("ME_NSY", 0);
    const long long work_enough = 32000;          // block-candidates per CTA beyond which a bigger CTA buys nothing
    const int forced_tmr = env_int("ME_TMR", 0);
    struct Plan { int nsx = 0, nsy = 0, su = 0, sd = 0, pitch = 0, tmb = 0, tmr = 1, smem = 0, copy = 0, box_rows = 0, n_row_boxes = 0; long long work = 0; bool under = false; double balance = 0; };
    Plan best;
    for (int tsx = 1; tsx <= 8; ++tsx) {
        int su = tsx == 1 ? U : ((U / tsx) & ~15);
        if (su < 16 && tsx > 1) break;
        const int nsx = tsx == 1 ? 1 : U / su;
        if (tsx > 1 && nsx != tsx) continue;                       // each slab count is tried once
        if (forced_nsx && nsx != forced_nsx) continue;
        const int wmax = U - (nsx - 1) * su;                        // the last slab is the widest
        for (int tsy = 1; tsy <= 16; ++tsy) {
            int sd = tsy == 1 ? Uy : (Uy / tsy) / NDY * NDY;
            if (sd < NDY && tsy > 1) break;
            const int nsy = tsy == 1 ? 1 : Uy / sd;
            if (tsy > 1 && nsy != tsy) continue;
            if (forced_nsy && nsy != forced_nsy) continue;
            const int hmax = Uy - (nsy - 1) * sd;
            // Block rows per tile: more than one only without slabs (a small range leaves little work per row,
            // and the window's 2R halo rows are then shared by the rows of the tile).
            for (int tmr = 1; tmr <= (nsx * nsy == 1 ? 8 : 1); ++tmr) {
            if (forced_tmr && tmr != forced_tmr) continue;
            if (tmr > 1 && (tmr > a.nby || tmr * B > 256)) break;
            // a row that already carries enough work stays a tile of its own (and can run in the resident kernel)
            if (tmr > 1 && !forced_tmr && (long long)((kPitches[2] - a.x_off - (wmax - 1)) / MB) * NSUB * wmax * hmax >= work_enough) break;
            const int win_rows = ceil_div(hmax, NDY) * NDY + tmr * B - 1;
            const int n_row_boxes = ceil_div(win_rows, 256), box_rows = ceil_div(win_rows, n_row_boxes);
            const int alloc_rows = n_row_boxes * box_rows;
            for (int pi = 0; pi < 3; ++pi) {
                const int pitch = kPitches[pi];
                if (forced_pitch && pitch != forced_pitch) continue;
                int tmb = (pitch - a.x_off - (wmax - 1)) / MB;
                if (tmb > 256 / MB) tmb = 256 / MB;                 // reference strip is 256 bytes wide
                if (tmb * NSUB > me::kMaxKeys) tmb = me::kMaxKeys / NSUB;
                if (tmb * NSUB * tmr > me::kMaxTileKeys) tmb = me::kMaxTileKeys / (NSUB * tmr);
                if (tmb > a.nmbx) tmb = a.nmbx;
                if (forced_tmb) { if (forced_tmb > tmb) continue; tmb = forced_tmb; }
                if (tmb < 1) continue;
                // small frames: keep at least ~2 CTAs per SM in flight for a 4-pair batch
                if (tmr > 1 && !forced_tmr &&
                    (long long)ceil_div(a.nmbx, tmb) * ceil_div(a.nby, tmr) * c->plan_pairs < 2ll * c->sm_count) continue;
                if (!forced_tmb)
                    while (tmb > 1 && (long long)ceil_div(a.nmbx, tmb) * ceil_div(a.nby, tmr) * nsx * nsy * c->plan_pairs < 2ll * c->sm_count) --tmb;
                // every tile width from the widest that fits down to half of it: a narrower tile can
                // divide its work items over the CTA's warps more evenly (r01: 8 blocks beat 9 by 1.2 %)
                const int tmb_lo = forced_tmb ? tmb : (tmb + 1) / 2;
                for (int t = tmb; t >= tmb_lo; --t) {
                    Plan p;
                    p.nsx = nsx; p.nsy = nsy; p.su = su; p.sd = sd; p.pitch = pitch; p.tmb = t; p.tmr = tmr;
                    p.box_rows = box_rows; p.n_row_boxes = n_row_boxes;
                    p.copy = round_up(alloc_rows * pitch, 128);
                    p.smem = a.n_bufs * p.copy + 128 + 256 + me::kRefPitch * B * tmr + me::kMaxTileKeys * 8 + 64;
                    if (p.smem < 20 * 1024) p.smem = 20 * 1024;          // the pair reduction's histograms live in the same memory
                    if (p.smem > 227 * 1024) continue;
                    p.under = p.smem <= smem_cap;
                    p.work = (long long)t * NSUB * tmr * wmax * hmax;
                    // work items of a full tile over 256 threads: fraction of warp passes that are full
                    const int nqw = ((wmax - 1) / 4 + 1) / NQ + 3 * (((wmax - 2) / 4 + 1) / NQ);   // ~ wide column groups
                    const long long items = (long long)t * tmr * ceil_div(hmax, NDY) * (nqw > 0 ? nqw : 1);
                    p.balance = (double)items / (double)(ceil_div((int)items, 256) * 256)
                              * (double)a.nmbx / (double)(ceil_div(a.nmbx, t) * t)       // ... and of tiles that are full
                              * (double)a.nby / (double)(ceil_div(a.nby, tmr) * tmr);
                    const long long pw = p.work < work_enough ? p.work : work_enough, bw = best.work < work_enough ? best.work : work_enough;
                    bool better;
                    if (!best.tmb) better = true;
                    else if (p.under != best.under) better = p.under;
                    else if (!p.under) better = p.smem < best.smem;
                    else if ((p.nsx * p.nsy == 1) != (best.nsx * best.nsy == 1)) better = p.nsx * p.nsy == 1;   // no slabs when one window fits
                    else if (pw != bw) better = pw > bw;
                    else if (p.balance > best.balance + 0.015 || p.balance < best.balance - 0.015) better = p.balance > best.balance;
                    else if (p.nsx * p.nsy != best.nsx * best.nsy) better = p.nsx * p.nsy < best.nsx * best.nsy;
                    else if (p.tmb != best.tmb) better = p.tmb > best.tmb;
                    else better = p.smem < best.smem;
                    if (better) best = p;
                }
            }
            }
        }
    }
    if (!best.tmb)
        return fail(ME_EUNSUPPORTED, "block %d with range %d: no slab of the candidate square fits one TMA box and 227 KB of "
                    "shared memory", B, R);
    a.nsx = best.nsx; a.nsy = best.nsy; a.su = best.su; a.sd = best.sd;
    a.tmb = best.tmb; a.tmr = best.tmr; a.copy_bytes = best.copy; a.box_rows = best.box_rows; a.n_row_boxes = best.n_row_boxes;
    c->win_pitch = best.pitch;
    c->smem = best.smem;
    c->kc = pick_kernel(B, best.pitch, k64);
    if (best.tmr > 1) c->kc.fn = c->kc.fn_multirow;
    a.tiles_x = ceil_div(a.nmbx, a.tmb);
    a.tiles_per_pair = a.tiles_x * ceil_div(a.nby, a.tmr);
    {
        auto magic = [](int d) { return d <= 1 ? 0u : (unsigned)(0x100000000ull / (unsigned)d + 1ull); };
        a.magic_nj[0] = magic(a.tmb);
        a.magic_nj[1] = magic(a.nmbx - (a.tiles_x - 1) * a.tmb);
    }

    // candidate tiling inside a slab of width w: per shift s the columns q with 4q + s < w, in NQ-wide
    // groups (16-byte aligned starts) plus single leftover columns.  Table 0: regular slab, 1: last slab.
    for (int t = 0; t < 2; ++t) {
        const int w = t == 0 ? (a.nsx > 1 ? a.su : U) : U - (a.nsx - 1) * a.su;
        int ng = 0;
        a.n_wide[t] = a.n_narrow[t] = 0;
        // Order: group index major, shift minor.  Consecutive lanes then take the four shifted copies of
        // two adjacent 16-byte column chunks; with the copies skewed 32 bytes apart (kCopySkew) the eight
        // lanes of a quarter-warp hit eight different bank groups.  Lanes walk the column groups fastest, so
        // the one padded dy group of a slab (U is rarely a multiple of NDY) stays in the last warp or two
        // and every other warp takes the fast, test-free argmin path.
        int nqs[4], nws[4], maxw = 0, maxn = 0;
        for (int s = 0; s < 4; ++s) {
            nqs[s] = (w - 1 - s) >= 0 ? (w - 1 - s) / 4 + 1 : 0;
            nws[s] = nqs[s] / NQ;
            if (nws[s] > maxw) maxw = nws[s];
            if (nqs[s] - nws[s] * NQ > maxn) maxn = nqs[s] - nws[s] * NQ;
        }
        for (int g = 0; g < maxw; ++g)
            for (int s = 0; s < 4; ++s)
                if (g < nws[s]) {
                    if (ng >= me::kMaxGroups) return fail(ME_EUNSUPPORTED, "range %d needs more than %d column groups", R, me::kMaxGroups);
                    a.group[t][ng++] = (unsigned short)((s << 12) | (g * NQ));
                    ++a.n_wide[t];
                }
        for (int k = 0; k < maxn; ++k)
            for (int s = 0; s < 4; ++s)
                if (nws[s] * NQ + k < nqs[s]) {
                    if (ng >= me::kMaxGroups) return fail(ME_EUNSUPPORTED, "range %d needs more than %d column groups", R, me::kMaxGroups);
                    a.group[t][ng++] = (unsigned short)((s << 12) | (nws[s] * NQ + k));
                    ++a.n_narrow[t];
                }
    }

    // resident (persistent) form: two stages of window copies + reduction scratch + control
    c->pstage_bytes = round_up(a.n_bufs * a.copy_bytes + 128 + 256 + me::kRefPitch * B, 128);
    c->smem_persistent = me::kStages * c->pstage_bytes + (2 * me::kSelBins + 64) * 4 + 1024;
    if (c->smem_persistent > 227 * 1024 || !c->kc.fn_persistent || env_int("ME_PERSISTENT", 1) == 0) c->smem_persistent = 0;
    c->force_persistent = env_int("ME_PERSISTENT", 1) == 2;
    if (c->smem_persistent && env_int("ME_PERSISTENT", 1) != 2) {
        // The resident kernel pays when searching a unit (tile x slab) takes the consumer warps clearly longer than
        // loading and shifting its window takes the producers; the realigning copy build of windows whose origin is
        // not 16-byte aligned (R not a multiple of 16) costs about three times the plain one.  Measured (r01f,
        // tools/cmp_persist.sh, one-row tiles): wins from a ratio of about 1.7 -- 16x16 +-32 +2.5 %, +-48 +4.6 %,
        // +-64 (2x3 slabs) +2.3 %, 4K +-64 +4.5 %; 8x8 +-32 +2.5 %, +-48 +8 %, 4K +-64 (1x2 slabs) +9 %; 32x32 +-32
        // +2 %, +-40 +12 % -- and loses below it (8x8 +-24: -13 %, 16x16 +-20: -2 %).
        const double cand = (double)U * Uy / (a.nsx * a.nsy);                                             // candidates of an average slab
        const double t_cons = (double)a.tmb * NSUB * cand * B * B / 256.0;                               // cycles at 64 lanes x 4 px
        const double build = (double)a.n_row_boxes * a.box_rows * (c->win_pitch / 16) / 32.0 * 120.0 / me::kProducerWarps;
        const double t_prod = 6000.0 + build * (a.x_off == 0 ? 1.0 : (a.x_off & 3) ? 3.0 : 1.5);   // whole-word offsets skip the selects
        if (t_cons < 1.7 * t_prod) c->smem_persistent = 0;
    }
    if (a.tmr > 1) c->smem_persistent = 0;          // the resident kernel takes one block row per unit

    {   // fastdiv magics of full slabs (see sad_search.cuh)
        auto magic = [](int d) { return d <= 1 ? 0u : (unsigned)(0x100000000ull / (unsigned)d + 1ull); };
        for (int ly = 0; ly < 2; ++ly) {
            const int h = ly == 0 ? (a.nsy > 1 ? a.sd : Uy) : Uy - (a.nsy - 1) * a.sd;
            a.full_nd[ly] = ceil_div(h, NDY);
            a.magic_nd[ly] = magic(a.full_nd[ly]);
            for (int lx = 0; lx < 2; ++lx) {
                a.magic_wide[ly][lx] = magic(a.full_nd[ly] * a.n_wide[lx]);
                a.magic_narrow[ly][lx] = magic(a.full_nd[ly] * a.n_narrow[lx]);
            }
        }
        for (int lx = 0; lx < 2; ++lx) { a.magic_nw[lx] = magic(a.n_wide[lx]); a.magic_nn[lx] = magic(a.n_narrow[lx]); }
    }

    // Threads: whole multiples of 4 warps so that every scheduler of the SM gets the same number of warps
    // (r01 sweep: 224 threads = 7 warps ran 13 % slower than 128 or 256).
    const int total = a.tmb * a.tmr * ceil_div(Uy - (a.nsy - 1) * a.sd, NDY) * a.n_wide[1];
    c->threads = env_int("ME_THREADS", total >= 384 ? 256 : 128);
    if (c->threads % 32 || c->threads < 32 || c->threads > c->kc.maxt)
        return fail(ME_EINVAL, "ME_THREADS=%d is not a multiple of 32 in [32,%d]", c->threads, c->kc.maxt);
    a.use_tma = env_int("ME_NO_TMA", 0) ? 0 : 1;
    a.repeat = env_int("ME_REPEAT", 1) > 0 ? env_int("ME_REPEAT", 1) : 1;

    // Work accounting: nominal = what the ALU executes, valid = in-frame candidates.
    c->cand_px = (long long)a.nb * B * B * U * Uy;
    long long valid = 0;
    for (int by = 0; by < a.nby; ++by) {
        const int y0 = by * B;
        const int ny = (Ry < H - B - y0 ? Ry : H - B - y0) - (-Ry > -y0 ? -Ry : -y0) + 1;
        long long sx = 0;
        for (int bx = 0; bx < a.nbx; ++bx) {
            const int x0 = bx * B;
            sx += (R < W - B - x0 ? R : W - B - x0) - (-R > -x0 ? -R : -x0) + 1;
        }
        valid += sx * ny;
    }
    c->cand_px_valid = valid * B * B;
    return ME_OK;
}

// The plan a launch of n_pairs runs with: the context's own (tiles small enough to fill the GPU with a few pairs), or
// the one made for max_pairs when the launch is large enough to fill the GPU with its bigger tiles.
const me_ctx *pick_plan(const me_ctx *c, int n_pairs)
{
    if (c->big && (long long)c->big->args.tiles_per_pair * c->big->args.nsx * c->big->args.nsy * n_pairs >= 2ll * c->sm_count)
        return c->big;
    return c;
}

// pl: the plan (c itself or c->big); frame_pitch 0 = the context's ring layout.
int launch_search(me_ctx *c, const me_ctx *pl, const uint8_t *frames, const CUtensorMap &tm_win, const CUtensorMap &tm_ref,
                  int n_pairs, cudaStream_t st, size_t frame_pitch = 0, int slot = 0)
{
    me::SearchArgs a = pl->args;
    a.frames = frames;
    a.pairs = c->d_pairs + (size_t)slot * 2 * c->max_pairs;
    a.pair_out = a.pair_out + (size_t)slot * c->max_pairs;
    if (frame_pitch) { a.frame_pitch = (int)frame_pitch; a.frame_stride = (long long)frame_pitch * c->H; }
    const long long units = (long long)a.tiles_per_pair * n_pairs * a.nsx * a.nsy;
    if (c->have_rec) { CU_TRY(cudaStreamWaitEvent(st, c->ev_rec, 0)); c->have_rec = false; }
    CU_TRY(cudaEventRecord(c->ev_k0, st));
    if (pl->smem_persistent && (units >= 2ll * c->sm_count || pl->force_persistent)) {
        // enough work to keep one resident CTA per SM busy: the warp-specialised kernel (sad_persistent.cuh)
        me::PersistArgs pa;
        pa.n_units = (int)units; pa.n_pairs = n_pairs; pa.stage_bytes = pl->pstage_bytes;
        pa.ready_pairs = c->d_ready; pa.error_flag = c->d_error;
        CU_TRY(cudaMemsetAsync(c->d_ready, 0xff, sizeof(int) * (size_t)(n_pairs + 2), st));   // pair queue + unit counter
        void *params[4] = {(void *)&tm_win, (void *)&tm_ref, (void *)&a, (void *)&pa};
        const unsigned grid = units < c->sm_count ? (unsigned)units : (unsigned)c->sm_count;
        CU_TRY(cudaLaunchKernel(pl->kc.fn_persistent, dim3(grid), dim3((unsigned)pl->kc.persist_threads), params,
                                (size_t)pl->smem_persistent, st));
    } else {
        void *params[3] = {(void *)&tm_win, (void *)&tm_ref, (void *)&a};
        CU_TRY(cudaLaunchKernel(pl->kc.fn, dim3((unsigned)units), dim3((unsigned)pl->threads), params, (size_t)pl->smem, st));
    }
    CU_TRY(cudaEventRecord(c->ev_k1, st));
    c->launches += 1;
    return ME_OK;
}

int check_pairs(me_ctx *c, const int32_t *pairs, int n_pairs, int n_frames)
{
    if (!pairs || n_pairs < 1 || n_pairs > c->max_pairs)
        return fail(ME_EINVAL, "n_pairs=%d outside 1..%d (or pairs is NULL)", n_pairs, c->max_pairs);
    for (int i = 0; i < 2 * n_pairs; ++i)
        if (pairs[i] < 0 || pairs[i] >= n_frames)
            return fail(ME_EINVAL, "pair %d references frame %d, outside 0..%d", i / 2, pairs[i], n_frames - 1);
    return ME_OK;
}

// The resident kernel gave up waiting for a pair (its 2 s guard; never seen in a correct run, but a debugger, a
// profiler replay or GPU time-slicing can stretch a launch that far).  The search that hit it is lost; the tickets
// and keys it left half-counted are put back so that the context's next search starts clean.
int report_search_timeout(me_ctx *c)
{
    const size_t nb = (size_t)c->args.nb * (size_t)c->max_pairs;
    CU_TRY(cudaStreamSynchronize(c->s_compute));
    CU_TRY(cudaMemset(c->args.pair_ticket, 0, sizeof(unsigned) * (size_t)c->max_pairs));
    if (c->args.tile_ticket) CU_TRY(cudaMemset(c->args.tile_ticket, 0, c->n_tile_tickets * sizeof(unsigned)));
    if (c->args.blk_key) CU_TRY(cudaMemset(c->args.blk_key, 0xff, nb * sizeof(unsigned long long)));
    CU_TRY(cudaMemset(c->d_error, 0, sizeof(int)));
    return fail(ME_ECUDA, "the resident search kernel gave up waiting for a frame pair to complete; the results of this "
                          "search are invalid, the context has been reset and can be used again");
}

// Rows of `width` bytes from one pitched buffer to another.  When both pitches equal the width the rows are one contiguous
// run and go as ONE linear copy: a 2-D copy is a DMA descriptor per row (1920-byte rows at 1080p), which held pinned
// host <-> device transfers to ~16 GB/s (r02 bench: the end-to-end arm was bound by exactly that).
cudaError_t copy_rows(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t rows, cudaMemcpyKind kind,
                      cudaStream_t st)
{
    if (dpitch == width && spitch == width) return cudaMemcpyAsync(dst, src, width * rows, kind, st);
    return cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, kind, st);
}

// Host rows -> pinned staging -> device slots, double-buffered on the copy stream.
int upload_impl(me_ctx *c, const uint8_t *src, size_t row_pitch, int first_slot, int n_frames, bool bgr)
{
    const size_t row_bytes = (size_t)c->W * (bgr ? 3 : 1);
    if (!src || n_frames < 1 || first_slot < 0 || first_slot + n_frames > c->max_frames)
        return fail(ME_EINVAL, "upload of %d frames at slot %d does not fit the %d-slot ring (or NULL source)", n_frames,
                    first_slot, c->max_frames);
    if (row_pitch < row_bytes) return fail(ME_EINVAL, "row_pitch %zu < %zu bytes per row", row_pitch, row_bytes);
    CU_TRY(cudaSetDevice(c->device));
    for (auto &r : c->recent)                        // do not overwrite frames a running search still reads
        if (r.used && r.lo < first_slot + n_frames && r.hi >= first_slot) CU_TRY(cudaStreamWaitEvent(c->s_copy, r.done, 0));
    const size_t frame_bytes = row_bytes * c->H;
    // A source that is already page-locked (cudaHostAlloc / cudaHostRegister / torch pin_memory) is
    // copied straight from where it lies; anything else goes through the pinned staging ring.
    cudaPointerAttributes pa;
    const bool pinned = cudaPointerGetAttributes(&pa, src) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    if (!pinned) (void)cudaGetLastError();
    if (pinned && !bgr) {
        uint8_t *dst = c->d_frames + (size_t)first_slot * c->frame_stride;
        CU_TRY(copy_rows(dst, c->pitch, src, row_pitch, row_bytes, (size_t)n_frames * c->H, cudaMemcpyHostToDevice, c->s_copy));
        CU_TRY(cudaEventRecord(c->ev_upload, c->s_copy));
        c->have_upload = true;
        return ME_OK;
    }
    const int per_chunk = (int)(c->stage_bytes / frame_bytes);
    if (per_chunk < 1) return fail(ME_ENOMEM, "staging buffer smaller than one frame");
    for (int f = 0; f < n_frames; f += per_chunk) {
        const int nf = (n_frames - f < per_chunk) ? n_frames - f : per_chunk;
        const int sb = c->stage_next;
        c->stage_next ^= 1;
        CU_TRY(cudaEventSynchronize(c->stage_free[sb]));
        const uint8_t *s = src + (size_t)f * c->H * row_pitch;
        if (row_pitch == row_bytes) {
            memcpy(c->h_stage[sb], s, frame_bytes * nf);
        } else {
            for (size_t r = 0; r < (size_t)nf * c->H; ++r) memcpy(c->h_stage[sb] + r * row_bytes, s + r * row_pitch, row_bytes);
        }
        uint8_t *dst = c->d_frames + (size_t)(first_slot + f) * c->frame_stride;
        if (!bgr) {
            CU_TRY(copy_rows(dst, c->pitch, c->h_stage[sb], row_bytes, row_bytes, (size_t)nf * c->H, cudaMemcpyHostToDevice, c->s_copy));
        } else {
            const size_t need = frame_bytes * nf;
            if (c->d_bgr_bytes < need) {
                if (c->d_bgr) { CU_TRY(cudaStreamSynchronize(c->s_copy)); CU_TRY(cudaFree(c->d_bgr)); }
                c->d_bgr = nullptr; c->d_bgr_bytes = 0;
                CU_TRY(cudaMalloc(&c->d_bgr, c->stage_bytes));
                c->d_bgr_bytes = c->stage_bytes;
            }
            CU_TRY(cudaMemcpyAsync(c->d_bgr, c->h_stage[sb], need, cudaMemcpyHostToDevice, c->s_copy));
            dim3 blk(128), grd((unsigned)ceil_div(ceil_div(c->W, 4), 128), (unsigned)(nf * c->H));
            me::bgr2gray_kernel<<<grd, blk, 0, c->s_copy>>>(c->d_bgr, row_bytes, dst, c->pitch, c->W, nf * c->H);
            CU_TRY(cudaGetLastError());
            c->launches += 1;
        }
        CU_TRY(cudaEventRecord(c->stage_free[sb], c->s_copy));
    }
    CU_TRY(cudaEventRecord(c->ev_upload, c->s_copy));
    c->have_upload = true;
    return ME_OK;
}

} // namespace

namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// OpenCV's almost_dist2weight table for one temporal window size (oracle/nlm_np.py: weight_table)
void nlm_weight_table(float h, int tws, int sws, int n_frames, std::vector<int> &table, int &shift)
{
    const long long max_est = (long long)n_frames * sws * sws * 255;
    const int fixed = (int)std::min<long long>(INT32_MAX / max_est, INT32_MAX);
    const int area = tws * tws;
    shift = 0;
    while ((1 << shift) < area) ++shift;
    const double mult = (double)(1 << shift) / area;
    const int n = (int)(65025 / mult + 1);
    const float hh = h * h;
    table.clear();
    int last = 0;
    for (int a = 0; a < n; ++a) {
        double w = std::exp(-(a * mult) / (double)hh);
        if (std::isnan(w)) w = 1.0;
        int wt = (int)std::nearbyint(fixed * w);                   // cvRound: round half to even
        if (wt < 0.001 * fixed) wt = 0;
        table.push_back(wt);
        if (wt) last = a;
    }
    table.resize((size_t)last + 1);
}

template <int TWS, int SH, bool LS> int nlm_launch(const me::NlmArgs &a, dim3 grid, size_t smem, cudaStream_t s)
{
    int rc = raise_smem_limit((const void *)me::nlm_kernel<TWS, SH, LS>, false);
    if (rc) return rc;
    me::nlm_kernel<TWS, SH, LS><<<grid, me::kNlmThreads, smem, s>>>(a);
    CU_TRY(cudaGetLastError());
    return ME_OK;
}

}  // namespace

extern "C" {

const char *me_last_error(void) { return g_err; }
const char *me_version(void) { return "me_b200 0.1 (sm_100a; TMA-staged VABSDIFF4 full search)"; }

int me_create(int device, int height, int width, const me_params *params, int max_frames, int max_pairs, me_ctx **out)
{
    if (!out) return fail(ME_EINVAL, "out is NULL");
    *out = nullptr;
    if (!params) return fail(ME_EINVAL, "params is NULL");
    const int B = params->block, R = params->range;
    if (B != 8 && B != 16 && B != 32) return fail(ME_EUNSUPPORTED, "block size %d: built sizes are 8, 16, 32", B);
    if (R < 1) return fail(ME_EINVAL, "range %d < 1", R);
    if (params->range_y < 0 || params->range_y > R) return fail(ME_EINVAL, "range_y %d outside 0..range (%d); 0 means the same as range", params->range_y, R);
    if (height < B || width < B) return fail(ME_EINVAL, "frame %dx%d smaller than one %d-pixel block", width, height, B);
    if (max_frames < 2 || max_pairs < 1) return fail(ME_EINVAL, "max_frames >= 2 and max_pairs >= 1 required");
    int ndev = 0;
    CU_TRY(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(ME_EINVAL, "device %d of %d", device, ndev);
    CU_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(ME_EUNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);

    me_ctx *c = new me_ctx();
    c->device = device; c->H = height; c->W = width; c->p = *params;
    c->max_frames = max_frames; c->max_pairs = max_pairs;
    c->sm_count = prop.multiProcessorCount;
    c->pitch = (size_t)round_up(width, 16);
    c->frame_stride = c->pitch * (size_t)height;
    {   // small frames: tiles small enough that a batch of a few pairs (the anchored scan's usual two or three) fills the GPU
        const int few = env_int("ME_PLAN_PAIRS", 2);
        c->plan_pairs = max_pairs > few ? few : max_pairs;
    }
    int rc = configure(c);
    if (rc) { delete c; return rc; }
    if (max_pairs > c->plan_pairs) {
        // ... and a second plan for launches of many pairs, where bigger tiles fill it as well (the 640x64 recordings: 2-block
        // tiles for 4 pairs, 8-block tiles for 64).  Kept only when it differs.
        me_ctx *b = new me_ctx();
        b->device = device; b->H = height; b->W = width; b->p = *params;
        b->max_frames = max_frames; b->max_pairs = max_pairs; b->sm_count = c->sm_count;
        b->pitch = c->pitch; b->frame_stride = c->frame_stride;
        b->plan_pairs = max_pairs;
        const me::SearchArgs &x = c->args, &y = b->args;
        if (configure(b) == ME_OK && (x.tmb != y.tmb || x.tmr != y.tmr || c->win_pitch != b->win_pitch || x.nsx != y.nsx || x.nsy != y.nsy))
            c->big = b;
        else
            delete b;
    }

#define CREATE_TRY(expr)                                                                                     \
    do {                                                                                                     \
        cudaError_t e_ = (expr);                                                                             \
        if (e_ != cudaSuccess) {                                                                             \
            fail(ME_ECUDA, "%s failed: %s", #expr, cudaGetErrorString(e_));                                  \
            me_destroy(c);                                                                                   \
            return ME_ECUDA;                                                                                 \
        }                                                                                                    \
    } while (0)

    const size_t ring = c->frame_stride * (size_t)max_frames + 256;
    CREATE_TRY(cudaMalloc(&c->d_frames, ring));
    CREATE_TRY(cudaMemset(c->d_frames, 0, ring));
    CREATE_TRY(cudaStreamCreateWithFlags(&c->s_copy, cudaStreamNonBlocking));
    CREATE_TRY(cudaStreamCreateWithFlags(&c->s_compute, cudaStreamNonBlocking));
    CREATE_TRY(cudaEventCreateWithFlags(&c->ev_upload, cudaEventDisableTiming));
    CREATE_TRY(cudaEventCreateWithFlags(&c->ev_search, cudaEventDisableTiming));
    CREATE_TRY(cudaEventCreate(&c->ev_k0));
    CREATE_TRY(cudaEventCreate(&c->ev_k1));
    CREATE_TRY(cudaEventCreate(&c->ev_t0));
    CREATE_TRY(cudaEventCreate(&c->ev_t1));
    CREATE_TRY(cudaEventCreate(&c->ev_c0));
    CREATE_TRY(cudaEventCreate(&c->ev_c1));
    CREATE_TRY(cudaEventCreateWithFlags(&c->ev_rec, cudaEventDisableTiming));
    for (auto &r : c->recent) CREATE_TRY(cudaEventCreateWithFlags(&r.done, cudaEventDisableTiming));
    // staging: two buffers, each >= one BGR frame and about 32 MiB for long uploads
    size_t want = (size_t)width * height * 3;
    if (want < (32u << 20)) want = 32u << 20;
    const size_t cap = (size_t)width * height * 3 * (size_t)max_frames;
    if (want > cap) want = cap;
    c->stage_bytes = want;
    for (int i = 0; i < 2; ++i) {
        CREATE_TRY(cudaHostAlloc(&c->h_stage[i], c->stage_bytes, cudaHostAllocDefault));
        CREATE_TRY(cudaEventCreateWithFlags(&c->stage_free[i], cudaEventDisableTiming));
    }
    CREATE_TRY(cudaStreamCreateWithFlags(&c->s_fetch, cudaStreamNonBlocking));
    for (auto &e : c->ev_slot) CREATE_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CREATE_TRY(cudaHostAlloc(&c->h_pairs, sizeof(int32_t) * 2 * (size_t)max_pairs * me_ctx::kSearchRing, cudaHostAllocDefault));
    CREATE_TRY(cudaHostAlloc(&c->h_pair_out, sizeof(me_pair_result) * ((size_t)max_pairs + 1), cudaHostAllocDefault));   // + the error word of the scan
    CREATE_TRY(cudaMalloc(&c->d_pairs, sizeof(int32_t) * 2 * (size_t)max_pairs * me_ctx::kSearchRing));
    const size_t nb = (size_t)c->args.nb * (size_t)max_pairs;
    CREATE_TRY(cudaMalloc(&c->args.blk_mv, nb * 2 * sizeof(int16_t)));
    CREATE_TRY(cudaMalloc(&c->args.blk_sad, nb * sizeof(uint32_t)));
    CREATE_TRY(cudaMalloc(&c->args.blk_sub, nb * 2 * sizeof(float)));
    CREATE_TRY(cudaMalloc(&c->args.blk_flags, nb));
    CREATE_TRY(cudaMalloc(&c->args.pair_out, sizeof(me_pair_result) * (size_t)max_pairs * me_ctx::kSearchRing));
    CREATE_TRY(cudaMalloc(&c->args.pair_ticket, sizeof(unsigned) * (size_t)max_pairs));
    CREATE_TRY(cudaMemset(c->args.pair_ticket, 0, sizeof(unsigned) * (size_t)max_pairs));
    const bool big_slabs = c->big && c->big->args.nsx * c->big->args.nsy > 1;
    if (c->args.nsx * c->args.nsy > 1 || big_slabs) {   // cross-slab argmin scratch: all-ones keys, zero tickets (self-restoring)
        const size_t nt = (size_t)std::max(c->args.tiles_per_pair, c->big ? c->big->args.tiles_per_pair : 0) * (size_t)max_pairs;
        CREATE_TRY(cudaMalloc(&c->args.blk_key, nb * sizeof(unsigned long long)));
        CREATE_TRY(cudaMemset(c->args.blk_key, 0xff, nb * sizeof(unsigned long long)));
        CREATE_TRY(cudaMalloc(&c->args.tile_ticket, nt * sizeof(unsigned)));
        CREATE_TRY(cudaMemset(c->args.tile_ticket, 0, nt * sizeof(unsigned)));
        c->n_tile_tickets = nt;
    }
    CREATE_TRY(cudaMalloc(&c->d_ready, sizeof(int) * (size_t)(max_pairs + 2)));
    CREATE_TRY(cudaMalloc(&c->d_error, sizeof(int)));
    CREATE_TRY(cudaMemset(c->d_error, 0, sizeof(int)));
    if (me_ctx *b = c->big) {                            // same outputs and scratch, its own kernels and launch shape
        b->args.blk_mv = c->args.blk_mv; b->args.blk_sad = c->args.blk_sad; b->args.blk_sub = c->args.blk_sub;
        b->args.blk_flags = c->args.blk_flags; b->args.pair_out = c->args.pair_out; b->args.pair_ticket = c->args.pair_ticket;
        b->args.blk_key = c->args.blk_key; b->args.tile_ticket = c->args.tile_ticket;
    }
#undef CREATE_TRY
    rc = make_tensor_map(&c->tm_win, c->d_frames, c->W, c->H, max_frames, c->pitch, c->win_pitch, c->args.box_rows);
    if (!rc) rc = make_tensor_map(&c->tm_ref, c->d_frames, c->W, c->H, max_frames, c->pitch, me::kRefPitch, B * c->args.tmr);
    if (!rc && c->big) {
        rc = make_tensor_map(&c->big->tm_win, c->d_frames, c->W, c->H, max_frames, c->pitch, c->big->win_pitch, c->big->args.box_rows);
        if (!rc) rc = make_tensor_map(&c->big->tm_ref, c->d_frames, c->W, c->H, max_frames, c->pitch, me::kRefPitch, B * c->big->args.tmr);
    }
    for (const me_ctx *pl : {static_cast<const me_ctx *>(c), static_cast<const me_ctx *>(c->big)}) {
        if (!pl || rc) continue;
        rc = raise_smem_limit(pl->kc.fn);
        if (!rc && pl->smem_persistent) rc = raise_smem_limit(pl->kc.fn_persistent);
    }
    if (rc) { me_destroy(c); return rc; }
    *out = c;
    return ME_OK;
}

int me_destroy(me_ctx *c)
{
    if (!c) return ME_OK;
    cudaSetDevice(c->device);
    if (c->s_copy) cudaStreamSynchronize(c->s_copy);
    if (c->s_compute) cudaStreamSynchronize(c->s_compute);
    cudaFree(c->d_frames); cudaFree(c->d_bgr); cudaFree(c->d_luts); cudaFree(c->d_nlm); cudaFree(c->d_nlm_meta); cudaFree(c->d_pairs);
    cudaFree(c->args.blk_mv); cudaFree(c->args.blk_sad); cudaFree(c->args.blk_sub); cudaFree(c->args.blk_flags);
    cudaFree(c->args.pair_out); cudaFree(c->args.pair_ticket); cudaFree(c->args.blk_key); cudaFree(c->args.tile_ticket);
    cudaFree(c->d_ready); cudaFree(c->d_error); cudaFree(c->d_scan);
    delete c->big;                                       // owns nothing: its pointers are the context's
    c->big = nullptr;
    for (int i = 0; i < 2; ++i) {
        if (c->h_stage[i]) cudaFreeHost(c->h_stage[i]);
        if (c->stage_free[i]) cudaEventDestroy(c->stage_free[i]);
    }
    if (c->h_pairs) cudaFreeHost(c->h_pairs);
    if (c->h_pair_out) cudaFreeHost(c->h_pair_out);
    if (c->ev_upload) cudaEventDestroy(c->ev_upload);
    if (c->ev_search) cudaEventDestroy(c->ev_search);
    for (auto &e : c->ev_slot) if (e) cudaEventDestroy(e);
    if (c->s_fetch) { cudaStreamSynchronize(c->s_fetch); cudaStreamDestroy(c->s_fetch); }
    if (c->ev_k0) cudaEventDestroy(c->ev_k0);
    if (c->ev_k1) cudaEventDestroy(c->ev_k1);
    if (c->ev_t0) cudaEventDestroy(c->ev_t0);
    if (c->ev_t1) cudaEventDestroy(c->ev_t1);
    if (c->ev_c0) cudaEventDestroy(c->ev_c0);
    if (c->ev_c1) cudaEventDestroy(c->ev_c1);
    if (c->ev_rec) cudaEventDestroy(c->ev_rec);
    for (auto &r : c->recent) if (r.done) cudaEventDestroy(r.done);
    if (c->s_copy) cudaStreamDestroy(c->s_copy);
    if (c->s_compute) cudaStreamDestroy(c->s_compute);
    delete c;
    return ME_OK;
}

int me_get_geometry(const me_ctx *c, me_geometry *g)
{
    if (!c || !g) return fail(ME_EINVAL, "NULL argument");
    g->height = c->H; g->width = c->W;
    g->nbx = c->args.nbx; g->nby = c->args.nby;
    g->max_frames = c->max_frames; g->max_pairs = c->max_pairs;
    g->tile_blocks_x = c->args.tmb * ((c->p.block < 16 ? 16 : c->p.block) / c->p.block); g->threads_per_cta = c->threads; g->smem_bytes = c->smem;
    g->tiles_per_pair = c->args.tiles_per_pair;
    g->cand_px_per_pair = c->cand_px; g->cand_px_per_pair_valid = c->cand_px_valid;
    g->sm_count = c->sm_count; g->uses_tma = c->args.use_tma;
    g->slabs_x = c->args.nsx; g->slabs_y = c->args.nsy; g->window_pitch = c->win_pitch; g->resident_kernel = c->smem_persistent ? 1 : 0; g->tile_blocks_y = c->args.tmr;
    return ME_OK;
}

int me_upload(me_ctx *c, const uint8_t *luma, size_t row_pitch, int first_slot, int n_frames)
{
    if (!c) return fail(ME_EINVAL, "ctx is NULL");
    return upload_impl(c, luma, row_pitch, first_slot, n_frames, false);
}

int me_upload_bgr(me_ctx *c, const uint8_t *bgr, size_t row_pitch, int first_slot, int n_frames)
{
    if (!c) return fail(ME_EINVAL, "ctx is NULL");
    return upload_impl(c, bgr, row_pitch, first_slot, n_frames, true);
}

int me_clahe(me_ctx *c, int first_slot, int n_frames, double clip_limit, int tiles_x, int tiles_y)
{
    if (!c) return fail(ME_EINVAL, "ctx is NULL");
    if (n_frames < 1 || first_slot < 0 || first_slot + n_frames > c->max_frames)
        return fail(ME_EINVAL, "%d frames at slot %d do not fit the %d-slot ring", n_frames, first_slot, c->max_frames);
    if (tiles_x < 1 || tiles_y < 1 || tiles_x > 64 || tiles_y > 64 || clip_limit < 0)
        return fail(ME_EINVAL, "tile grid %dx%d (1..64 each) / clip limit %g", tiles_x, tiles_y, clip_limit);
    CU_TRY(cudaSetDevice(c->device));
    me::ClaheArgs a;
    a.frames = c->d_frames; a.frame_stride = (long long)c->frame_stride; a.pitch = (int)c->pitch; a.W = c->W; a.H = c->H;
    a.tx = tiles_x; a.ty = tiles_y;
    int We = c->W, He = c->H;
    if (c->W % tiles_x || c->H % tiles_y) { We = c->W + tiles_x - c->W % tiles_x; He = c->H + tiles_y - c->H % tiles_y; }   // OpenCV's border rule
    a.tw = We / tiles_x; a.th = He / tiles_y;
    if (We - c->W >= c->W || He - c->H >= c->H) return fail(ME_EINVAL, "frame %dx%d too small for a %dx%d tile grid", c->W, c->H, tiles_x, tiles_y);
    const int area = a.tw * a.th;
    a.lut_scale = 255.0f / (float)area;
    a.clip = 0;
    if (clip_limit > 0) { a.clip = (int)(clip_limit * area / 256); if (a.clip < 1) a.clip = 1; }
    a.inv_tw = 1.0f / (float)a.tw; a.inv_th = 1.0f / (float)a.th;
    // Tables live only between the kernels of one chunk of frames: per frame 256 B per tile + 2 KB per interpolation cell.
    const int chunk = n_frames < 64 ? n_frames : 64;
    const size_t lut_bytes = (size_t)chunk * tiles_x * tiles_y * 256;
    const size_t need = align_up(lut_bytes, 256) + (size_t)chunk * (tiles_x + 1) * (tiles_y + 1) * 2048;
    if (c->d_luts_bytes < need) {
        CU_TRY(cudaStreamSynchronize(c->s_copy));
        if (c->d_luts) CU_TRY(cudaFree(c->d_luts));
        c->d_luts = nullptr; c->d_luts_bytes = 0;
        CU_TRY(cudaMalloc(&c->d_luts, need));
        c->d_luts_bytes = need;
    }
    a.luts = c->d_luts;
    uint2 *packed = reinterpret_cast<uint2 *>(c->d_luts + align_up(lut_bytes, 256));
    for (auto &r : c->recent)                        // the frames are rewritten in place: wait for searches that read them
        if (r.used && r.lo < first_slot + n_frames && r.hi >= first_slot) CU_TRY(cudaStreamWaitEvent(c->s_copy, r.done, 0));
    for (int f = 0; f < n_frames; f += chunk) {
        const int nf = n_frames - f < chunk ? n_frames - f : chunk;
        me::clahe_lut_kernel<<<dim3((unsigned)(tiles_x * tiles_y), (unsigned)nf), 256, 0, c->s_copy>>>(a, first_slot + f);
        CU_TRY(cudaGetLastError());
        me::clahe_pack_kernel<<<dim3((unsigned)((tiles_x + 1) * (tiles_y + 1)), (unsigned)nf), 256, 0, c->s_copy>>>(a, packed);
        CU_TRY(cudaGetLastError());
        me::clahe_apply_kernel<<<dim3((unsigned)ceil_div(c->H, me::kClaheRows), (unsigned)nf), 256, 0, c->s_copy>>>(a, first_slot + f, packed);
        CU_TRY(cudaGetLastError());
        c->launches += 3;
    }
    CU_TRY(cudaEventRecord(c->ev_upload, c->s_copy));   // same stream as the uploads: ordered after them, searches wait for this event
    c->have_upload = true;
    return ME_OK;
}

int me_denoise(me_ctx *c, int first_slot, int n_frames, const int32_t *temporal_window, float h, int template_window, int search_window)
{
    if (!c || !temporal_window) return fail(ME_EINVAL, "NULL argument");
    if (n_frames < 1 || first_slot < 0 || first_slot + n_frames > c->max_frames)
        return fail(ME_EINVAL, "%d frames at slot %d do not fit the %d-slot ring", n_frames, first_slot, c->max_frames);
    const int tws = 2 * (template_window / 2) + 1, sws = 2 * (search_window / 2) + 1, th = tws / 2, sh = sws / 2;
    if (template_window < 1 || search_window < 1 || !(h >= 0)) return fail(ME_EINVAL, "template %d / search %d / h %g", template_window, search_window, (double)h);
    if (tws != 7 && tws != 19) return fail(ME_EUNSUPPORTED, "template window %d: 7 and 19 are built", template_window);
    if (sws > 41) return fail(ME_EUNSUPPORTED, "search window %d: up to 41 is built", search_window);
    for (int k = 0; k < n_frames; ++k) {
        const int t = temporal_window[k];
        if (t == 0) continue;
        if (t > 1 && (template_window % 2 == 0 || search_window % 2 == 0))
            return fail(ME_EINVAL, "frame %d: temporal windows need odd template / search windows (%d / %d), as in OpenCV", k, template_window, search_window);
        if (t < 1 || t % 2 == 0 || k - t / 2 < 0 || k + t / 2 >= n_frames)
            return fail(ME_EINVAL, "frame %d: temporal window %d must be 0 or odd and stay inside the %d frames of the call", k, t, n_frames);
    }
    CU_TRY(cudaSetDevice(c->device));

    // Weight tables, one per distinct temporal window, + the per-frame table index: [twin | lut_off | lut_len | luts] in a
    // device buffer of their own.  The same call as last time (a recording is denoised chunk after chunk with the same
    // windows) finds them there: no rebuild (46k exp() per table), no upload, no stream synchronisation.
    std::vector<int> key;
    key.push_back(n_frames); key.push_back(tws); key.push_back(sws);
    { int hb; memcpy(&hb, &h, sizeof hb); key.push_back(hb); }
    key.insert(key.end(), temporal_window, temporal_window + n_frames);
    int shift = c->nlm_meta_shift;
    if (key != c->nlm_meta) {
        std::vector<int> luts, off(n_frames), len(n_frames);
        std::vector<std::pair<int, std::pair<int, int>>> seen;
        for (int k = 0; k < n_frames; ++k) {
            const int t = temporal_window[k];
            if (t == 0) { off[k] = len[k] = 0; continue; }
            bool found = false;
            for (auto &e : seen) if (e.first == t) { off[k] = e.second.first; len[k] = e.second.second; found = true; }
            if (found) continue;
            std::vector<int> tab;
            nlm_weight_table(h, tws, sws, t, tab, shift);
            off[k] = (int)luts.size(); len[k] = (int)tab.size();
            seen.push_back({t, {off[k], len[k]}});
            luts.insert(luts.end(), tab.begin(), tab.end());
            luts.push_back(0);                       // entry [len]: the kernel clamps the table index instead of testing it
        }
        std::vector<int> meta;
        meta.insert(meta.end(), temporal_window, temporal_window + n_frames);
        meta.insert(meta.end(), off.begin(), off.end());
        meta.insert(meta.end(), len.begin(), len.end());
        meta.insert(meta.end(), luts.begin(), luts.end());
        const size_t meta_bytes = meta.size() * sizeof(int);
        CU_TRY(cudaStreamSynchronize(c->s_copy));    // an earlier call may still read the old tables; meta is pageable host memory
        if (c->d_nlm_meta_bytes < meta_bytes) {
            if (c->d_nlm_meta) CU_TRY(cudaFree(c->d_nlm_meta));
            c->d_nlm_meta = nullptr; c->d_nlm_meta_bytes = 0;
            CU_TRY(cudaMalloc(&c->d_nlm_meta, meta_bytes));
            c->d_nlm_meta_bytes = meta_bytes;
        }
        CU_TRY(cudaMemcpy(c->d_nlm_meta, meta.data(), meta_bytes, cudaMemcpyHostToDevice));
        c->nlm_meta = key;
        c->nlm_meta_shift = shift;
        c->nlm_lut_small = true;                     // every table short enough and every weight 16-bit: shared-memory form usable
        for (int k = 0; k < n_frames; ++k) if (len[k] + 1 > me::kNlmLutSmem) c->nlm_lut_small = false;
        for (int w : luts) if (w > 65535) c->nlm_lut_small = false;
    }

    me::NlmArgs a{};
    a.src = c->d_frames; a.frame_stride = (long long)c->frame_stride; a.pitch = (int)c->pitch; a.W = c->W; a.H = c->H;
    a.first_slot = first_slot; a.sh = sh; a.shift = shift;
    const int b = sh + th;
    a.itp = me::nlm_region_pitch(sh, th);
    a.it_rows = me::kNlmTR + 2 * b;
    const int nr = me::kNlmTR + 2 * th;
    const size_t smem = 2 * (size_t)nr * (me::kNlmTW + 4) * 4 + (size_t)a.it_rows * a.itp + 16;   // two buffers of horizontal sums + the region
    const int tiles = ceil_div(c->W, me::kNlmTW) * ceil_div(c->H, me::kNlmTR);
    a.nsplit = 1;
    if ((long long)tiles * n_frames < 2ll * c->sm_count) a.nsplit = std::min(sws, (int)ceil_div(2 * c->sm_count, tiles * n_frames));
    a.nsplit = env_int("ME_NLM_SPLIT", a.nsplit);
    if (a.nsplit < 1) a.nsplit = 1;

    // device scratch: [dst frames | acc]
    const size_t dst_bytes = (size_t)n_frames * c->H * c->pitch;
    const size_t acc_bytes = a.nsplit > 1 ? (size_t)n_frames * c->H * c->W * 8 : 0;
    const size_t need = align_up(dst_bytes, 256) + align_up(acc_bytes, 256);
    if (c->d_nlm_bytes < need) {
        CU_TRY(cudaStreamSynchronize(c->s_copy));
        if (c->d_nlm) CU_TRY(cudaFree(c->d_nlm));
        c->d_nlm = nullptr; c->d_nlm_bytes = 0;
        CU_TRY(cudaMalloc(&c->d_nlm, need));
        c->d_nlm_bytes = need;
    }
    a.dst = c->d_nlm;
    a.acc = reinterpret_cast<int *>(c->d_nlm + align_up(dst_bytes, 256));
    const int *d_meta = c->d_nlm_meta;
    a.twin = d_meta; a.lut_off = d_meta + n_frames; a.lut_len = d_meta + 2 * n_frames; a.luts = d_meta + 3 * n_frames;

    for (auto &r : c->recent)                        // the frames are rewritten in place: wait for searches that read them
        if (r.used && r.lo < first_slot + n_frames && r.hi >= first_slot) CU_TRY(cudaStreamWaitEvent(c->s_copy, r.done, 0));
    if (a.nsplit > 1) CU_TRY(cudaMemsetAsync(a.acc, 0, acc_bytes, c->s_copy));
    const dim3 grid((unsigned)ceil_div(c->W, me::kNlmTW), (unsigned)ceil_div(c->H, me::kNlmTR), (unsigned)(n_frames * a.nsplit));
    // the reference's windows (19 / 41) with a table that fits 16-bit shared-memory entries take the specialised instantiation
    const bool lut_smem = tws == 19 && sh == 20 && c->nlm_lut_small;
    int rc = tws == 19 ? (lut_smem ? nlm_launch<19, 20, true>(a, grid, smem + 2 * me::kNlmLutSmem, c->s_copy)
                                   : nlm_launch<19, 0, false>(a, grid, smem, c->s_copy))
                       : nlm_launch<7, 0, false>(a, grid, smem, c->s_copy);
    if (rc) return rc;
    c->launches += 1;
    if (a.nsplit > 1) {
        me::NlmArgs f = a;
        const long long n_px = (long long)n_frames * c->H * c->W;
        me::nlm_finish_kernel<<<(unsigned)((n_px + 255) / 256), 256, 0, c->s_copy>>>(f, n_px);
        CU_TRY(cudaGetLastError());
        c->launches += 1;
    }
    CU_TRY(copy_rows(c->d_frames + (size_t)first_slot * c->frame_stride, c->pitch, c->d_nlm, c->pitch, (size_t)c->W,
                     (size_t)n_frames * c->H, cudaMemcpyDeviceToDevice, c->s_copy));
    CU_TRY(cudaEventRecord(c->ev_upload, c->s_copy));
    c->have_upload = true;
    return ME_OK;
}

int me_download(me_ctx *c, uint8_t *luma, size_t row_pitch, int first_slot, int n_frames)
{
    if (!c || !luma) return fail(ME_EINVAL, "NULL argument");
    if (n_frames < 1 || first_slot < 0 || first_slot + n_frames > c->max_frames)
        return fail(ME_EINVAL, "%d frames at slot %d do not fit the %d-slot ring", n_frames, first_slot, c->max_frames);
    if (row_pitch < (size_t)c->W) return fail(ME_EINVAL, "row_pitch %zu < width %d", row_pitch, c->W);
    CU_TRY(cudaSetDevice(c->device));
    CU_TRY(cudaStreamSynchronize(c->s_copy));
    cudaPointerAttributes pa;
    const bool pinned = cudaPointerGetAttributes(&pa, luma) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    if (!pinned) (void)cudaGetLastError();
    const size_t frame_bytes = (size_t)c->W * c->H;
    const int per_chunk = (int)(c->stage_bytes / frame_bytes);
    if (pinned || per_chunk < 1) {                   // a page-locked destination takes the copy directly
        CU_TRY(copy_rows(luma, row_pitch, c->d_frames + (size_t)first_slot * c->frame_stride, c->pitch, (size_t)c->W,
                         (size_t)n_frames * c->H, cudaMemcpyDeviceToHost, c->s_fetch));
        CU_TRY(cudaStreamSynchronize(c->s_fetch));
        return ME_OK;
    }
    // Pageable destination: device -> pinned staging (async, the fetch stream) while the chunk before it is copied from the
    // other staging buffer into the caller's memory -- the mirror image of me_upload (a direct cudaMemcpy2D into pageable
    // memory ran at a third of the link's rate: r02 bench, CLAHE end to end).
    for (int sb = 0; sb < 2; ++sb) CU_TRY(cudaEventSynchronize(c->stage_free[sb]));      // no upload still reads the staging buffers
    int pend_f = -1, pend_n = 0, pend_sb = 0;
    auto drain = [&](int f, int nf, int sb) -> int {
        CU_TRY(cudaEventSynchronize(c->stage_free[sb]));
        uint8_t *dst = luma + (size_t)f * c->H * row_pitch;
        if (row_pitch == (size_t)c->W) memcpy(dst, c->h_stage[sb], frame_bytes * nf);
        else for (size_t r = 0; r < (size_t)nf * c->H; ++r) memcpy(dst + r * row_pitch, c->h_stage[sb] + r * c->W, (size_t)c->W);
        return ME_OK;
    };
    for (int f = 0; f < n_frames; f += per_chunk) {
        const int nf = n_frames - f < per_chunk ? n_frames - f : per_chunk;
        const int sb = c->stage_next;
        c->stage_next ^= 1;
        CU_TRY(copy_rows(c->h_stage[sb], (size_t)c->W, c->d_frames + (size_t)(first_slot + f) * c->frame_stride, c->pitch,
                         (size_t)c->W, (size_t)nf * c->H, cudaMemcpyDeviceToHost, c->s_fetch));
        CU_TRY(cudaEventRecord(c->stage_free[sb], c->s_fetch));
        if (pend_f >= 0) { int rc = drain(pend_f, pend_n, pend_sb); if (rc) return rc; }
        pend_f = f; pend_n = nf; pend_sb = sb;
    }
    return drain(pend_f, pend_n, pend_sb);
}

int me_copy_frames_device(me_ctx *c, void *d_dst, size_t dst_pitch, int first_slot, int n_frames, void *stream)
{
    if (!c || !d_dst) return fail(ME_EINVAL, "NULL argument");
    if (n_frames < 1 || first_slot < 0 || first_slot + n_frames > c->max_frames)
        return fail(ME_EINVAL, "%d frames at slot %d do not fit the %d-slot ring", n_frames, first_slot, c->max_frames);
    if (dst_pitch < (size_t)c->W) return fail(ME_EINVAL, "dst_pitch %zu < width %d", dst_pitch, c->W);
    CU_TRY(cudaSetDevice(c->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : c->s_copy;
    if (st != c->s_copy && c->have_upload) CU_TRY(cudaStreamWaitEvent(st, c->ev_upload, 0));
    CU_TRY(copy_rows(d_dst, dst_pitch, c->d_frames + (size_t)first_slot * c->frame_stride, c->pitch, (size_t)c->W,
                     (size_t)n_frames * c->H, cudaMemcpyDeviceToDevice, st));
    return ME_OK;
}

int me_search(me_ctx *c, const int32_t *pairs, int n_pairs)
{
    if (!c) return fail(ME_EINVAL, "ctx is NULL");
    int rc = check_pairs(c, pairs, n_pairs, c->max_frames);
    if (rc) return rc;
    CU_TRY(cudaSetDevice(c->device));
    const int slot = (c->slot + 1) % me_ctx::kSearchRing;
    if (c->slot_used[slot]) CU_TRY(cudaEventSynchronize(c->ev_slot[slot]));   // the search four launches ago: its pair list is about to be rewritten
    int32_t *hp = c->h_pairs + (size_t)slot * 2 * c->max_pairs, *dp = c->d_pairs + (size_t)slot * 2 * c->max_pairs;
    memcpy(hp, pairs, sizeof(int32_t) * 2 * (size_t)n_pairs);
    CU_TRY(cudaMemcpyAsync(dp, hp, sizeof(int32_t) * 2 * (size_t)n_pairs, cudaMemcpyHostToDevice, c->s_compute));
    if (c->have_upload) CU_TRY(cudaStreamWaitEvent(c->s_compute, c->ev_upload, 0));
    const me_ctx *pl = pick_plan(c, n_pairs);
    rc = launch_search(c, pl, c->d_frames, pl->tm_win, pl->tm_ref, n_pairs, c->s_compute, 0, slot);
    if (rc) return rc;
    CU_TRY(cudaEventRecord(c->ev_slot[slot], c->s_compute));
    c->slot = slot; c->slot_used[slot] = true; c->slot_pairs[slot] = n_pairs;
    CU_TRY(cudaEventRecord(c->ev_search, c->s_compute));
    {
        me_ctx::SearchRecord &r = c->recent[c->recent_next];
        c->recent_next = (c->recent_next + 1) % 4;
        r.lo = pairs[0]; r.hi = pairs[0];
        for (int i = 0; i < 2 * n_pairs; ++i) { if (pairs[i] < r.lo) r.lo = pairs[i]; if (pairs[i] > r.hi) r.hi = pairs[i]; }
        CU_TRY(cudaEventRecord(r.done, c->s_compute));
        r.used = true;
    }
    c->have_search = true;
    c->last_pairs = n_pairs;
    return ME_OK;
}

int me_search_device(me_ctx *c, const uint8_t *d_frames, size_t pitch, int n_frames, const int32_t *pairs, int n_pairs,
                     void *stream)
{
    if (!c) return fail(ME_EINVAL, "ctx is NULL");
    if (!d_frames || (reinterpret_cast<uintptr_t>(d_frames) & 15) || pitch % 16 || pitch < (size_t)c->W)
        return fail(ME_EINVAL, "device frames need a 16-byte aligned base and a pitch >= width that is a multiple of 16");
    int rc = check_pairs(c, pairs, n_pairs, n_frames);
    if (rc) return rc;
    CU_TRY(cudaSetDevice(c->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : c->s_compute;
    const me_ctx *pl = pick_plan(c, n_pairs);
    CUtensorMap tw, tr;
    rc = make_tensor_map(&tw, d_frames, c->W, c->H, n_frames, pitch, pl->win_pitch, pl->args.box_rows);
    if (!rc) rc = make_tensor_map(&tr, d_frames, c->W, c->H, n_frames, pitch, me::kRefPitch, c->p.block * pl->args.tmr);
    if (rc) return rc;
    const int slot = (c->slot + 1) % me_ctx::kSearchRing;
    if (c->slot_used[slot]) CU_TRY(cudaEventSynchronize(c->ev_slot[slot]));
    int32_t *hp = c->h_pairs + (size_t)slot * 2 * c->max_pairs, *dp = c->d_pairs + (size_t)slot * 2 * c->max_pairs;
    memcpy(hp, pairs, sizeof(int32_t) * 2 * (size_t)n_pairs);
    CU_TRY(cudaMemcpyAsync(dp, hp, sizeof(int32_t) * 2 * (size_t)n_pairs, cudaMemcpyHostToDevice, st));
    rc = launch_search(c, pl, d_frames, tw, tr, n_pairs, st, pitch, slot);
    if (rc) return rc;
    CU_TRY(cudaEventRecord(c->ev_slot[slot], st));
    c->slot = slot; c->slot_used[slot] = true; c->slot_pairs[slot] = n_pairs;
    CU_TRY(cudaEventRecord(c->ev_search, st));
    if (st != c->s_compute) CU_TRY(cudaStreamWaitEvent(c->s_compute, c->ev_search, 0));
    c->have_search = true;
    c->last_pairs = n_pairs;
    return ME_OK;
}

int me_last_ticket(const me_ctx *c) { return c ? c->slot : 0; }

int me_fetch_pairs_ticket(me_ctx *c, int ticket, me_pair_result *out, int n_pairs)
{
    if (!c || !out) return fail(ME_EINVAL, "NULL argument");
    if (!c->have_search) return fail(ME_ESTATE, "me_fetch_pairs before any me_search");
    if (ticket < 0 || ticket >= me_ctx::kSearchRing || !c->slot_used[ticket])
        return fail(ME_EINVAL, "ticket %d does not name one of the last %d searches", ticket, me_ctx::kSearchRing);
    if (n_pairs < 1 || n_pairs > c->slot_pairs[ticket]) return fail(ME_EINVAL, "n_pairs=%d, that search had %d", n_pairs, c->slot_pairs[ticket]);
    CU_TRY(cudaSetDevice(c->device));
    // on a stream of its own: a fetch queued on the compute stream would wait for the searches launched after the one it wants
    int *h_err = reinterpret_cast<int *>(c->h_pair_out + c->max_pairs);
    CU_TRY(cudaStreamWaitEvent(c->s_fetch, c->ev_slot[ticket], 0));
    CU_TRY(cudaMemcpyAsync(c->h_pair_out, c->args.pair_out + (size_t)ticket * c->max_pairs, sizeof(me_pair_result) * (size_t)n_pairs,
                           cudaMemcpyDeviceToHost, c->s_fetch));
    CU_TRY(cudaMemcpyAsync(h_err, c->d_error, sizeof(int), cudaMemcpyDeviceToHost, c->s_fetch));
    CU_TRY(cudaStreamSynchronize(c->s_fetch));
    if (*h_err) return report_search_timeout(c);
    memcpy(out, c->h_pair_out, sizeof(me_pair_result) * (size_t)n_pairs);
    return ME_OK;
}

int me_fetch_pairs(me_ctx *c, me_pair_result *out, int n_pairs)
{
    if (!c) return fail(ME_EINVAL, "NULL argument");
    return me_fetch_pairs_ticket(c, c->slot, out, n_pairs);
}

int me_fetch_blocks(me_ctx *c, int pair_index, int16_t *mv, uint32_t *sad, float *sub, uint8_t *flags)
{
    if (!c) return fail(ME_EINVAL, "ctx is NULL");
    if (!c->have_search) return fail(ME_ESTATE, "me_fetch_blocks before any me_search");
    if (pair_index < 0 || pair_index >= c->last_pairs) return fail(ME_EINVAL, "pair_index %d, last search had %d pairs", pair_index, c->last_pairs);
    CU_TRY(cudaSetDevice(c->device));
    const size_t nb = (size_t)c->args.nb, o = nb * (size_t)pair_index;
    cudaStream_t st = c->s_compute;
    if (mv) CU_TRY(cudaMemcpyAsync(mv, c->args.blk_mv + o * 2, nb * 2 * sizeof(int16_t), cudaMemcpyDeviceToHost, st));
    if (sad) CU_TRY(cudaMemcpyAsync(sad, c->args.blk_sad + o, nb * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    if (sub) CU_TRY(cudaMemcpyAsync(sub, c->args.blk_sub + o * 2, nb * 2 * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (flags) CU_TRY(cudaMemcpyAsync(flags, c->args.blk_flags + o, nb, cudaMemcpyDeviceToHost, st));
    int err = 0;
    CU_TRY(cudaMemcpyAsync(&err, c->d_error, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    if (err) return report_search_timeout(c);
    return ME_OK;
}

int me_estimate_pair(me_ctx *c, const uint8_t *prev, const uint8_t *cur, size_t row_pitch, me_pair_result *out)
{
    if (!c || !prev || !cur || !out) return fail(ME_EINVAL, "NULL argument");
    int rc = upload_impl(c, prev, row_pitch, 0, 1, false);
    if (!rc) rc = upload_impl(c, cur, row_pitch, 1, 1, false);
    const int32_t pr[2] = {0, 1};
    if (!rc) rc = me_search(c, pr, 1);
    if (!rc) rc = me_fetch_pairs(c, out, 1);
    return rc;
}

} // extern "C"

namespace {

// The whole walk as one cooperative launch of sad_scan_kernel (sad_scan.cuh).  Returns ME_OK with *done = false when this
// context's plan cannot use it (64-bit keys, multi-row tiles) -- the caller then drives the walk from the host.
int anchored_scan_device(me_ctx *c, int anchor_slot, int first_slot, int n_frames, float scale, float gate,
                         me_scan_record *out, int *searches, int *launches, bool *done)
{
    *done = false;
    if (!c->kc.fn_scan || c->args.tmr != 1 || env_int("ME_SCAN_DEVICE", 1) == 0) return ME_OK;
    if (c->max_frames >= (1 << me::kScanSlotBits)) return ME_OK;      // frame slots travel in 20-bit fields of the barrier word
    int coop = 0;
    CU_TRY(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device));
    if (!coop) return ME_OK;
    cudaStream_t st = c->s_compute;
    const size_t rec_bytes = align_up(sizeof(me_scan_record) * (size_t)c->max_frames, 256);
    if (!c->d_scan) CU_TRY(cudaMalloc(&c->d_scan, rec_bytes + 256));
    int rc = raise_smem_limit(c->kc.fn_scan);
    if (rc) return rc;
    int per_sm = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, c->kc.fn_scan, c->threads, (size_t)c->smem));
    if (per_sm < 1) return ME_OK;
    me::SearchArgs a = c->args;
    a.frames = c->d_frames;
    a.pairs = c->d_pairs;
    me::ScanArgs sc;
    sc.first_slot = first_slot; sc.n_frames = n_frames; sc.anchor_slot = anchor_slot;
    sc.scale = scale; sc.gate = gate;
    sc.max_look = c->max_pairs < 256 ? c->max_pairs : 256;
    sc.out = reinterpret_cast<me_scan_record *>(c->d_scan);
    sc.pairs = c->d_pairs;
    sc.state = reinterpret_cast<int *>(c->d_scan + rec_bytes);
    sc.sync = reinterpret_cast<unsigned *>(c->d_scan + rec_bytes + 64);
    sc.trace = env_int("ME_SCAN_TRACE", 0) ? reinterpret_cast<unsigned long long *>(c->d_scan + rec_bytes + 96) : nullptr;
    sc.trace_cta = env_int("ME_SCAN_TRACE", 0) - 1;
    // every CTA of the grid arrives at every barrier, with or without work: no more CTAs per SM than ME_SCAN_CTAS (tuned: see DESIGN.md 4)
    const int cap = env_int("ME_SCAN_CTAS", 4);
    if (cap > 0 && per_sm > cap) per_sm = cap;
    long long grid = (long long)per_sm * c->sm_count;
    // a batch that fits the grid is one unit per CTA; one pair more and a quarter of the CTAs run two while the rest wait (r02 trace,
    // 640x64: 553 units on 444 CTAs), so the look-ahead stops where the grid is full -- when that still leaves a batch of two or more
    const long long per_pair_units = (long long)a.tiles_per_pair * a.nsx * a.nsy;
    if (env_int("ME_SCAN_FIT", 1) && grid / per_pair_units >= 2 && grid / per_pair_units < sc.max_look) sc.max_look = (int)(grid / per_pair_units);
    const long long most = per_pair_units * sc.max_look;                                   // units of the largest batch
    if (grid > most) grid = most;
    CU_TRY(cudaMemsetAsync(c->d_scan + rec_bytes, 0, 256, st));
    void *params[4] = {(void *)&c->tm_win, (void *)&c->tm_ref, (void *)&a, (void *)&sc};
    CU_TRY(cudaEventRecord(c->ev_k0, st));
    CU_TRY(cudaLaunchCooperativeKernel(c->kc.fn_scan, dim3((unsigned)grid), dim3((unsigned)c->threads), params, (size_t)c->smem, st));
    CU_TRY(cudaEventRecord(c->ev_k1, st));
    c->launches += 1;
    int state[8];
    CU_TRY(cudaMemcpyAsync(state, sc.state, sizeof state, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaMemcpyAsync(out, sc.out, sizeof(me_scan_record) * (size_t)n_frames, cudaMemcpyDeviceToHost, st));
    CU_TRY(cudaStreamSynchronize(st));
    c->have_search = true;
    c->last_pairs = 1;
    if (state[7]) {
        report_search_timeout(c);
        return fail(ME_ECUDA, "the resident scan kernel gave up waiting at a grid barrier; the context has been reset");
    }
    if (sc.trace) {
        unsigned long long tr[20];
        CU_TRY(cudaMemcpy(tr, sc.trace, sizeof tr, cudaMemcpyDeviceToHost));
        const double n = state[6] ? state[6] : 1;
        fprintf(stderr, "[me_b200] scan trace: %d steps (mean batch %.2f pairs), per step: CTA 0 %.2f us in search_tile + %.2f us at the barrier; "
                "step start -> slowest unit has written its records %.2f us -> last CTA arrives %.2f us, the gate %.2f us (grid %lld x %d threads, %d units per pair)\n",
                state[6], tr[5] / n, tr[0] * 1e-3 / n, tr[1] * 1e-3 / n, tr[7] * 1e-3 / n, tr[2] * 1e-3 / n, tr[3] * 1e-3 / n, grid, c->threads,
                a.tiles_per_pair * a.nsx * a.nsy);
        fprintf(stderr, "[me_b200] scan trace: CTA %d, SM-clock cycles per step: load %.0f, shifted copies %.0f, candidates %.0f, finalize %.0f, "
                "ticket %.0f; pair reduction %.0f cycles each (%llu times)\n", sc.trace_cta, tr[8] / n, tr[9] / n, tr[10] / n, tr[11] / n,
                tr[12] / n, tr[14] ? (double)tr[13] / (double)tr[14] : 0.0, tr[14]);
        if (tr[14])
            fprintf(stderr, "[me_b200] scan trace: inside a pair reduction: clear + key loads %.0f, sums %.0f, counting + selection %.0f, record %.0f cycles\n",
                    (double)tr[15] / tr[14], (double)tr[16] / tr[14], (double)tr[17] / tr[14], (double)tr[18] / tr[14]);
    }
    if (searches) *searches = state[5];
    if (launches) *launches = state[6];
    *done = true;
    return ME_OK;
}

} // namespace

extern "C" {

int me_anchored_scan(me_ctx *c, int anchor_slot, int first_slot, int n_frames, float scale, float gate,
                     me_scan_record *out, int *searches, int *launches)
{
    if (!c || !out) return fail(ME_EINVAL, "NULL argument");
    if (n_frames < 1 || first_slot < 0 || first_slot + n_frames > c->max_frames || anchor_slot < 0 || anchor_slot >= c->max_frames)
        return fail(ME_EINVAL, "scan of %d frames at slot %d with the anchor in slot %d does not fit the %d-slot ring", n_frames,
                    first_slot, anchor_slot, c->max_frames);
    CU_TRY(cudaSetDevice(c->device));
    if (c->have_search) CU_TRY(cudaEventSynchronize(c->ev_search));   // h_pairs / h_pair_out are about to be reused
    if (c->have_upload) CU_TRY(cudaStreamWaitEvent(c->s_compute, c->ev_upload, 0));
    cudaStream_t st = c->s_compute;
    {
        bool done = false;
        int rc = anchored_scan_device(c, anchor_slot, first_slot, n_frames, scale, gate, out, searches, launches, &done);
        if (rc) return rc;
        if (done) {
            c->slot = 0; c->slot_used[0] = true; c->slot_pairs[0] = c->last_pairs;      // the walk uses ring slot 0
            CU_TRY(cudaEventRecord(c->ev_slot[0], st));
            CU_TRY(cudaEventRecord(c->ev_search, st));
            me_ctx::SearchRecord &r = c->recent[c->recent_next];
            c->recent_next = (c->recent_next + 1) % 4;
            r.lo = anchor_slot < first_slot ? anchor_slot : first_slot;
            r.hi = anchor_slot > first_slot + n_frames - 1 ? anchor_slot : first_slot + n_frames - 1;
            CU_TRY(cudaEventRecord(r.done, st));
            r.used = true;
            return ME_OK;
        }
    }
    int *h_err = reinterpret_cast<int *>(c->h_pair_out + c->max_pairs);
    const int end = first_slot + n_frames, max_look = c->max_pairs < 256 ? c->max_pairs : 256;
    int anchor = anchor_slot, k = first_slot, look = max_look < 4 ? max_look : 4, since = 0, n_search = 0, n_launch = 0, last = 0;
    int lo = anchor_slot < first_slot ? anchor_slot : first_slot, hi = anchor_slot > end - 1 ? anchor_slot : end - 1;
    while (k < end) {
        const int L = std::min(std::min(look, end - k), c->max_pairs);
        for (int i = 0; i < L; ++i) { c->h_pairs[2 * i] = anchor; c->h_pairs[2 * i + 1] = k + i; }
        CU_TRY(cudaMemcpyAsync(c->d_pairs, c->h_pairs, sizeof(int32_t) * 2 * (size_t)L, cudaMemcpyHostToDevice, st));
        const me_ctx *pl = pick_plan(c, L);
        int rc = launch_search(c, pl, c->d_frames, pl->tm_win, pl->tm_ref, L, st);
        if (rc) return rc;
        CU_TRY(cudaMemcpyAsync(c->h_pair_out, c->args.pair_out, sizeof(me_pair_result) * (size_t)L, cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaMemcpyAsync(h_err, c->d_error, sizeof(int), cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        c->have_search = true;
        c->last_pairs = last = L;
        if (*h_err) return report_search_timeout(c);
        ++n_launch;
        n_search += L;
        int acc = -1;
        for (int i = 0; i < L; ++i) {
            const me_pair_result &r = c->h_pair_out[i];
            const bool skip = (r.pdx * scale < gate) && r.flag == 1;          // estimate_vel_vid.py:531-534
            me_scan_record &o = out[k + i - first_slot];
            o.anchor_slot = anchor; o.accepted = skip ? 0 : 1; o.pair = r;
            if (!skip) { acc = i; break; }                                     // :575 -- the rest of the batch compared with a stale anchor
        }
        if (acc >= 0) {
            const int dist = since + acc + 1;                                  // frames from the old anchor to the new one
            anchor = k + acc; k += acc + 1; since = 0;
            look = std::min(std::max(dist + 1, 2), max_look);                  // expect the next accept about as far away
        } else {
            since += L; k += L;
            look = std::min(2 * look, max_look);                               // same anchor: look further
        }
    }
    c->slot = 0; c->slot_used[0] = true; c->slot_pairs[0] = c->last_pairs;              // the walk uses ring slot 0
    CU_TRY(cudaEventRecord(c->ev_slot[0], st));
    CU_TRY(cudaEventRecord(c->ev_search, st));
    {
        me_ctx::SearchRecord &r = c->recent[c->recent_next];
        c->recent_next = (c->recent_next + 1) % 4;
        r.lo = lo; r.hi = hi;
        CU_TRY(cudaEventRecord(r.done, st));
        r.used = true;
    }
    (void)last;
    if (searches) *searches = n_search;
    if (launches) *launches = n_launch;
    return ME_OK;
}

int me_sync(me_ctx *c)
{
    if (!c) return fail(ME_EINVAL, "ctx is NULL");
    CU_TRY(cudaSetDevice(c->device));
    CU_TRY(cudaStreamSynchronize(c->s_copy));
    CU_TRY(cudaStreamSynchronize(c->s_compute));
    return ME_OK;
}

int me_last_search_ms(me_ctx *c, float *ms)
{
    if (!c || !ms) return fail(ME_EINVAL, "NULL argument");
    if (!c->have_search) return fail(ME_ESTATE, "no search has been launched");
    CU_TRY(cudaSetDevice(c->device));
    CU_TRY(cudaEventSynchronize(c->ev_k1));
    CU_TRY(cudaEventElapsedTime(ms, c->ev_k0, c->ev_k1));
    return ME_OK;
}

int64_t me_launch_count(const me_ctx *c) { return c ? c->launches : 0; }

int me_timer_start(me_ctx *c)
{
    if (!c) return fail(ME_EINVAL, "ctx is NULL");
    CU_TRY(cudaSetDevice(c->device));
    CU_TRY(cudaEventRecord(c->ev_t0, c->s_compute));
    return ME_OK;
}

int me_timer_stop(me_ctx *c, float *ms)
{
    if (!c || !ms) return fail(ME_EINVAL, "NULL argument");
    CU_TRY(cudaSetDevice(c->device));
    CU_TRY(cudaEventRecord(c->ev_t1, c->s_compute));
    CU_TRY(cudaEventSynchronize(c->ev_t1));
    CU_TRY(cudaEventElapsedTime(ms, c->ev_t0, c->ev_t1));
    return ME_OK;
}

int me_copy_timer_start(me_ctx *c)
{
    if (!c) return fail(ME_EINVAL, "ctx is NULL");
    CU_TRY(cudaSetDevice(c->device));
    CU_TRY(cudaEventRecord(c->ev_c0, c->s_copy));
    return ME_OK;
}

int me_copy_timer_stop(me_ctx *c, float *ms)
{
    if (!c || !ms) return fail(ME_EINVAL, "NULL argument");
    CU_TRY(cudaSetDevice(c->device));
    CU_TRY(cudaEventRecord(c->ev_c1, c->s_copy));
    CU_TRY(cudaEventSynchronize(c->ev_c1));
    CU_TRY(cudaEventElapsedTime(ms, c->ev_c0, c->ev_c1));
    return ME_OK;
}

int me_copy_pair_records_device(me_ctx *c, void *d_dst, int n_pairs, void *stream)
{
    if (!c || !d_dst) return fail(ME_EINVAL, "NULL argument");
    if (!c->have_search) return fail(ME_ESTATE, "me_copy_pair_records_device before any me_search");
    if (n_pairs < 1 || n_pairs > c->last_pairs) return fail(ME_EINVAL, "n_pairs=%d, last search had %d", n_pairs, c->last_pairs);
    CU_TRY(cudaSetDevice(c->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : c->s_compute;
    if (st != c->s_compute) CU_TRY(cudaStreamWaitEvent(st, c->ev_search, 0));
    CU_TRY(cudaMemcpyAsync(d_dst, c->args.pair_out + (size_t)c->slot * c->max_pairs, sizeof(me_pair_result) * (size_t)n_pairs,
                           cudaMemcpyDeviceToDevice, st));
    if (st != c->s_compute) {                        // later searches on the compute stream wait for this copy (launch_search)
        CU_TRY(cudaEventRecord(c->ev_rec, st));
        c->have_rec = true;
    }
    return ME_OK;
}

int me_microbench_vabsdiff4(int device, int mode, double *instr_per_clk_per_sm, double *sm_mhz_seen, int *sm_count)
{
    if (!instr_per_clk_per_sm || !sm_mhz_seen || !sm_count) return fail(ME_EINVAL, "NULL argument");
    CU_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    const int sms = prop.multiProcessorCount, threads = 1024, blocks = sms, iters = 8192;
    uint32_t *d_out = nullptr;
    long long *d_clk = nullptr;
    CU_TRY(cudaMalloc(&d_out, sizeof(uint32_t) * (size_t)blocks * threads));
    CU_TRY(cudaMalloc(&d_clk, sizeof(long long) * (size_t)blocks));
    cudaEvent_t e0, e1;
    CU_TRY(cudaEventCreate(&e0));
    CU_TRY(cudaEventCreate(&e1));
    float ms = 0;
    for (int rep = 0; rep < 3; ++rep) {           // the last repetition is the measurement
        CU_TRY(cudaEventRecord(e0));
        if (mode == 1) me::microbench_kernel<1><<<blocks, threads>>>(d_out, iters, d_clk, 12345u + rep);
        else me::microbench_kernel<0><<<blocks, threads>>>(d_out, iters, d_clk, 12345u + rep);
        CU_TRY(cudaEventRecord(e1));
        CU_TRY(cudaEventSynchronize(e1));
        CU_TRY(cudaEventElapsedTime(&ms, e0, e1));
    }
    std::vector<long long> clk((size_t)blocks);
    CU_TRY(cudaMemcpy(clk.data(), d_clk, sizeof(long long) * (size_t)blocks, cudaMemcpyDeviceToHost));
    double mean_clk = 0;
    for (long long v : clk) mean_clk += (double)v;
    mean_clk /= blocks;
    // one 1024-thread CTA owns its SM for the whole run: thread-instructions / that CTA's clocks
    *instr_per_clk_per_sm = (double)threads * 16.0 * iters / mean_clk;
    *sm_mhz_seen = mean_clk / (ms * 1e3);
    *sm_count = sms;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d_out); cudaFree(d_clk);
    return ME_OK;
}

} // extern "C"

// =============================================================================================
// me_search_static_scratch -- the one entry point that allocates nothing.
//
// Every other search call works on a context, and creating a context allocates (device ring,
// result buffers, pinned staging).  Some hosts of this library cannot allocate when they need a
// search: a plug-in living inside a framework whose memory pool already owns the whole GPU, a
// real-time loop in which cudaMalloc's implicit device-wide synchronisation would stall the other
// streams, or a worker that was handed its buffers by a parent process and has no budget of its
// own.  Such a caller brings the frames and the output array (device memory it already has) and
// this function supplies the kernel's working memory -- the per-block records the fused median
// reads back and the per-pair tickets -- from a fixed array that is part of the loaded module.
//
// The array is ME_STATIC_SCRATCH_BYTES (4 MiB) and every GPU that loads the library carries it
// until the process exits.  Requests that need more are refused with ME_EINVAL and a message that
// states need and limit (1080p/16x16: 137 KB per pair => up to 30 pairs per call; 4K/8x8: one pair).
//
// All calls share that one array and it keeps whatever the previous call left in it.  The tickets,
// the only part that must start at zero, are cleared on the caller's stream before each launch;
// everything else is overwritten before it is read.  Two calls whose kernels can overlap on the GPU
// (different streams, or different host threads) race on the array: issue them on one stream or
// order them yourself.
//
// The first call on a device does the one-time set-up for that device (it resolves the array's
// device address and raises the kernel's shared-memory limit).  Make that first call outside any
// stream capture.  The call is asynchronous: results are in d_out once `stream` has drained.
// =============================================================================================
static __device__ __align__(256) unsigned char g_static_scratch[ME_STATIC_SCRATCH_BYTES];

namespace {

struct StaticDevice {
    std::once_flag once;                // two host threads may make their first call on a device together
    int rc = ME_OK;
    char err[256] = "";
    bool ready = false;
    unsigned char *scratch = nullptr;   // device address of g_static_scratch on this device
    int sm_count = 0;
};
StaticDevice g_static_dev[64];

inline size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

// Once per device, with that device current.  In host code `g_static_scratch` is only a handle for
// the symbol; the address a kernel can use comes from cudaGetSymbolAddress.
int static_device_setup(int dev)
{
    StaticDevice &d = g_static_dev[dev];
    std::call_once(d.once, [&d, dev]() {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (e == cudaSuccess && !ptr) e = cudaErrorInvalidSymbol;
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) {
            d.rc = ME_ECUDA;
            snprintf(d.err, sizeof d.err, "static scratch set-up on device %d failed: %s", dev, cudaGetErrorString(e));
            return;
        }
        d.scratch = static_cast<unsigned char *>(ptr);
        d.ready = true;
    });
    return d.ready ? ME_OK : fail(d.rc ? d.rc : ME_ECUDA, "%s", d.err);
}

} // namespace

extern "C" int me_search_static_scratch(const uint8_t *d_frames, size_t pitch, int n_frames, int height, int width,
                                        const me_params *params, const int32_t *pairs, int n_pairs,
                                        me_pair_result *d_out, void *stream)
{
    if (!d_frames || !params || !pairs || !d_out) return fail(ME_EINVAL, "NULL argument");
    const int B = params->block, R = params->range;
    if (B != 8 && B != 16 && B != 32) return fail(ME_EUNSUPPORTED, "block size %d: built sizes are 8, 16, 32", B);
    if (R < 1) return fail(ME_EINVAL, "range %d < 1", R);
    if (params->range_y < 0 || params->range_y > R) return fail(ME_EINVAL, "range_y %d outside 0..range (%d); 0 means the same as range", params->range_y, R);
    if (height < B || width < B) return fail(ME_EINVAL, "frame %dx%d smaller than one %d-pixel block", width, height, B);
    if ((reinterpret_cast<uintptr_t>(d_frames) & 15) || pitch % 16 || pitch < (size_t)width)
        return fail(ME_EINVAL, "device frames need a 16-byte aligned base and a pitch >= width that is a multiple of 16");
    if (n_frames < 1 || n_pairs < 1) return fail(ME_EINVAL, "n_frames and n_pairs must be positive");
    for (int i = 0; i < 2 * n_pairs; ++i)
        if (pairs[i] < 0 || pairs[i] >= n_frames)
            return fail(ME_EINVAL, "pair %d references frame %d, outside 0..%d", i / 2, pairs[i], n_frames - 1);

    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(ME_EUNSUPPORTED, "device ordinal %d outside the static table", dev);
    int rc = static_device_setup(dev);
    if (rc) return rc;

    // Launch shape: the same planner the contexts use, run on a throw-away host-side description.
    me_ctx plan;
    plan.device = dev; plan.H = height; plan.W = width; plan.p = *params;
    plan.max_frames = n_frames; plan.max_pairs = n_pairs; plan.plan_pairs = n_pairs;
    plan.sm_count = g_static_dev[dev].sm_count;
    plan.pitch = pitch; plan.frame_stride = pitch * (size_t)height;
    rc = configure(&plan);
    if (rc) return rc;

    // Carve the static array: tickets | pairs | mv | sad | sub | flags | (slabbed searches: tile tickets | keys)
    const size_t nb = (size_t)plan.args.nb * (size_t)n_pairs;
    const bool slabbed = plan.args.nsx * plan.args.nsy > 1;
    const size_t nt = (size_t)plan.args.tiles_per_pair * (size_t)n_pairs;
    const size_t o_ticket = 0;
    const size_t o_pairs = o_ticket + align256(sizeof(unsigned) * (size_t)n_pairs);
    const size_t o_mv = o_pairs + align256(sizeof(int32_t) * 2 * (size_t)n_pairs);
    const size_t o_sad = o_mv + align256(nb * 2 * sizeof(int16_t));
    const size_t o_sub = o_sad + align256(nb * sizeof(uint32_t));
    const size_t o_flags = o_sub + align256(nb * 2 * sizeof(float));
    const size_t o_tt = o_flags + align256(nb);
    const size_t o_key = o_tt + (slabbed ? align256(nt * sizeof(unsigned)) : 0);
    const size_t need = o_key + (slabbed ? align256(nb * sizeof(unsigned long long)) : 0);
    if (need > (size_t)ME_STATIC_SCRATCH_BYTES)
        return fail(ME_EINVAL, "%d pair(s) of %d blocks need %zu bytes of working memory; the static array holds %u "
                    "(ME_STATIC_SCRATCH_BYTES). Split the batch or use a context (me_create / me_search_device).",
                    n_pairs, plan.args.nb, need, (unsigned)ME_STATIC_SCRATCH_BYTES);

    unsigned char *base = g_static_dev[dev].scratch;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CUtensorMap tw, tr;
    rc = make_tensor_map(&tw, d_frames, width, height, n_frames, pitch, plan.win_pitch, plan.args.box_rows);
    if (!rc) rc = make_tensor_map(&tr, d_frames, width, height, n_frames, pitch, me::kRefPitch, B * plan.args.tmr);
    if (rc) return rc;
    rc = raise_smem_limit(plan.kc.fn);
    if (rc) return rc;

    me::SearchArgs a = plan.args;
    a.frames = d_frames;
    a.pair_ticket = reinterpret_cast<unsigned int *>(base + o_ticket);
    a.pairs = reinterpret_cast<const int32_t *>(base + o_pairs);
    a.blk_mv = reinterpret_cast<int16_t *>(base + o_mv);
    a.blk_sad = reinterpret_cast<uint32_t *>(base + o_sad);
    a.blk_sub = reinterpret_cast<float *>(base + o_sub);
    a.blk_flags = base + o_flags;
    a.pair_out = d_out;
    // the tickets must start at zero (and, for slabbed searches, the keys at all-ones); the rest of the
    // array is written before it is read
    CU_TRY(cudaMemsetAsync(base + o_ticket, 0, sizeof(unsigned) * (size_t)n_pairs, st));
    if (slabbed) {
        a.tile_ticket = reinterpret_cast<unsigned int *>(base + o_tt);
        a.blk_key = reinterpret_cast<unsigned long long *>(base + o_key);
        CU_TRY(cudaMemsetAsync(base + o_tt, 0, nt * sizeof(unsigned), st));
        CU_TRY(cudaMemsetAsync(base + o_key, 0xff, nb * sizeof(unsigned long long), st));
    }
    CU_TRY(cudaMemcpyAsync(base + o_pairs, pairs, sizeof(int32_t) * 2 * (size_t)n_pairs, cudaMemcpyHostToDevice, st));
    void *kargs[3] = {(void *)&tw, (void *)&tr, (void *)&a};
    CU_TRY(cudaLaunchKernel(plan.kc.fn, dim3((unsigned)(a.tiles_per_pair * n_pairs * a.nsx * a.nsy)), dim3((unsigned)plan.threads), kargs,
                            (size_t)plan.smem, st));
    return ME_OK;
}

