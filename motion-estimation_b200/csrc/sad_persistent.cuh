// sad_persistent.cuh -- persistent, warp-specialised form of the search kernel (sm_100a only).
//
// Why: with one tile per CTA (sad_search_kernel) the load / copy-build / finalize phases of a CTA are
// not hidden behind other CTAs' search phases -- r01 measured 0.909 of the VABSDIFF4 roofline for the
// candidate phase alone against 0.838 for the whole kernel, a constant ~3.8 us per CTA.  Here one CTA
// per SM stays resident and walks a strided list of work units (tile x candidate slab):
//
//   producer warp   per unit: waits for a free stage, issues the TMA boxes into it, waits for them,
//                   derives the byte-shifted copies, publishes the unit (mbarrier "ready").
//   consumer warps  claim 32-item chunks of the unit's candidate patches from a shared counter (so the
//                   warps stay balanced whatever the tile size), run the same run_item() as the
//                   one-tile kernel, and go straight on to the next stage.  The warp that completes
//                   a unit's last chunk closes it: cross-slab keys / tile ticket, 8-lane finalize,
//                   pair ticket.  Nobody waits at a CTA-wide barrier between units.
//   two stages      of window copies in shared memory (2 x ~71 KB at 1080p/16/+-32), so the producer
//                   prepares unit k+1 while the consumers search unit k.
//   pair medians    pairs whose last tile closed are queued; after its units each CTA takes queued
//                   pairs in turn and reduces them with the whole CTA (reduce_pair()).
//
// Results are bit-identical to sad_search_kernel (same arithmetic, order-independent argmin).
#pragma once
#include "sad_search.cuh"

namespace me {

constexpr int kStages = 2;
#ifndef ME_CW
#define ME_CW 20
#endif
constexpr int kConsumerWarps = ME_CW;        // 8/16-pixel blocks (80 registers per thread)
constexpr int kConsumerWarps32 = 16;      // 32-pixel blocks: the patch needs ~110 registers
#ifndef ME_NP
#define ME_NP 2
#endif
constexpr int kProducerWarps = ME_NP;      // warps that build the shifted copies of a stage
constexpr int kPersistThreads = 32 * (kProducerWarps + kConsumerWarps);

struct UnitInfo {                  // one per stage: written by the producer, read by the consumers
    int pair, tile_id, tx, ty, prev_slot, cur_slot;
    int X0, Y0, nj, nblk;
    int ubase, dbase, slab_w, lastx;
    int dyi_lo, dyi_hi, tile_interior, nd, dy_cap;
    int n_wide, n_narrow, per_j_w, per_j_n, total_w, total_n;
    unsigned magic_pjw, magic_pjn;
    int chunks_w, chunks_n;        // 32-item chunks of wide / narrow patches
};

struct PersistArgs {
    int n_units;                   // tiles_per_pair * nslab * n_pairs
    int n_pairs;
    int stage_bytes;               // window copies (+raw) + slack + reference strip, multiple of 128
    int *ready_pairs;              // [n_pairs + 2]: [0] = counter (starts at -1), then pair ids (-1 = not yet),
                                   // [n_pairs + 1] = the launch's unit counter (starts at -1)
    int *error_flag;               // set when a wait timed out (never in a correct run)
};

__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// for waits that are expected to take microseconds (the producer waiting for a stage): back off between polls
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t *bar, uint32_t parity)
{
    uint32_t done, addr = smem_u32(bar);
    for (;;) {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) break;
        __nanosleep(256);
    }
}
__device__ __forceinline__ void mbar_init_only(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

template <int B, int NDY>
__device__ __forceinline__ UnitInfo decode_unit(const SearchArgs &a, int u)
{
    using G = Geo<B>;
    UnitInfo ui;
    const int nslab = a.nsx * a.nsy;
    ui.tile_id = u / nslab;
    const int slab = u - ui.tile_id * nslab;
    ui.pair = ui.tile_id / a.tiles_per_pair;
    const int tr = ui.tile_id - ui.pair * a.tiles_per_pair;
    ui.ty = tr / a.tiles_x;
    ui.tx = tr - ui.ty * a.tiles_x;
    const int sy = slab / a.nsx, sx = slab - sy * a.nsx;
    ui.prev_slot = __ldg(a.pairs + 2 * ui.pair);
    ui.cur_slot = __ldg(a.pairs + 2 * ui.pair + 1);
    ui.X0 = ui.tx * a.tmb * G::MB;
    ui.Y0 = ui.ty * B;
    ui.nj = min(a.tmb, a.nmbx - ui.tx * a.tmb);
    ui.nblk = min(ui.nj * G::NSUB, a.nbx - ui.tx * a.tmb * G::NSUB);
    ui.ubase = sx * a.su;
    ui.dbase = a.dy0 + sy * a.sd;
    ui.lastx = sx == a.nsx - 1;
    const int lasty = sy == a.nsy - 1;
    ui.slab_w = ui.lastx ? a.U - ui.ubase : a.su;
    const int slab_h = lasty ? a.dy0 + a.Uy - ui.dbase : a.sd;
    const int twoR = 2 * a.R;
    ui.dyi_lo = max(0, a.R - ui.Y0 - ui.dbase);
    ui.dyi_hi = min(slab_h - 1, min(twoR, a.R + a.H - B - ui.Y0) - ui.dbase);
    const int xlast = ui.X0 + (ui.nblk - 1) * B;
    ui.tile_interior = ui.dyi_lo == 0 && ui.dyi_hi == slab_h - 1 && a.R - ui.X0 - ui.ubase <= 0 &&
                       min(twoR, a.R + a.W - B - xlast) - ui.ubase >= ui.slab_w - 1;
    ui.dy_cap = max(ui.dyi_hi - NDY + 1, 0);        // see sad_search_kernel: groups start at dyi_lo, the last one ends at dyi_hi
    ui.nd = ui.dyi_hi >= ui.dyi_lo ? (ui.dyi_hi - ui.dyi_lo) / NDY + 1 : 0;
    ui.n_wide = a.n_wide[ui.lastx];
    ui.n_narrow = a.n_narrow[ui.lastx];
    ui.per_j_w = ui.nd * ui.n_wide;
    ui.per_j_n = ui.nd * ui.n_narrow;
    ui.total_w = ui.nj * ui.per_j_w;
    ui.total_n = ui.nj * ui.per_j_n;
    const bool usual = ui.nd == a.full_nd[lasty];
    ui.magic_pjw = usual ? a.magic_wide[lasty][ui.lastx] : fastdiv_magic((unsigned)ui.per_j_w);
    ui.magic_pjn = usual ? a.magic_narrow[lasty][ui.lastx] : fastdiv_magic((unsigned)ui.per_j_n);
    ui.chunks_w = (ui.total_w + 31) >> 5;
    ui.chunks_n = (ui.total_n + 31) >> 5;
    return ui;
}

// The warp that completed a unit's last chunk (or the producer, for a unit without candidates).
// `keys` = the stage's block minima in shared memory.  Returns after the stage may be reused by this warp.
// `release` (may be null): the stage's "empty" barrier; it is signalled as soon as the keys are in registers,
// long before the global-memory part of the finalize step is over.
// `s_win` / `s_ref` (may be null): the stage's window copies and reference strip.  A single-slab unit finalizes from
// them (finalize_block_smem) and releases the stage after that short loop instead of before the global-memory one.
template <int B, int PITCH, bool K64, int RB>
__device__ __forceinline__ void close_unit(const SearchArgs &a, const PersistArgs &pa, const UnitInfo &ui,
                                           typename KeyOf<K64>::T *keys, int lane, uint64_t *release,
                                           const uint8_t *s_win = nullptr, const uint8_t *s_ref = nullptr)
{
    using KeyT = typename KeyOf<K64>::T;
    using G = Geo<B>;
    constexpr KeyT kNone = ~KeyT(0);
    const int nslab = a.nsx * a.nsy;
    KeyT mykey[(kMaxKeys + 31) / 32];
#pragma unroll
    for (int i = 0; i < (kMaxKeys + 31) / 32; ++i) mykey[i] = lane + 32 * i < ui.nblk ? keys[lane + 32 * i] : kNone;
    __syncwarp();
    // ... for 16- and 32-pixel blocks (r01f: 0.861 -> 0.870, 32x32 0.867 -> 0.888).  Units of 8x8 blocks hold 12-24
    // blocks; keeping the stage through their finalize stalls the producer (+-32: 0.71 -> 0.64, +-48: 0.68 -> 0.65), so
    // they release first and finalize from global memory as before.
    const bool from_smem = B >= 16 && nslab == 1 && s_win != nullptr;
    if (!from_smem && release && lane == 0) mbar_arrive(release);
    if (nslab > 1) {
        const size_t kbase = (size_t)ui.pair * a.nb + (size_t)ui.ty * a.nbx + (size_t)ui.tx * a.tmb * G::NSUB;
        KeyT *gkey = reinterpret_cast<KeyT *>(a.blk_key);
#pragma unroll
        for (int i = 0; i < (kMaxKeys + 31) / 32; ++i)
            if (lane + 32 * i < ui.nblk && mykey[i] != kNone) atomicMin(&gkey[kbase + lane + 32 * i], mykey[i]);
        __threadfence();
        __syncwarp();
        unsigned old = 0;
        if (lane == 0) old = atomicAdd(&a.tile_ticket[ui.tile_id], 1u);
        old = __shfl_sync(kFull, old, 0);
        if (old != (unsigned)nslab - 1u) return;                       // another slab's CTA will finalize this tile
        __threadfence();
#pragma unroll
        for (int i = 0; i < (kMaxKeys + 31) / 32; ++i)
            if (lane + 32 * i < ui.nblk) {
                mykey[i] = __ldcg(&gkey[kbase + lane + 32 * i]);
                gkey[kbase + lane + 32 * i] = kNone;                    // leave the scratch as we found it
            }
        if (lane == 0) a.tile_ticket[ui.tile_id] = 0;
    }
    // finalize: 8 lanes per block, 4 blocks per pass; keys travel by shuffle
    const uint8_t *curf = a.frames + (long long)ui.cur_slot * a.frame_stride;
    const uint8_t *prevf = a.frames + (long long)ui.prev_slot * a.frame_stride;
    constexpr KeyT rmask = (KeyT(1) << RB) - 1;
    constexpr int per_warp = 32 / kFinLanes;
    const int grp = lane / kFinLanes;
    for (int jb0 = 0; jb0 < ui.nblk; jb0 += per_warp) {
        const int jb = jb0 + grp;
        const int src = jb & 31;
        KeyT key = kNone;
#pragma unroll
        for (int i = 0; i < (kMaxKeys + 31) / 32; ++i) {
            const KeyT k = __shfl_sync(kFull, mykey[i], src);
            if ((jb >> 5) == i) key = k;
        }
        if (jb < ui.nblk) {
            const unsigned fmask = ((1u << kFinLanes) - 1u) << (grp * kFinLanes);
            if (from_smem)
                finalize_block_smem<B, PITCH>(a, ui.pair, s_win, s_ref, 0, jb, ui.tx * a.tmb * G::NSUB + jb, ui.ty,
                                              (uint32_t)(key >> RB), (uint32_t)(key & rmask), lane % kFinLanes, fmask);
            else
                finalize_block<B>(a, ui.pair, prevf, curf, ui.tx * a.tmb * G::NSUB + jb, ui.ty, (uint32_t)(key >> RB),
                                  (uint32_t)(key & rmask), lane % kFinLanes, fmask);
        }
    }
    __syncwarp();
    if (from_smem && release && lane == 0) mbar_arrive(release);
    if (lane == 0) {
        __threadfence();
        const unsigned old = atomicAdd(&a.pair_ticket[ui.pair], 1u);
        if (old == (unsigned)a.tiles_per_pair - 1u) {                   // the pair is complete: queue it for reduction
            __threadfence();
            const int slot = atomicAdd(&pa.ready_pairs[0], 1) + 1;      // counter starts at -1
            atomicExch(&pa.ready_pairs[1 + slot], ui.pair);
        }
    }
}

template <int B, int NDY, int NQ, int PITCH, bool K64, int RB, int CW, int NP>
__global__ void __launch_bounds__(32 * (NP + CW), 1)
sad_search_persistent(const __grid_constant__ CUtensorMap tm_win, const __grid_constant__ CUtensorMap tm_ref,
                      const SearchArgs a, const PersistArgs pa)
{
    using KeyT = typename KeyOf<K64>::T;
    using G = Geo<B>;
    constexpr KeyT kNone = ~KeyT(0);
    constexpr int LWW = (NQ % 4 == 0) ? 4 : 1;
    constexpr int ref_bytes = kRefPitch * B;
    extern __shared__ __align__(1024) uint8_t smem[];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // shared-memory map: [stage 0][stage 1][reduction scratch 16.5 KB][control]
    unsigned *red_scratch = reinterpret_cast<unsigned *>(smem + (size_t)kStages * pa.stage_bytes);
    uint8_t *ctl = reinterpret_cast<uint8_t *>(red_scratch) + (2 * kSelBins + 64) * 4;
    UnitInfo *info = reinterpret_cast<UnitInfo *>(ctl);                              // [kStages]
    KeyT *keys = reinterpret_cast<KeyT *>(ctl + 2 * 128);                            // [kStages][kMaxKeys], 8 B slots
    unsigned *claim = reinterpret_cast<unsigned *>(ctl + 2 * 128 + kStages * kMaxKeys * 8);   // [kStages][2]: next, done
    uint64_t *bars = reinterpret_cast<uint64_t *>(ctl + 2 * 128 + kStages * kMaxKeys * 8 + 32);   // full[2], ready[2], empty[2]
    int *s_pair = reinterpret_cast<int *>(bars + 6);
    static_assert(sizeof(UnitInfo) <= 128, "UnitInfo must fit its 128-byte slot");

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init_only(&bars[s], 1);                       // full: one arrive.expect_tx + the TMA bytes
            mbar_init_only(&bars[2 + s], NP);                  // ready: every producer warp
            mbar_init_only(&bars[4 + s], CW);                  // empty: every consumer warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // Units are handed out through one global counter (not a fixed stride per CTA): slabs of different size, border
    // tiles and skipped slabs make unit costs uneven, and a static split left SMs idle at the end of slabbed launches
    // (r01f ncu: ALU pipe 95 % busy while active, yet 3 % slower than one CTA per tile).
    const int stride = gridDim.x;
    int *unit_counter = pa.ready_pairs + pa.n_pairs + 1;
    int *s_unit = reinterpret_cast<int *>(bars + 6) + 4;          // [kStages] unit of the stage, shared by the producer warps
    if (warp < NP) {
        // ================================================================ producers: warp 0 issues the TMA boxes and
        // writes the unit's control words, all NP warps build the shifted copies (chunk i goes to warp i mod NP)
        unsigned full_phase[kStages] = {0, 0};     // the TMA barrier only advances on units that load something
        for (int k = 0;; ++k) {
            const int s = k & 1, use = k >> 1;
            if (use > 0) mbar_wait_relaxed(&bars[4 + s], (use - 1) & 1);
            int u = 0;
            if (NP == 1) {
                if (lane == 0) u = atomicAdd(unit_counter, 1) + 1;
                u = __shfl_sync(kFull, u, 0);
            } else {
                if (warp == 0 && lane == 0) s_unit[s] = atomicAdd(unit_counter, 1) + 1;
                asm volatile("bar.sync 1, %0;" ::"n"(32 * NP) : "memory");      // producer warps only
                u = s_unit[s];
            }
            if (u >= pa.n_units) {                                         // nothing left: tell the consumers and leave
                if (warp == 0 && lane == 0) info[s].pair = -1;
                __threadfence_block();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[2 + s]);
                break;
            }
            const UnitInfo ui = decode_unit<B, NDY>(a, u);
            uint8_t *s_win = smem + (size_t)s * pa.stage_bytes;      // address arithmetic on the __shared__ symbol keeps LDS/STS
            uint8_t *s_raw = a.x_off ? s_win + 4 * (size_t)a.copy_bytes + 128 : s_win;
            uint8_t *s_ref = s_win + (size_t)a.n_bufs * a.copy_bytes + 128 + 256;
            KeyT *skeys = keys + s * kMaxKeys;
            if (warp == 0) {
                if (lane == 0) { info[s] = ui; claim[2 * s] = 0; claim[2 * s + 1] = 0; }
                if (lane < kMaxKeys) skeys[lane] = kNone;
            }
            if (ui.nd > 0) {
                const int wx0 = ui.X0 - a.R - a.x_off + ui.ubase, wy0 = ui.Y0 - a.R + ui.dbase;
                if (warp == 0 && lane == 0) {
                    const uint32_t bytes = (uint32_t)a.n_row_boxes * a.box_rows * PITCH + (uint32_t)ref_bytes;
                    mbar_expect_tx(&bars[s], bytes);
                    for (int rb = 0; rb < a.n_row_boxes; ++rb)
                        tma_load_3d(s_raw + (size_t)rb * a.box_rows * PITCH, &tm_win, wx0, wy0 + rb * a.box_rows,
                                    ui.cur_slot, &bars[s]);
                    tma_load_3d(s_ref, &tm_ref, ui.X0, ui.Y0, ui.prev_slot, &bars[s]);
                }
                mbar_wait(&bars[s], full_phase[s] & 1);
                ++full_phase[s];
                // byte-shifted copies: chunk i goes to producer thread i mod (32 * NP)
                build_shifted_copies<PITCH>(a, s_raw, s_win, warp * 32 + lane, 32 * NP);
            }
            __threadfence_block();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars[2 + s]);                      // this warp's share of the unit is published
            if (warp == 0 && ui.chunks_w + ui.chunks_n == 0)               // no candidates here: the producer closes the unit itself
                close_unit<B, PITCH, K64, RB>(a, pa, ui, skeys, lane, nullptr);
        }
    } else {
        // ================================================================ consumers
        for (int k = 0;; ++k) {
            const int s = k & 1, use = k >> 1;
            mbar_wait(&bars[2 + s], use & 1);
            const UnitInfo &ui = info[s];
            if (ui.pair < 0) break;                                        // the producer found the unit counter exhausted
            const uint8_t *s_win = smem + (size_t)s * pa.stage_bytes;
            const uint8_t *s_ref = s_win + (size_t)a.n_bufs * a.copy_bytes + 128 + 256;
            KeyT *skeys = keys + s * kMaxKeys;
            const int chunks_w = ui.chunks_w, nchunks = chunks_w + ui.chunks_n;
            const unsigned short *grp = a.group[ui.lastx];
            bool released = false;
            while (nchunks > 0) {
                unsigned c = 0;
                if (lane == 0) c = atomicAdd(&claim[2 * s], 1u);
                c = __shfl_sync(kFull, c, 0);
                if (c >= (unsigned)nchunks) break;
                if (c < (unsigned)chunks_w) {
                    const int w_raw = (int)c * 32 + lane, total = ui.total_w, per_j = ui.per_j_w, n_wide = ui.n_wide;
                    const bool active = w_raw < total;
                    const int w = active ? w_raw : total - 1;
                    const int j = fastdiv(w, ui.magic_pjw), r = w - j * per_j;
                    const int dq = fastdiv(r, a.magic_nw[ui.lastx]), cg = r - dq * n_wide, dyi0 = min(ui.dyi_lo + dq * NDY, ui.dy_cap);
                    const unsigned g = grp[cg];
                    run_item<B, NDY, NQ, PITCH, LWW, K64, RB>(a, active, j, dyi0, (int)(g >> 12), (int)(g & 0xfffu), s_win, s_ref,
                                                              skeys, ui.tile_interior != 0, ui.dyi_lo, ui.dyi_hi, ui.X0, lane,
                                                              ui.ubase, ui.dbase, ui.slab_w);
                } else {
                    const int w_raw = ((int)c - chunks_w) * 32 + lane, total = ui.total_n, per_j = ui.per_j_n, n_narrow = ui.n_narrow;
                    const bool active = w_raw < total;
                    const int w = active ? w_raw : total - 1;
                    const int j = fastdiv(w, ui.magic_pjn), r = w - j * per_j;
                    const int dq = fastdiv(r, a.magic_nn[ui.lastx]), cg = r - dq * n_narrow, dyi0 = min(ui.dyi_lo + dq * NDY, ui.dy_cap);
                    const unsigned g = grp[ui.n_wide + cg];
                    run_item<B, NDY, 1, PITCH, 1, K64, RB>(a, active, j, dyi0, (int)(g >> 12), (int)(g & 0xfffu), s_win, s_ref,
                                                           skeys, ui.tile_interior != 0, ui.dyi_lo, ui.dyi_hi, ui.X0, lane,
                                                           ui.ubase, ui.dbase, ui.slab_w);
                }
                __threadfence_block();
                __syncwarp();
                unsigned old = 0;
                if (lane == 0) old = atomicAdd(&claim[2 * s + 1], 1u);
                old = __shfl_sync(kFull, old, 0);
                if (old == (unsigned)nchunks - 1u) {                       // this warp finished the unit's last chunk
                    __threadfence_block();
                    const UnitInfo mine = ui;                              // private copy: the stage is released inside close_unit
                    close_unit<B, PITCH, K64, RB>(a, pa, mine, skeys, lane, &bars[4 + s], s_win, s_ref);
                    released = true;
                    break;
                }
            }
            __syncwarp();
            if (!released && lane == 0) mbar_arrive(&bars[4 + s]);         // this warp is done with the stage
        }
    }
    __syncthreads();

    // ================================================================ queued pair medians, whole CTA
    for (int i = blockIdx.x; i < pa.n_pairs; i += stride) {
        if (tid == 0) {
            int p = -1;
            unsigned long long t0, t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            while ((p = __ldcg(&pa.ready_pairs[1 + i])) < 0) {
                __nanosleep(200);
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 2000000000ull) { atomicExch(pa.error_flag, 1); break; }   // 2 s: a bug, not a hang
            }
            __threadfence();
            *s_pair = p;
        }
        __syncthreads();
        const int p = *s_pair;
        if (p >= 0) reduce_pair(a, p, red_scratch);
        __syncthreads();
    }
}

} // namespace me
