// sad_scan.cuh -- the anchored pair walk of the video script as ONE resident kernel (sm_100a only).
//
// estimate_vel_vid.py:506-534, :575 compares frame k with the last ACCEPTED frame: which pair comes next depends on the
// result of the pair before.  Driven from the host that is one launch + copy + synchronise per dependent step (~25 us
// on top of the search itself, r02a: 1803 steps for a 4096-frame recording).  Here the whole walk is one cooperative
// launch; per step
//     every CTA      searches its share of the batch (anchor, k) ... (anchor, k + L - 1)   [search_tile, as ever]
//     grid barrier   the CTA that arrives LAST (every pair record of the batch is complete by then: each pair's last CTA
//                    reduced it) applies the gate to the L records in order, writes the scan records of the frames it
//                    decided and the next batch -- its size from the distance of the last accept, as the host loop does --
//                    and only then releases the others: one barrier per step.  The next batch (L, anchor slot, first
//                    frame slot) travels IN the barrier word the others spin on, so a released CTA needs no further
//                    read before it issues its loads.
// The results are byte-identical to the host-driven loop (same batches, same records, same gate arithmetic).
#pragma once
#include "sad_search.cuh"

namespace me {

struct ScanArgs {
    int first_slot, n_frames, anchor_slot;     // frames first_slot .. first_slot + n_frames - 1 are classified in order
    float scale, gate;                         // skip (keep the anchor) when pdx * scale < gate and flag == 1
    int max_look;                              // largest batch
    me_scan_record *out;                       // [n_frames], device
    int32_t *pairs;                            // the search's pair list (SearchArgs::pairs); unused by this kernel (the batch is in the barrier word)
    unsigned *sync;                            // [0] arrivals, [2..3] the 64-bit barrier word (scan_pack); zero on entry
    int *state;                                // [0] batch size L (0 = done) [1] k [2] anchor [3] look [4] since [5] searches [6] steps [7] error
    unsigned long long *trace;                 // debug (ME_SCAN_TRACE=n): [0] ns CTA 0 spent in search_tile, [1] ns it spent in barriers, else null
    int trace_cta;                             // the CTA whose phases are timed (n - 1)
};

// The barrier word every CTA spins on carries the next batch, so a released CTA starts its unit without another L2
// round trip for the batch size or the pair list: generation [63:49] | L [48:40] | anchor slot [39:20] | first frame slot k [19:0].
constexpr int kScanSlotBits = 20;
__device__ __forceinline__ unsigned long long scan_pack(unsigned gen, int L, int anchor, int k)
{
    return ((unsigned long long)(gen & 0x7fffu) << 49) | ((unsigned long long)L << 40) | ((unsigned long long)anchor << kScanSlotBits) |
           (unsigned long long)k;
}

// All CTAs of the (cooperative, co-resident) grid; `last_arriver()` runs in one thread of the CTA that completes the barrier,
// before anybody is released, and returns the next batch (L, anchor, k packed as above, generation bits zero).  Returns the
// barrier word of the new generation, or 0 with the error flag set: a 2 s guard turns a lost CTA into an error instead of a hang.
template <class F>
__device__ __forceinline__ unsigned long long scan_grid_sync(const ScanArgs &sc, F last_arriver)
{
    __shared__ unsigned long long s_word;
    volatile unsigned long long *word = reinterpret_cast<volatile unsigned long long *>(sc.sync + 2);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long w = *word;
        const unsigned gen = (unsigned)(w >> 49);
        __threadfence();
        if (atomicAdd(sc.sync, 1u) == gridDim.x - 1) {
            *reinterpret_cast<volatile unsigned *>(sc.sync) = 0u;
            __threadfence();
            w = last_arriver() | ((unsigned long long)((gen + 1) & 0x7fffu) << 49);
            __threadfence();
            *word = w;
        } else {
            unsigned long long t0, t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            while ((unsigned)((w = *word) >> 49) == gen) {
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 2000000000ull || *reinterpret_cast<volatile int *>(sc.state + 7)) { atomicExch(sc.state + 7, 1); w = 0; break; }
            }
        }
        __threadfence();
        s_word = w;
    }
    __syncthreads();
    return s_word;
}

// The batch that starts at frame slot k with `look` as the wanted size (bookkeeping in state[], the batch itself in the return value).
__device__ __forceinline__ unsigned long long scan_next_batch(const ScanArgs &sc, int anchor, int k, int look)
{
    const int end = sc.first_slot + sc.n_frames;
    int L = min(min(look, end - k), sc.max_look);
    if (k >= end) L = 0;
    sc.state[0] = L; sc.state[1] = k; sc.state[2] = anchor; sc.state[3] = look;
    if (L) { sc.state[5] += L; sc.state[6] += 1; }
    return scan_pack(0, L, anchor, k);
}

template <int B, int NDY, int NQ, int PITCH, int RB, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB)
sad_scan_kernel(const __grid_constant__ CUtensorMap tm_win, const __grid_constant__ CUtensorMap tm_ref, const SearchArgs a,
                const ScanArgs sc)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    unsigned long long word = scan_grid_sync(sc, [&] {
        sc.state[4] = 0; sc.state[5] = 0; sc.state[6] = 0;
        return scan_next_batch(sc, sc.anchor_slot, sc.first_slot, min(4, sc.max_look));
    });
    const int per_pair = a.tiles_per_pair * a.nsx * a.nsy;
    // ME_SCAN_TRACE = n: CTA n - 1 adds up the cycles of its phases in trace[8..14] (search_tile's SCAN build)
    unsigned long long *phase = sc.trace && (int)blockIdx.x == sc.trace_cta ? sc.trace + 8 : nullptr;
    for (;;) {
        const int L = (int)((word >> 40) & 0x1ffu), anchor_now = (int)((word >> kScanSlotBits) & ((1u << kScanSlotBits) - 1u)),
                  k_now = (int)(word & ((1u << kScanSlotBits) - 1u));
        if (L == 0) break;                                              // finished, or a barrier gave up (word 0)
        unsigned long long t0 = 0, t1 = 0, t2 = 0;
        if (sc.trace && blockIdx.x == 0 && threadIdx.x == 0) { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0)); sc.trace[4] = t0; }
        TileTrace tt;
        tt.phase = phase; tt.unit_max = sc.trace ? sc.trace + 6 : nullptr; tt.reduce = sc.trace ? sc.trace + 13 : nullptr; tt.t_step = 0;
        if (sc.trace) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt.t_step));
        for (int u = blockIdx.x; u < per_pair * L; u += gridDim.x) {
            search_tile<B, NDY, NQ, PITCH, false, RB, false, true>(tm_win, tm_ref, a, u, smem, sc.trace ? &tt : nullptr, anchor_now, k_now + u / per_pair);
            __syncthreads();                                           // the next unit reuses the shared memory
        }
        if (sc.trace && blockIdx.x == 0 && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        word = scan_grid_sync(sc, [&] {
            unsigned long long ta = 0, tb = 0;
            if (sc.trace) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ta));
            const int k = k_now, anchor = anchor_now, look = sc.state[3], since = sc.state[4];
            unsigned long long next;
            int acc = -1;
            static_assert(sizeof(me_pair_result) == 32, "pair record is read as two 16-byte words");
            for (int i0 = 0; i0 < L && acc < 0; i0 += 4) {                                     // four records in flight (L2 reads: written by other CTAs)
                int4 w[4][2];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (i0 + u < L) {
                        const int4 *p = reinterpret_cast<const int4 *>(a.pair_out + i0 + u);
                        w[u][0] = __ldcg(p); w[u][1] = __ldcg(p + 1);
                    }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int i = i0 + u;
                    if (i >= L || acc >= 0) break;
                    me_pair_result r;
                    memcpy(&r, &w[u][0], 16);
                    memcpy(reinterpret_cast<char *>(&r) + 16, &w[u][1], 16);
                    const bool skip = (__fmul_rn(r.pdx, sc.scale) < sc.gate) && r.flag == 1;  // estimate_vel_vid.py:531-534
                    me_scan_record o;
                    o.anchor_slot = anchor; o.accepted = skip ? 0 : 1; o.pair = r;
                    sc.out[k + i - sc.first_slot] = o;
                    if (!skip) acc = i;                                                        // :575
                }
            }
            if (acc >= 0) {
                const int dist = since + acc + 1;
                sc.state[4] = 0;
                next = scan_next_batch(sc, k + acc, k + acc + 1, min(max(dist + 1, 2), sc.max_look));
            } else {
                sc.state[4] = since + L;
                next = scan_next_batch(sc, anchor, k + L, min(2 * look, sc.max_look));
            }
            if (sc.trace) {                                             // [2] step start (CTA 0) -> last arrival, [3] the gate, [5] sum of L
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tb));
                sc.trace[2] += ta - *reinterpret_cast<volatile unsigned long long *>(sc.trace + 4);
                sc.trace[3] += tb - ta;
                sc.trace[5] += (unsigned long long)L;
                sc.trace[7] += *reinterpret_cast<volatile unsigned long long *>(sc.trace + 6);      // the step's slowest unit
                *reinterpret_cast<volatile unsigned long long *>(sc.trace + 6) = 0ull;
            }
            return next;
        });
        if (sc.trace && blockIdx.x == 0 && threadIdx.x == 0) {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t2));
            sc.trace[0] += t1 - t0; sc.trace[1] += t2 - t1;
        }
    }
}

} // namespace me
