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
//                    and only then releases the others: one barrier per step
// The results are byte-identical to the host-driven loop (same batches, same records, same gate arithmetic).
#pragma once
#include "sad_search.cuh"

namespace me {

struct ScanArgs {
    int first_slot, n_frames, anchor_slot;     // frames first_slot .. first_slot + n_frames - 1 are classified in order
    float scale, gate;                         // skip (keep the anchor) when pdx * scale < gate and flag == 1
    int max_look;                              // largest batch
    me_scan_record *out;                       // [n_frames], device
    int32_t *pairs;                            // the search's pair list (SearchArgs::pairs), rewritten per step
    unsigned *sync;                            // [0] arrivals, [1] generation; zero on entry
    int *state;                                // [0] batch size L (0 = done) [1] k [2] anchor [3] look [4] since [5] searches [6] steps [7] error
    unsigned long long *trace;                 // debug (ME_SCAN_TRACE=1): [0] ns CTA 0 spent in search_tile, [1] ns it spent in barriers, else null
};

// All CTAs of the (cooperative, co-resident) grid; `last_arriver()` runs in one thread of the CTA that completes the barrier,
// before anybody is released.  A 2 s guard turns a lost CTA into an error flag instead of a hang.
template <class F>
__device__ __forceinline__ bool scan_grid_sync(const ScanArgs &sc, F last_arriver)
{
    __shared__ int s_ok;
    __syncthreads();
    if (threadIdx.x == 0) {
        int ok = 1;
        const unsigned gen = *reinterpret_cast<volatile unsigned *>(sc.sync + 1);
        __threadfence();
        if (atomicAdd(sc.sync, 1u) == gridDim.x - 1) {
            *reinterpret_cast<volatile unsigned *>(sc.sync) = 0u;
            __threadfence();
            last_arriver();
            __threadfence();
            atomicAdd(sc.sync + 1, 1u);
        } else {
            unsigned long long t0, t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            while (*reinterpret_cast<volatile unsigned *>(sc.sync + 1) == gen) {
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 2000000000ull || *reinterpret_cast<volatile int *>(sc.state + 7)) { atomicExch(sc.state + 7, 1); ok = 0; break; }
            }
        }
        __threadfence();
        s_ok = ok;
    }
    __syncthreads();
    return s_ok != 0;
}

// The batch that starts at frame slot k with `look` as the wanted size.
__device__ __forceinline__ void scan_write_batch(const ScanArgs &sc, int anchor, int k, int look)
{
    const int end = sc.first_slot + sc.n_frames;
    int L = min(min(look, end - k), sc.max_look);
    if (k >= end) L = 0;
    for (int i = 0; i < L; ++i) { sc.pairs[2 * i] = anchor; sc.pairs[2 * i + 1] = k + i; }
    sc.state[0] = L; sc.state[1] = k; sc.state[2] = anchor; sc.state[3] = look;
    if (L) { sc.state[5] += L; sc.state[6] += 1; }
}

template <int B, int NDY, int NQ, int PITCH, int RB, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB)
sad_scan_kernel(const __grid_constant__ CUtensorMap tm_win, const __grid_constant__ CUtensorMap tm_ref, const SearchArgs a,
                const ScanArgs sc)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        sc.state[4] = 0; sc.state[5] = 0; sc.state[6] = 0;
        scan_write_batch(sc, sc.anchor_slot, sc.first_slot, min(4, sc.max_look));
    }
    if (!scan_grid_sync(sc, [] {})) return;
    const int per_pair = a.tiles_per_pair * a.nsx * a.nsy;
    for (;;) {
        const int L = *reinterpret_cast<volatile int *>(sc.state);
        if (L == 0) break;
        unsigned long long t0 = 0, t1 = 0, t2 = 0;
        if (sc.trace && blockIdx.x == 0 && threadIdx.x == 0) { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0)); sc.trace[4] = t0; }
        for (int u = blockIdx.x; u < per_pair * L; u += gridDim.x) {
            search_tile<B, NDY, NQ, PITCH, false, RB, false>(tm_win, tm_ref, a, u, smem);
            __syncthreads();                                           // the next unit reuses the shared memory
        }
        if (sc.trace && blockIdx.x == 0 && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        const bool ok = scan_grid_sync(sc, [&] {
            unsigned long long ta = 0, tb = 0;
            if (sc.trace) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ta));
            const int k = sc.state[1], anchor = sc.state[2], look = sc.state[3], since = sc.state[4];
            int acc = -1;
            for (int i = 0; i < L; ++i) {
                me_pair_result r;                                                              // written by another CTA: read through L2
                {
                    const int4 *p = reinterpret_cast<const int4 *>(a.pair_out + i);
                    const int4 lo = __ldcg(p), hi = __ldcg(p + 1);
                    static_assert(sizeof(me_pair_result) == 32, "pair record is read as two 16-byte words");
                    memcpy(&r, &lo, 16);
                    memcpy(reinterpret_cast<char *>(&r) + 16, &hi, 16);
                }
                const bool skip = (__fmul_rn(r.pdx, sc.scale) < sc.gate) && r.flag == 1;      // estimate_vel_vid.py:531-534
                me_scan_record o;
                o.anchor_slot = anchor; o.accepted = skip ? 0 : 1; o.pair = r;
                sc.out[k + i - sc.first_slot] = o;
                if (!skip) { acc = i; break; }                                                 // :575
            }
            if (acc >= 0) {
                const int dist = since + acc + 1;
                sc.state[4] = 0;
                scan_write_batch(sc, k + acc, k + acc + 1, min(max(dist + 1, 2), sc.max_look));
            } else {
                sc.state[4] = since + L;
                scan_write_batch(sc, anchor, k + L, min(2 * look, sc.max_look));
            }
            if (sc.trace) {                                             // [2] step start (CTA 0) -> last arrival, [3] the gate, [5] sum of L
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tb));
                sc.trace[2] += ta - *reinterpret_cast<volatile unsigned long long *>(sc.trace + 4);
                sc.trace[3] += tb - ta;
                sc.trace[5] += (unsigned long long)L;
            }
        });
        if (sc.trace && blockIdx.x == 0 && threadIdx.x == 0) {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t2));
            sc.trace[0] += t1 - t0; sc.trace[1] += t2 - t1;
        }
        if (!ok) return;
    }
}

} // namespace me
