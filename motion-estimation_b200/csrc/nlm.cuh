// nlm.cuh -- non-local-means denoising of ring frames (single image and temporal window).
//
// Replaces cv.fastNlMeansDenoising / cv.fastNlMeansDenoisingMulti(h=15, templateWindowSize=19,
// searchWindowSize=41, temporalWindowSize=5|9) as denoise_hf5 / denoising call them
// (estimate_vel_vid.py:226-232, 246-262) -- 99 % of the reference's wall time (1.9 s per 640x64 frame).
// Bit-exact with OpenCV's 8-bit implementation (restated in oracle/nlm_np.py and pinned against cv2 there):
// squared-difference patch distance summed exactly over the template window, weight table indexed by
// dist >> shift, 32-bit fixed-point accumulation, one rounded division.
//
// nlm_kernel<TWS>: one CTA of 256 threads owns a 64x128 tile of output pixels and walks the whole search
// volume (frame, dy, dx).  The displaced frame's region (tile + template + search halo, reflect-101 at the
// frame edge) sits in shared memory; per offset
//   pass A  (row, 16-column segment) tasks: |a-b| of 4 pixels per VABSDIFF4, squares with IMAD, a sliding
//           TWS-wide horizontal sum, 16 sums stored as 4 x STS.128; the words of the frame being denoised
//           stay in registers for the whole kernel (they do not depend on the offset);
//   pass B  one thread per (column, 32-row half): sliding TWS-high vertical sum of pass A's rows with the
//           window's rows in registers (each sum is read once), weight lookup, est += w * pixel,
//           wsum += w with the 64 accumulators in registers.
// Integer-ALU/shared-memory bound: about 20 thread-instructions per output pixel per offset; HBM traffic
// is the tile regions once per temporal frame (nothing per offset).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace me {

struct NlmArgs {
    const uint8_t *src;         // ring base
    long long frame_stride;
    int pitch, W, H;
    int first_slot;             // ring slot of frame 0 of this call
    uint8_t *dst;               // [n][H][pitch] denoised frames (scratch, copied back by the host)
    const int *twin;            // [n] temporal window (odd) of each frame; 0 = copy the frame through
    const int *lut_off;         // [n] offset of the frame's weight table in luts
    const int *lut_len;         // [n] entries before the table turns to zero
    const int *luts;
    int sh;                     // search window half size
    int shift;                  // dist >> shift indexes the table
    int itp, it_rows;           // shared region: row pitch in bytes (odd number of words), rows
    int nsplit;                 // CTAs per tile along the (frame, dy) axis; > 1 accumulates through acc
    int *acc;                   // [n][H][W][2] (est, wsum), zeroed by the host when nsplit > 1
};

__device__ __forceinline__ int reflect101_iter(int p, int n)
{
    if (n == 1) return 0;
    while ((unsigned)p >= (unsigned)n) p = p < 0 ? -p : 2 * n - 2 - p;     // OpenCV iterates for borders wider than the frame
    return p;
}

constexpr int kNlmK = 32, kNlmRS = 2, kNlmTW = 128, kNlmTR = kNlmK * kNlmRS, kNlmS = 16, kNlmThreads = 256;

// Shared-region row pitch for a search half-size: whole words, an odd number of them (lanes on consecutive rows hit distinct banks).
__host__ __device__ constexpr int nlm_region_pitch(int sh, int th)
{
    int p = (kNlmTW + 2 * (sh + th) + 3) / 4 * 4;
    return (p / 4) % 2 == 0 ? p + 4 : p;
}
__host__ __device__ constexpr int nlm_shift(int tws)            // OpenCV: dist >> shift with 2^shift >= template area
{
    int s = 0;
    while ((1 << s) < tws * tws) ++s;
    return s;
}

// SH > 0: the search half-size is a compile-time constant (the reference's 41-pixel window: SH = 20), so the region pitch
// and every row offset are immediates; SH = 0 reads them from the arguments (other window sizes).
//
// Schedule: the horizontal pass of offset k+1 and the vertical pass of offset k run between the same two barriers
// (two buffers of horizontal sums), so an offset costs one __syncthreads instead of two and a thread's light and heavy
// phases even out (r02 ncu of the two-barrier form: 0.51 barrier stalls per issued instruction, 59.7 % issue-active).
// LS: the frame's weight table is staged in shared memory as 16-bit entries (the host checks that it fits: the
// reference's parameters give weights <= 5009 and ~1100 entries): a table read is then SHF + VIMNMX + LDS instead of a
// 64-bit address computation and a global load (r02b ncu: 4 of 25.5 thread-instructions per offset and pixel).
constexpr int kNlmLutSmem = 1280;          // entries of the shared-memory table (2.5 KB)
template <int TWS, int SH, bool LS>
__global__ void __launch_bounds__(kNlmThreads, 2) nlm_kernel(const NlmArgs a)
{
    constexpr int TH = TWS / 2, K = kNlmK, TW = kNlmTW, TR = kNlmTR, NR = TR + 2 * TH, S = kNlmS, NSEG = TW / S,
                  NTASK = NSEG * NR, ROUNDS = (NTASK + kNlmThreads - 1) / kNlmThreads, NPX = S + TWS - 1, NW = (NPX + 3) / 4,
                  HP = TW + 4;                                   // HP = 4 mod 32: conflict-free STS.128 from lanes on consecutive rows
    constexpr int SHIFT = nlm_shift(TWS);
    static_assert(kNlmThreads == TW * (TR / K), "vertical pass: one thread per (column, K-row band)");
    extern __shared__ __align__(16) unsigned char nlm_smem[];
    int *s_h0 = reinterpret_cast<int *>(nlm_smem);              // 2 x (NR x HP) horizontal sums
    unsigned char *s_it = nlm_smem + 2 * NR * HP * 4;           // it_rows x itp displaced-frame region
    unsigned short *s_lut = reinterpret_cast<unsigned short *>(nlm_smem + 2 * NR * HP * 4 + a.it_rows * a.itp);   // LS only

    const int tid = threadIdx.x;
    const int n = blockIdx.z / a.nsplit, sp = blockIdx.z - n * a.nsplit;
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TR;
    const int sh = SH > 0 ? SH : a.sh, sws = 2 * sh + 1, b = sh + TH;
    const int itp = SH > 0 ? nlm_region_pitch(SH, TH) : a.itp, it_rows = TR + 2 * b, itw = itp >> 2;
    const int T = a.twin[n], half = T / 2;
    const int *lut = a.luts + a.lut_off[n];
    const int lut_n = a.lut_len[n];                              // lut[lut_n] == 0: indices are clamped, not tested
    const uint8_t *main_f = a.src + (long long)(a.first_slot + n) * a.frame_stride;

    if (T == 0) {                                                 // window 0: the frame is only a neighbour, pass it through
        if (sp == 0)
            for (int idx = tid; idx < TR * TW; idx += kNlmThreads) {
                const int gy = y0 + idx / TW, gx = x0 + idx % TW;
                if (gy < a.H && gx < a.W) a.dst[((long long)n * a.H + gy) * a.pitch + gx] = main_f[(size_t)gy * a.pitch + gx];
            }
        return;
    }

    // words of the frame being denoised that this thread's horizontal-pass tasks compare against (offset independent)
    uint32_t i0[ROUNDS][NW];
#pragma unroll
    for (int rd = 0; rd < ROUNDS; ++rd) {
        const int task = tid + rd * kNlmThreads;
        const int seg = task / NR, r = task - seg * NR;
        const uint8_t *row = main_f + (size_t)reflect101_iter(y0 - TH + r, a.H) * a.pitch;
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            uint32_t w = 0;
            if (task < NTASK)
#pragma unroll
                for (int q = 0; q < 4; ++q) w |= (uint32_t)row[reflect101_iter(x0 - TH + seg * S + 4 * k + q, a.W)] << (8 * q);
            i0[rd][k] = w;
        }
    }

    int est[K], wsum[K];                                        // row i of the thread's column
#pragma unroll
    for (int i = 0; i < K; ++i) est[i] = wsum[i] = 0;

    const int rows_all = T * sws;
    const int per = (rows_all + a.nsplit - 1) / a.nsplit;
    const int row_lo = sp * per, row_hi = min(rows_all, row_lo + per);
    // vertical pass: a thread owns one column and K rows; the TWS rows of its sliding window stay in registers, so every
    // horizontal sum is read from shared memory once (K + TWS - 1 word loads per K outputs)
    const int j = tid & (TW - 1), rh = tid / TW;
    if (LS) {
        for (int i = tid; i <= lut_n; i += kNlmThreads) s_lut[i] = (unsigned short)lut[i];
        // visible after the first __syncthreads below (the region staging)
    }

    // ---- horizontal pass of one offset into buffer `buf`: TWS-wide sums of squared differences
    auto pass_a = [&](int dy, int dx, int buf) {
        int *s_h = s_h0 + buf * (NR * HP);
#pragma unroll
        for (int rd = 0; rd < ROUNDS; ++rd) {
            const int task = tid + rd * kNlmThreads;
            if (task < NTASK) {
                const int seg = task / NR, r = task - seg * NR;
                const int start = seg * S + dx + sh;
                const uint32_t *row = reinterpret_cast<const uint32_t *>(s_it + (r + sh + dy) * itp) + (start >> 2);
                const int sft = (start & 3) * 8;
                uint32_t w[NW + 1], d[NW];
#pragma unroll
                for (int k = 0; k <= NW; ++k) w[k] = row[k];
#pragma unroll
                for (int k = 0; k < NW; ++k) d[k] = __vabsdiffu4(i0[rd][k], __funnelshift_r(w[k], w[k + 1], sft));
                // squares first (independent of one another), then the sliding sum as one three-input add per output: a short
                // dependent chain instead of 49 chained multiply-adds (the kernel is latency-bound at 4 warps per scheduler)
                int sq[NPX];
#pragma unroll
                for (int x = 0; x < NPX; ++x) {
                    const int px = (int)__byte_perm(d[x >> 2], 0u, 0x4440u | (unsigned)(x & 3));                      // one PRMT per pixel
                    sq[x] = px * px;
                }
                int part[4] = {0, 0, 0, 0};
#pragma unroll
                for (int x = 0; x < TWS; ++x) part[x & 3] += sq[x];
                int out[S];
                out[0] = (part[0] + part[1]) + (part[2] + part[3]);
#pragma unroll
                for (int c = 1; c < S; ++c) out[c] = out[c - 1] + sq[c + TWS - 1] - sq[c - 1];
                int4 *dstp = reinterpret_cast<int4 *>(s_h + r * HP + seg * S);
#pragma unroll
                for (int e = 0; e < S / 4; ++e) dstp[e] = make_int4(out[4 * e], out[4 * e + 1], out[4 * e + 2], out[4 * e + 3]);
            }
        }
    };
    // ---- vertical pass of one offset from buffer `buf`: TWS-high sums, weights, accumulation
    auto pass_b = [&](int dy, int dx, int buf) {
        const int *hp = s_h0 + buf * (NR * HP) + (rh * K) * HP + j;
        const unsigned char *pp = s_it + (rh * K + b + dy) * itp + (j + b + dx);
        int hw[TWS], v = 0;
#pragma unroll
        for (int r = 0; r < TWS; ++r) { hw[r] = hp[r * HP]; v += hw[r]; }
#pragma unroll
        for (int i = 0; i < K; ++i) {
            int w;
            if (LS) w = s_lut[min(v >> SHIFT, lut_n)];               // clamped: the entry after the table is 0
            else w = __ldg(lut + min(v >> SHIFT, lut_n));
            est[i] += w * (int)pp[i * itp];
            wsum[i] += w;
            if (i + 1 < K) {
                const int in = hp[(i + TWS) * HP];
                v += in - hw[i % TWS];
                hw[i % TWS] = in;
            }
        }
    };

    int cur_t = -1, pend_dy = 0, pend_dx = 0, buf = 0;
    bool pending = false;
    for (int orow = row_lo; orow < row_hi; ++orow) {
        const int ti = orow / sws, dy = orow - ti * sws - sh;
        if (ti != cur_t) {                                        // stage the next temporal frame's region
            cur_t = ti;
            if (pending) { pass_b(pend_dy, pend_dx, buf ^ 1); pending = false; }    // the last offset of the old frame still reads the old region
            const uint8_t *f = main_f + (long long)(ti - half) * a.frame_stride;
            __syncthreads();
            for (int idx = tid; idx < it_rows * itw; idx += kNlmThreads) {
                const int rr = idx / itw, wc = idx - rr * itw;
                const uint8_t *row = f + (size_t)reflect101_iter(y0 - b + rr, a.H) * a.pitch;
                uint32_t w = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q) w |= (uint32_t)row[reflect101_iter(x0 - b + 4 * wc + q, a.W)] << (8 * q);
                reinterpret_cast<uint32_t *>(s_it)[idx] = w;
            }
            __syncthreads();
        }
        for (int dx = -sh; dx <= sh; ++dx) {
            pass_a(dy, dx, buf);
            if (pending) pass_b(pend_dy, pend_dx, buf ^ 1);
            __syncthreads();
            pend_dy = dy; pend_dx = dx; pending = true;
            buf ^= 1;
        }
    }
    if (pending) pass_b(pend_dy, pend_dx, buf ^ 1);

#pragma unroll
    for (int i = 0; i < K; ++i) {
        const int gy = y0 + rh * K + i, gx = x0 + j;
        if (gy >= a.H || gx >= a.W) continue;
        if (a.nsplit > 1) {
            int *acc = a.acc + (((long long)n * a.H + gy) * a.W + gx) * 2;
            if (wsum[i]) { atomicAdd(acc, est[i]); atomicAdd(acc + 1, wsum[i]); }
        } else {
            const unsigned q = ((unsigned)est[i] + (unsigned)wsum[i] / 2u) / (unsigned)wsum[i];
            a.dst[((long long)n * a.H + gy) * a.pitch + gx] = (uint8_t)min(q, 255u);
        }
    }
}

// nsplit > 1: turn the accumulated (est, wsum) into pixels
__global__ void __launch_bounds__(256) nlm_finish_kernel(const NlmArgs a, long long n_px)
{
    const long long p = (long long)blockIdx.x * 256 + threadIdx.x;
    if (p >= n_px) return;
    const long long row = p / a.W;
    const int x = (int)(p - row * a.W);
    if (a.twin[row / a.H] == 0) return;                          // passed through by nlm_kernel
    const unsigned est = (unsigned)a.acc[2 * p], ws = (unsigned)a.acc[2 * p + 1];
    a.dst[row * a.pitch + x] = (uint8_t)min((est + ws / 2u) / ws, 255u);
}

}  // namespace me
