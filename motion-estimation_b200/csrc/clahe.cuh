// clahe.cuh -- contrast-limited adaptive histogram equalisation on the device frame ring.
//
// Replaces the per-frame cv.createCLAHE(clipLimit=2.0, tileGridSize=(8,8)).apply(z) of the reference's
// adaptive_histogram_equalization (estimate_vel_vid.py:201-219), the step right before the pair loop.
// Bit-exact with OpenCV's 8-bit CLAHE (restated in oracle/clahe_np.py and pinned against cv2 there):
//   clahe_lut_kernel    one CTA per (tile, frame): 256-bin histogram in shared memory (reflect-101
//                       extension when the frame does not divide into the tile grid), clip, batch +
//                       strided-residual redistribution, inclusive scan, LUT = round(cumsum * 255 / area)
//   clahe_pack_kernel   the four tables of every interpolation cell packed into one word per grey level
//   clahe_apply_kernel  bilinear blend of the four surrounding tile LUTs in binary32 with OpenCV's operation
//                       order (no FMA contraction), one rounding, in place; one table word per pixel.
// Both are HBM-bound streaming kernels: 3 bytes of traffic per pixel (read, read, write) + 16 KB of LUTs per frame.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace me {

struct ClaheArgs {
    uint8_t *frames;            // ring base
    long long frame_stride;
    int pitch, W, H;
    int tx, ty;                 // tile grid
    int tw, th;                 // tile size in the (possibly extended) frame
    int clip;                   // 0 = no clipping
    float lut_scale;            // 255 / (tw*th)
    float inv_tw, inv_th;
    uint8_t *luts;              // [n_frames][ty][tx][256]
};

__device__ __forceinline__ int reflect101(int i, int n)
{
    i = i >= n ? 2 * (n - 1) - i : i;
    return i < 0 ? -i : i;
}

__global__ void __launch_bounds__(256) clahe_lut_kernel(const ClaheArgs a, int first_slot)
{
    // one histogram per warp: an eighth of the atomic contention (four copies per warp, by lane quarter, were tried and lost:
    // 68 -> 80 us, r02h -- the extra clearing and summing costs more than the same-address atomics it avoids)
    __shared__ __align__(1024) unsigned whist[8][256];          // 1 KB per warp, 1 KB aligned: bin address = base | (byte << 2)
    __shared__ unsigned wsum[8];
    __shared__ unsigned s_clipped;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tj = blockIdx.x / a.tx, ti = blockIdx.x - tj * a.tx, frame = blockIdx.y;
    const uint8_t *f = a.frames + (long long)(first_slot + frame) * a.frame_stride;
#pragma unroll
    for (int w = 0; w < 8; ++w) whist[w][tid] = 0;
    if (tid == 0) s_clipped = 0;
    __syncthreads();
    const int xb = ti * a.tw, xe = xb + a.tw;             // the tile's columns; beyond W only in the reflect-extended case
    const unsigned hbase = (unsigned)__cvta_generic_to_shared(&whist[warp][0]);
    // one histogram update: shift the byte to bits 2..9, mask and merge the warp's base in one LOP3, one shared-memory reduction
    auto bump = [&](uint32_t v, int q) {
        const unsigned sh = q == 0 ? v << 2 : v >> (8 * q - 2);
        asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(hbase | (sh & 0x3fcu)) : "memory");
    };
    if (xe <= a.W) {
        // four pixels per load: aligned words that cover [xb, xe), bytes outside it masked off.  The tile's (row, word) pairs
        // are spread over the whole CTA and four loads are in flight per thread before the first histogram update (r02a ncu
        // of the one-row-per-warp form: 8.5 long-scoreboard stalls per issued instruction, 1.5 TB/s).
        const int w0 = xb >> 2, w1 = (xe + 3) >> 2, nw = w1 - w0, total = a.th * nw;
        // (row, word) of this thread's first pair, then steps of 256 pairs: incremental, no division per load
        const int step_r = 256 / nw, step_c = 256 - step_r * nw;
        int r = tid / nw, cw = tid - r * nw;
        const bool inside = (tj + 1) * a.th <= a.H;                      // no reflected rows in this tile (all but a ragged last row of tiles)
        const bool aligned = ((xb | xe) & 3) == 0;                       // tile edges on word boundaries (the usual case): no per-byte tests
        int p0 = tid;
        if (inside && aligned) {
            // the usual tile: byte offset of the thread's word kept incrementally (r02 SASS of the general loop: 28 instructions
            // of address arithmetic and guards per load, 50 per word in all), no guards until the last partial round
            unsigned off = (unsigned)(tj * a.th + r) * (unsigned)a.pitch + 4u * (unsigned)(w0 + cw);
            const unsigned d_step = (unsigned)step_r * (unsigned)a.pitch + 4u * (unsigned)step_c;
            const unsigned d_wrap = (unsigned)a.pitch - 4u * (unsigned)nw;
            for (; p0 + 3 * 256 < total; p0 += 4 * 256) {
                uint32_t v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    v[u] = __ldg(reinterpret_cast<const uint32_t *>(f + off));
                    off += d_step; cw += step_c;
                    if (cw >= nw) { cw -= nw; off += d_wrap; }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
#pragma unroll
                    for (int q = 0; q < 4; ++q) bump(v[u], q);
            }
            for (; p0 < total; p0 += 256) {
                const uint32_t v = __ldg(reinterpret_cast<const uint32_t *>(f + off));
                off += d_step; cw += step_c;
                if (cw >= nw) { cw -= nw; off += d_wrap; }
#pragma unroll
                for (int q = 0; q < 4; ++q) bump(v, q);
            }
        } else {
            for (; p0 < total; p0 += 4 * 256) {
                uint32_t v[4];
                int wi[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    wi[u] = -1;
                    if (p0 + u * 256 < total) {
                        wi[u] = w0 + cw;
                        const int y = inside ? tj * a.th + r : reflect101(tj * a.th + r, a.H);
                        v[u] = __ldg(reinterpret_cast<const uint32_t *>(f + (size_t)y * a.pitch) + wi[u]);
                    }
                    cw += step_c; r += step_r;
                    if (cw >= nw) { cw -= nw; ++r; }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (wi[u] < 0) continue;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int x = 4 * wi[u] + q;
                        if (aligned || (x >= xb && x < xe)) bump(v[u], q);
                    }
                }
            }
        }
    } else {
        const int area = a.tw * a.th;
        for (int p = tid; p < area; p += 256) {
            const int r = p / a.tw, c = p - r * a.tw;
            const int y = reflect101(tj * a.th + r, a.H), x = reflect101(xb + c, a.W);
            atomicAdd(&whist[warp][f[(size_t)y * a.pitch + x]], 1u);
        }
    }
    __syncthreads();
    unsigned h = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) h += whist[w][tid];
    if (a.clip > 0) {
        unsigned excess = h > (unsigned)a.clip ? h - (unsigned)a.clip : 0u;
        h = h > (unsigned)a.clip ? (unsigned)a.clip : h;
        excess = __reduce_add_sync(0xffffffffu, excess);
        if (lane == 0 && excess) atomicAdd(&s_clipped, excess);
        __syncthreads();
        const unsigned clipped = s_clipped, batch = clipped >> 8, resid = clipped - (batch << 8);
        h += batch;
        if (resid) {
            const unsigned step = 256u / resid > 1u ? 256u / resid : 1u;
            if (tid % step == 0 && tid / step < resid) h += 1u;
        }
    }
    // inclusive scan of the 256 bins
    unsigned incl = h;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    unsigned base = 0;
    for (int w = 0; w < warp; ++w) base += wsum[w];
    const float v = __fmul_rn((float)(base + incl), a.lut_scale);
    int q = __float2int_rn(v);                                        // cvRound: round half to even
    q = q < 0 ? 0 : (q > 255 ? 255 : q);
    a.luts[((size_t)(frame * a.ty + tj) * a.tx + ti) * 256 + tid] = (uint8_t)q;
}

// ---------------------------------------------------------------- packed interpolation tables
// A pixel blends the tables of the four tiles around it.  All pixels of one interpolation CELL (the rectangle between
// four tile centres; (tx + 1) x (ty + 1) cells per frame) use the same four tables, so those are packed once per frame
// into one 8-byte entry per grey level: (top-left, top-right, bottom-left, bottom-right) as binary16.  The apply kernel then
// makes ONE 64-bit lookup per pixel instead of four byte lookups and four integer conversions (r02a ncu of the byte form: 49 M
// shared wavefronts, 31 M of them bank conflicts, 48 thread-instructions per pixel).
__global__ void __launch_bounds__(256) clahe_pack_kernel(const ClaheArgs a, uint2 *packed)
{
    const int cells_x = a.tx + 1, cy = blockIdx.x / cells_x, cx = blockIdx.x - cy * cells_x, frame = blockIdx.y, v = threadIdx.x;
    const int tx1 = max(cx - 1, 0), tx2 = min(cx, a.tx - 1), ty1 = max(cy - 1, 0), ty2 = min(cy, a.ty - 1);
    const uint8_t *l = a.luts + (size_t)frame * a.ty * a.tx * 256;
    // as binary16 (0..255 are exact): half -> float is one full-rate instruction in the apply kernel, where byte -> float costs
    // either the 16-lane conversion unit or a PRMT on the ALU pipe that kernel is bound by; 16-byte binary32 entries were tried
    // and lost (r02f: 185 us against 155 us, four-bank lookups and a 72 KB table to stage per CTA)
    const __half2 top = __floats2half2_rn((float)l[(ty1 * a.tx + tx1) * 256 + v], (float)l[(ty1 * a.tx + tx2) * 256 + v]);
    const __half2 bot = __floats2half2_rn((float)l[(ty2 * a.tx + tx1) * 256 + v], (float)l[(ty2 * a.tx + tx2) * 256 + v]);
    packed[((size_t)frame * (a.ty + 1) * cells_x + blockIdx.x) * 256 + v] =
        make_uint2(*reinterpret_cast<const unsigned *>(&top), *reinterpret_cast<const unsigned *>(&bot));
}

// Rows per CTA of clahe_apply_kernel; the packed tables of the (at most two) cell rows those image rows touch are staged
// in shared memory when they fit (kClaheSmemCells cells; the reference's 8x8 grid needs 2 x 9).
constexpr int kClaheRows = 16, kClaheSmemCells = 20;

// One CTA = kClaheRows rows of one frame.  A thread owns one 4-pixel word column at a time and walks the rows: the
// x-dependent interpolation terms (OpenCV computes them per column: cell index, xa, 1 - xa) are computed once per word
// column, the y-dependent ones once per row and CTA, and a pixel costs one table word, four conversions, nine binary32
// operations in OpenCV's order (no FMA contraction) and one rounding.
__global__ void __launch_bounds__(256) clahe_apply_kernel(const ClaheArgs a, int first_slot, const uint2 *__restrict__ packed)
{
    __shared__ __align__(16) uint2 s_tab[kClaheSmemCells * 256];                        // 40 KB
    __shared__ float s_ya[kClaheRows], s_ya1[kClaheRows];
    __shared__ int s_cy[kClaheRows];
    const int frame = blockIdx.y, y_lo = blockIdx.x * kClaheRows, tid = threadIdx.x;
    const int cells_x = a.tx + 1;
    uint8_t *fbase = a.frames + (long long)(first_slot + frame) * a.frame_stride;
    const uint2 *gtab = packed + (size_t)frame * (a.ty + 1) * cells_x * 256;
    const int rows = min(kClaheRows, a.H - y_lo);
    if (tid < rows) {
        const float tyf = __fsub_rn(__fmul_rn((float)(y_lo + tid), a.inv_th), 0.5f);
        const int t1 = (int)floorf(tyf);
        const float ya = __fsub_rn(tyf, (float)t1);
        s_ya[tid] = ya; s_ya1[tid] = __fsub_rn(1.0f, ya); s_cy[tid] = t1 + 1;          // cell row 0 .. ty
    }
    __syncthreads();
    const int cy_lo = s_cy[0], n_cy = s_cy[rows - 1] - cy_lo + 1;                       // 1 or 2 cell rows (more only for tiles under 8 rows)
    const bool staged = n_cy * cells_x <= kClaheSmemCells;
    if (staged) {
        const uint4 *src = reinterpret_cast<const uint4 *>(gtab + (size_t)cy_lo * cells_x * 256);
        for (int i = tid; i < n_cy * cells_x * 128; i += 256) reinterpret_cast<uint4 *>(s_tab)[i] = __ldg(src + i);
        __syncthreads();
    }
    // The blend in binary32, OpenCV's operation order (no FMA contraction).  The result becomes a byte without the conversion
    // unit (16 lanes per clock and SM; the I2F / F2I form was bound by it: r02b, 191 us per 64 frames): adding 1.5 * 2^23 rounds
    // to the nearest even integer exactly as cvRound / __float2int_rn does and leaves it in the low mantissa bits.  0 <= v <
    // 255.5 by construction (table entries are 0..255, xa + (1 - xa) <= 1 + 2^-24), so no clamp is needed.
    auto blend = [](uint2 t, float xa, float xa1, float ya, float ya1) {
        const float2 tt = __half22float2(*reinterpret_cast<const __half2 *>(&t.x)), tb = __half22float2(*reinterpret_cast<const __half2 *>(&t.y));
        const float top = __fadd_rn(__fmul_rn(tt.x, xa1), __fmul_rn(tt.y, xa));
        const float bot = __fadd_rn(__fmul_rn(tb.x, xa1), __fmul_rn(tb.y, xa));
        const float v = __fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya));
        return __float_as_uint(__fadd_rn(v, 12582912.0f)) - 0x4B400000u;
    };
    // the same loop twice, so that the staged form reads the table with shared-memory loads (a pointer chosen at run time
    // compiles to generic loads): `tab` is indexed [(cy - cy_lo) * cells_x + cx][v]
    auto run = [&](const uint2 *tab) {
        const int words = a.W >> 2;
        for (int cw = tid; cw < words; cw += 256) {
            float xa[4], xa1[4];
            int cell[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float txf = __fsub_rn(__fmul_rn((float)(4 * cw + i), a.inv_tw), 0.5f);
                const int t1 = (int)floorf(txf);
                xa[i] = __fsub_rn(txf, (float)t1); xa1[i] = __fsub_rn(1.0f, xa[i]);
                cell[i] = (t1 + 1) * 256;
            }
            uint32_t in[kClaheRows];                             // all rows' words in flight before the first one is needed
            uint32_t *p0 = reinterpret_cast<uint32_t *>(fbase + (size_t)y_lo * a.pitch) + cw;
#pragma unroll
            for (int r = 0; r < kClaheRows; ++r)
                if (r < rows) in[r] = *reinterpret_cast<const uint32_t *>(reinterpret_cast<const uint8_t *>(p0) + (size_t)r * a.pitch);
#pragma unroll
            for (int r = 0; r < kClaheRows; ++r) {
                if (r >= rows) break;
                const float ya = s_ya[r], ya1 = s_ya1[r];
                const uint2 *trow = tab + (size_t)(s_cy[r] - cy_lo) * cells_x * 256;
                uint32_t q[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) q[i] = blend(trow[cell[i] + ((in[r] >> (8 * i)) & 255u)], xa[i], xa1[i], ya, ya1);
                const uint32_t out = __byte_perm(__byte_perm(q[0], q[1], 0x0040), __byte_perm(q[2], q[3], 0x0040), 0x5410);
                *reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(p0) + (size_t)r * a.pitch) = out;
            }
        }
        for (int x = 4 * words + (tid & 3); x < a.W && tid < 4 * rows; x += 4) {       // ragged right edge (W not a multiple of 4)
            const int r = tid >> 2;
            const float txf = __fsub_rn(__fmul_rn((float)x, a.inv_tw), 0.5f);
            const int t1 = (int)floorf(txf);
            const float xa = __fsub_rn(txf, (float)t1);
            uint8_t *px = fbase + (size_t)(y_lo + r) * a.pitch + x;
            const uint2 *trow = tab + (size_t)(s_cy[r] - cy_lo) * cells_x * 256;
            *px = (uint8_t)blend(trow[(t1 + 1) * 256 + *px], xa, __fsub_rn(1.0f, xa), s_ya[r], s_ya1[r]);
        }
    };
    if (staged) run(s_tab);
    else run(gtab + (size_t)cy_lo * cells_x * 256);
}

} // namespace me
