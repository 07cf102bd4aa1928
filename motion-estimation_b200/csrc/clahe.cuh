// clahe.cuh -- contrast-limited adaptive histogram equalisation on the device frame ring.
//
// Replaces the per-frame cv.createCLAHE(clipLimit=2.0, tileGridSize=(8,8)).apply(z) of the reference's
// adaptive_histogram_equalization (estimate_vel_vid.py:201-219), the step right before the pair loop.
// Bit-exact with OpenCV's 8-bit CLAHE (restated in oracle/clahe_np.py and pinned against cv2 there):
//   clahe_lut_kernel    one CTA per (tile, frame): 256-bin histogram in shared memory (reflect-101
//                       extension when the frame does not divide into the tile grid), clip, batch +
//                       strided-residual redistribution, inclusive scan, LUT = round(cumsum * 255 / area)
//   clahe_apply_kernel  4 pixels per thread: bilinear blend of the four surrounding tile LUTs in binary32
//                       with OpenCV's operation order (no FMA contraction), one rounding, in place.
// Both are HBM-bound streaming kernels: 3 bytes of traffic per pixel (read, read, write) + 16 KB of LUTs per frame.
#pragma once
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
    __shared__ unsigned whist[8][256];                   // one histogram per warp: an eighth of the atomic contention
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
    if (xe <= a.W) {
        // one warp per row, four pixels per load: aligned words that cover [xb, xe), bytes outside it masked off
        const int w0 = xb >> 2, w1 = (xe + 3) >> 2;
        for (int r = warp; r < a.th; r += 8) {
            const uint32_t *row = reinterpret_cast<const uint32_t *>(f + (size_t)reflect101(tj * a.th + r, a.H) * a.pitch);
            for (int w = w0 + lane; w < w1; w += 32) {
                const uint32_t v = __ldg(row + w);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int x = 4 * w + q;
                    if (x >= xb && x < xe) atomicAdd(&whist[warp][(v >> (8 * q)) & 255u], 1u);
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

// Rows per CTA of clahe_apply_kernel and the largest tile grid whose lookup tables it keeps in shared memory.
constexpr int kClaheRows = 8, kClaheSmemTiles = 64;

// One CTA = kClaheRows rows of one frame, 16 pixels (one 16-byte load and store) per thread and step.  The frame's
// lookup tables (16 KB for the 8x8 grid) are staged in shared memory once per CTA, so the four table reads per pixel
// are shared-memory byte loads instead of scattered global ones (r01: the global form ran at 0.75 TB/s).
__global__ void __launch_bounds__(256) clahe_apply_kernel(const ClaheArgs a, int first_slot)
{
    __shared__ __align__(16) uint8_t s_lut[kClaheSmemTiles * 256];
    const int frame = blockIdx.y, y_lo = blockIdx.x * kClaheRows, tid = threadIdx.x;
    uint8_t *fbase = a.frames + (long long)(first_slot + frame) * a.frame_stride;
    const uint8_t *glut = a.luts + (size_t)frame * a.ty * a.tx * 256;
    const int ntile = a.tx * a.ty;
    const uint8_t *lut = glut;
    if (ntile <= kClaheSmemTiles) {
        const uint4 *src = reinterpret_cast<const uint4 *>(glut);
        for (int i = tid; i < ntile * 16; i += 256) reinterpret_cast<uint4 *>(s_lut)[i] = __ldg(src + i);
        __syncthreads();
        lut = s_lut;
    }
    const int chunks = (a.W + 15) >> 4;                                   // 16-pixel chunks per row
    for (int it = tid; it < kClaheRows * chunks; it += 256) {
        const int ry = it / chunks, cx = it - ry * chunks, y = y_lo + ry, x16 = cx << 4;
        if (y >= a.H) break;
        uint8_t *row = fbase + (size_t)y * a.pitch;
        const float tyf = __fsub_rn(__fmul_rn((float)y, a.inv_th), 0.5f);
        int ty1 = (int)floorf(tyf);
        const float ya = __fsub_rn(tyf, (float)ty1), ya1 = __fsub_rn(1.0f, ya);
        const int ty2 = min(ty1 + 1, a.ty - 1);
        ty1 = max(ty1, 0);
        const uint8_t *l1 = lut + (size_t)ty1 * a.tx * 256, *l2 = lut + (size_t)ty2 * a.tx * 256;
        // one pixel: bilinear blend of the four surrounding tables, OpenCV's operation order, one rounding
        auto pixel = [&](int x, int v) {
            const float txf = __fsub_rn(__fmul_rn((float)x, a.inv_tw), 0.5f);
            int tx1 = (int)floorf(txf);
            const float xa = __fsub_rn(txf, (float)tx1), xa1 = __fsub_rn(1.0f, xa);
            const int tx2 = min(tx1 + 1, a.tx - 1);
            tx1 = max(tx1, 0);
            const float top = __fadd_rn(__fmul_rn((float)l1[tx1 * 256 + v], xa1), __fmul_rn((float)l1[tx2 * 256 + v], xa));
            const float bot = __fadd_rn(__fmul_rn((float)l2[tx1 * 256 + v], xa1), __fmul_rn((float)l2[tx2 * 256 + v], xa));
            const int q = __float2int_rn(__fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya)));
            return (uint32_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
        };
        if (x16 + 16 <= a.W) {
            const uint4 v = *reinterpret_cast<const uint4 *>(row + x16);
            const uint32_t in[4] = {v.x, v.y, v.z, v.w};
            uint32_t out[4] = {0, 0, 0, 0};
#pragma unroll
            for (int i = 0; i < 16; ++i) out[i >> 2] |= pixel(x16 + i, (int)((in[i >> 2] >> (8 * (i & 3))) & 255u)) << (8 * (i & 3));
            *reinterpret_cast<uint4 *>(row + x16) = make_uint4(out[0], out[1], out[2], out[3]);
        } else {
            for (int x = x16; x < a.W; ++x) row[x] = (uint8_t)pixel(x, row[x]);
        }
    }
}

} // namespace me
