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
    __shared__ unsigned hist[256];
    __shared__ unsigned wsum[8];
    __shared__ unsigned s_clipped;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tj = blockIdx.x / a.tx, ti = blockIdx.x - tj * a.tx, frame = blockIdx.y;
    const uint8_t *f = a.frames + (long long)(first_slot + frame) * a.frame_stride;
    hist[tid] = 0;
    if (tid == 0) s_clipped = 0;
    __syncthreads();
    const int area = a.tw * a.th;
    for (int p = tid; p < area; p += 256) {
        const int r = p / a.tw, c = p - r * a.tw;
        const int y = reflect101(tj * a.th + r, a.H), x = reflect101(ti * a.tw + c, a.W);
        atomicAdd(&hist[f[(size_t)y * a.pitch + x]], 1u);
    }
    __syncthreads();
    unsigned h = hist[tid];
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

__global__ void __launch_bounds__(128) clahe_apply_kernel(const ClaheArgs a, int first_slot)
{
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y, frame = blockIdx.z;
    if (x4 >= a.W) return;
    uint8_t *row = a.frames + (long long)(first_slot + frame) * a.frame_stride + (size_t)y * a.pitch;
    const uint8_t *lut = a.luts + (size_t)frame * a.ty * a.tx * 256;
    const float tyf = __fsub_rn(__fmul_rn((float)y, a.inv_th), 0.5f);
    int ty1 = (int)floorf(tyf);
    const float ya = __fsub_rn(tyf, (float)ty1), ya1 = __fsub_rn(1.0f, ya);
    const int ty2 = min(ty1 + 1, a.ty - 1);
    ty1 = max(ty1, 0);
    const uint8_t *l1 = lut + (size_t)ty1 * a.tx * 256, *l2 = lut + (size_t)ty2 * a.tx * 256;
    const int n = min(4, a.W - x4);
    uint8_t in[4] = {0, 0, 0, 0}, out[4];
    if (n == 4) { const uchar4 v = *reinterpret_cast<const uchar4 *>(row + x4); in[0] = v.x; in[1] = v.y; in[2] = v.z; in[3] = v.w; }
    else for (int i = 0; i < n; ++i) in[i] = row[x4 + i];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float txf = __fsub_rn(__fmul_rn((float)(x4 + i), a.inv_tw), 0.5f);
        int tx1 = (int)floorf(txf);
        const float xa = __fsub_rn(txf, (float)tx1), xa1 = __fsub_rn(1.0f, xa);
        const int tx2 = min(tx1 + 1, a.tx - 1);
        tx1 = max(tx1, 0);
        const int v = in[i];
        const float top = __fadd_rn(__fmul_rn((float)l1[tx1 * 256 + v], xa1), __fmul_rn((float)l1[tx2 * 256 + v], xa));
        const float bot = __fadd_rn(__fmul_rn((float)l2[tx1 * 256 + v], xa1), __fmul_rn((float)l2[tx2 * 256 + v], xa));
        int q = __float2int_rn(__fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya)));
        out[i] = (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
    }
    if (n == 4) *reinterpret_cast<uchar4 *>(row + x4) = make_uchar4(out[0], out[1], out[2], out[3]);
    else for (int i = 0; i < n; ++i) row[x4 + i] = out[i];
}

} // namespace me
