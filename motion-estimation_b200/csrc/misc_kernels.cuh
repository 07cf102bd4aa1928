// misc_kernels.cuh -- the two small kernels outside the search: BGR -> luma of the ingest path and the
// VABSDIFF4 issue-rate microbenchmark behind the roofline denominator.  Included by me_api.cu only.
#pragma once
#include "sad_search.cuh"

namespace me {

// ---------------------------------------------------------------- BGR -> luma (ingest, a9)
// cv.cvtColor(BGR2GRAY) for 8-bit: Y = (B*3735 + G*19235 + R*9798 + 16384) >> 15 (cv2 4.13, checked in tests).
__global__ void bgr2gray_kernel(const uint8_t *__restrict__ bgr, size_t bgr_pitch, uint8_t *__restrict__ gray,
                                size_t gray_pitch, int W, int rows)
{
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y = blockIdx.y;
    if (x4 >= W || y >= rows) return;
    const uint8_t *p = bgr + (size_t)y * bgr_pitch + (size_t)x4 * 3;
    uint8_t o[4] = {0, 0, 0, 0};
    const int n = min(4, W - x4);
    for (int i = 0; i < n; ++i)
        o[i] = (uint8_t)((p[3 * i] * 3735u + p[3 * i + 1] * 19235u + p[3 * i + 2] * 9798u + 16384u) >> 15);
    uint8_t *g = gray + (size_t)y * gray_pitch + x4;
    if (n == 4) *reinterpret_cast<uchar4 *>(g) = make_uchar4(o[0], o[1], o[2], o[3]);
    else for (int i = 0; i < n; ++i) g[i] = o[i];
}

// ---------------------------------------------------------------- issue-rate microbenchmark
// Register-only 4x4 patch of VABSDIFF4.U8.ACC (the search kernel's operand pattern), one
// 1024-thread CTA per SM.  MODE 1 adds one LDS.128 per 16 SAD ops (the kernel's load mix).
template <int MODE>
__global__ void __launch_bounds__(1024, 1) microbench_kernel(uint32_t *out, int iters, long long *clk, uint32_t seed)
{
    __shared__ uint4 sbuf[1024];
    sbuf[threadIdx.x] = make_uint4(seed, seed * 3u, seed * 5u, seed * 7u);
    __syncthreads();
    uint32_t acc[16], a[4], b[4];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = i;
#pragma unroll
    for (int i = 0; i < 4; ++i) { a[i] = seed * (i + 1) + threadIdx.x; b[i] = seed ^ (i * 0x01010101u); }
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int d = 0; d < 4; ++d) acc[d * 4 + q] = vsad4(a[d], b[q], acc[d * 4 + q]);
        if (MODE == 1) {
            const uint4 v = sbuf[(threadIdx.x + it) & 1023];
            b[0] ^= v.x; b[1] += v.y; b[2] ^= v.z; b[3] += v.w;
        }
    }
    const long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

} // namespace me
