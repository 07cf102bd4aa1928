// Issue-rate probes for the integer SAD op on sm_100a (not part of the product).
// Each kernel runs `iters` iterations of an unrolled body of 32 ops per thread and reports
// thread-instructions / clk / SM from clock64() deltas.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(2); } } while (0)
__device__ __forceinline__ uint32_t vsad4(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }

// V0: one varying register (acc), a and b loop-invariant and shared by all ops
// V1: a_i distinct per accumulator (16 regs), b shared                      (2 fresh regs / op)
// V2: a_i and b_i distinct                                                  (3 fresh regs / op)
// V3: like the search kernel: acc[d][q] += sad(ref[d], cur[q]) 4x4 patch    (outer product reuse)
// V4: scalar IADD3 acc_i = acc_i + a_i + b  (ALU reference)
// V5: IMAD acc_i = a_i * b + acc_i          (FMA-pipe reference)
// V6: V3 plus one IMAD per 4 SAD ops
template <int V>
__global__ void __launch_bounds__(1024, 1) probe(uint32_t* out, int iters, long long* clk, uint32_t seed) {
    uint32_t acc[16], a[16], b[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { acc[i] = i; a[i] = seed * (i + 1) + threadIdx.x; b[i] = seed ^ (i * 0x01010101u); }
    uint32_t z = seed;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 2; ++rep) {
            if (V == 0) {
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[i] = vsad4(a[0], b[0], acc[i]);
            } else if (V == 1) {
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[i] = vsad4(a[i], b[0], acc[i]);
            } else if (V == 2) {
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[i] = vsad4(a[i], b[i], acc[i]);
            } else if (V == 3 || V == 6) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int d = 0; d < 4; ++d) {
                        acc[d * 4 + q] = vsad4(a[d], b[q], acc[d * 4 + q]);
                        if (V == 6 && d == 0) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(z) : "r"(a[5]), "r"(b[5]));
                    }
            } else if (V == 4) {
#pragma unroll
                for (int i = 0; i < 16; ++i) asm volatile("{ .reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2; }" : "+r"(acc[i]) : "r"(a[i]), "r"(b[0]));
            } else if (V == 5) {
#pragma unroll
                for (int i = 0; i < 16; ++i) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(acc[i]) : "r"(a[i]), "r"(b[0]));
            }
        }
    }
    long long t1 = clock64();
    uint32_t s = z;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}
template <int V> void run(const char* name, int sms, int threads, uint32_t* d_out, long long* d_clk) {
    const int iters = 8192, blocks = sms;
    for (int rep = 0; rep < 2; ++rep) { probe<V><<<blocks, threads>>>(d_out, iters, d_clk, 1234u + rep); CK(cudaDeviceSynchronize()); }
    std::vector<long long> c(blocks); CK(cudaMemcpy(c.data(), d_clk, sizeof(long long) * blocks, cudaMemcpyDeviceToHost));
    double m = 0; for (auto v : c) m += (double)v; m /= blocks;
    printf("%-40s threads/SM %4d : %7.2f thread-instr/clk/SM\n", name, threads, (double)threads * 32.0 * iters / m);
}
int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    uint32_t* d_out; long long* d_clk; CK(cudaMalloc(&d_out, 4ull * sms * 1024)); CK(cudaMalloc(&d_clk, 8ull * sms));
    for (int threads : {128, 256, 512, 1024}) {
        run<0>("V0 sad(a,b,acc_i) 1 fresh reg", sms, threads, d_out, d_clk);
        run<1>("V1 sad(a_i,b,acc_i) 2 fresh regs", sms, threads, d_out, d_clk);
        run<2>("V2 sad(a_i,b_i,acc_i) 3 fresh regs", sms, threads, d_out, d_clk);
        run<3>("V3 4x4 outer-product patch", sms, threads, d_out, d_clk);
        run<6>("V6 patch + 1 IMAD per 4 SAD (SAD rate)", sms, threads, d_out, d_clk);
        run<4>("V4 IADD3 reference", sms, threads, d_out, d_clk);
        run<5>("V5 IMAD reference", sms, threads, d_out, d_clk);
    }
    return 0;
}
