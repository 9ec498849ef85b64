// sad_microbench.cu - measures the per-SM issue rates that bound the integer-SAD roofline:
// packed SAD (VABSDIFF4 with accumulate), and its co-issue with IMAD / min / shared-memory loads.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sad_microbench sad_microbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t vsad4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

template <int MODE>
__global__ void __launch_bounds__(256) bench(uint32_t *out, int iters, uint32_t seed) {
    __shared__ uint4 sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = make_uint4(i * seed, i + seed, i ^ seed, seed);
    __syncthreads();
    uint32_t acc[16];
    uint32_t a[4], b[4];
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = i;
#pragma unroll
    for (int i = 0; i < 4; i++) { a[i] = seed * (threadIdx.x + i); b[i] = seed ^ (threadIdx.x * 77 + i); }
    uint32_t m = 0xffffffffu, k = seed;
    const int lane = threadIdx.x & 31;
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) { // 64 vsad4
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int i = 0; i < 16; i++) acc[i] = vsad4(a[i & 3], b[r], acc[i]);
        } else if (MODE == 1) { // 64 vsad4 + 16 IMAD + 16 min
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int i = 0; i < 16; i++) acc[i] = vsad4(a[i & 3], b[r], acc[i]);
#pragma unroll
            for (int i = 0; i < 16; i++) { uint32_t key = acc[i] * 65536u + k; m = min(m, key); }
            k += 3;
        } else if (MODE == 2) { // 64 vsad4 + 4 LDS.128 conflict-free (b from smem each row)
#pragma unroll
            for (int r = 0; r < 4; r++) {
                uint4 v = sm[(threadIdx.x + r * 33 + it) & 1023];
                b[0] = v.x; b[1] = v.y; b[2] = v.z; b[3] = v.w;
#pragma unroll
                for (int i = 0; i < 16; i++) acc[i] = vsad4(a[i & 3], b[i >> 2], acc[i]);
            }
        } else if (MODE == 3) { // 64 vsad4 + 8 LDS.128 (4 distinct + 4 broadcast)
#pragma unroll
            for (int r = 0; r < 4; r++) {
                uint4 v = sm[(threadIdx.x + r * 33 + it) & 1023];
                uint4 w = sm[(r * 5 + it) & 1023];
                b[0] = v.x; b[1] = v.y; b[2] = v.z; b[3] = v.w;
                a[0] = w.x; a[1] = w.y; a[2] = w.z; a[3] = w.w;
#pragma unroll
                for (int i = 0; i < 16; i++) acc[i] = vsad4(a[i & 3], b[i >> 2], acc[i]);
            }
        } else if (MODE == 4) { // 64 x (vabsdiff4 without acc + dp4a)
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int i = 0; i < 16; i++) acc[i] = __dp4a(__vabsdiffu4(a[i & 3], b[r] + i), 0x01010101u, acc[i]);
        } else if (MODE == 5) { // 64 IADD3-ish (full-rate ALU reference)
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int i = 0; i < 16; i++) acc[i] = (acc[i] + a[i & 3]) ^ b[r];
        } else if (MODE == 6) { // 16 redux.min per 64 vsad4
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int i = 0; i < 16; i++) acc[i] = vsad4(a[i & 3], b[r], acc[i]);
#pragma unroll
            for (int i = 0; i < 16; i++) { uint32_t r = __reduce_min_sync(0xffffffffu, acc[i]); if (lane == i) m = min(m, r); }
        } else if (MODE == 7) { // 64 vsad4 + 16 funnel shifts (unaligned operand build)
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    uint32_t x = (i & 3) == 0 ? __funnelshift_r(a[r], b[(i >> 2)], 8 * (seed & 3)) : a[i & 3];
                    acc[i] = vsad4(x, b[r], acc[i]);
                }
        } else if (MODE == 8) { // 64 LDS.128 broadcast only
#pragma unroll
            for (int r = 0; r < 16; r++) {
                uint4 w = sm[(r * 5 + it) & 1023];
                acc[r] += w.x ^ w.y ^ w.z ^ w.w;
            }
        } else if (MODE == 9) { // 16 LDS.128 distinct (conflict-free) only
#pragma unroll
            for (int r = 0; r < 16; r++) {
                uint4 w = sm[(threadIdx.x + r * 37 + it) & 1023];
                acc[r] += w.x ^ w.y ^ w.z ^ w.w;
            }
        }
    }
    uint32_t s = m;
#pragma unroll
    for (int i = 0; i < 16; i++) s += acc[i];
    if (s == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
int run(const char *name, double ops_per_iter, int nsm, double clock_ghz, uint32_t *out) {
    const int iters = 4096;
    for (int wps = 1; wps <= 16; wps *= 2) { // resident CTAs (of 8 warps) per SM: 1 -> 8 warps/SM ...
        if (wps > 8) break;
        dim3 grid(nsm * wps), block(256);
        bench<MODE><<<grid, block>>>(out, 64, 1234567u);
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0));
        bench<MODE><<<grid, block>>>(out, iters, 1234567u);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        double total = ops_per_iter * iters * (double)grid.x * block.x; // thread-level ops
        double rate = total / (ms * 1e-3);                                 // ops / s
        printf("%-34s warps/SM %3d  %8.3f ms  %8.2f Tops/s  %7.2f lane-ops/clk/SM (at %.3f GHz nominal)\n", name, wps * 8, ms,
               rate * 1e-12, rate / nsm / (clock_ghz * 1e9), clock_ghz);
    }
    return 0;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    double ghz = clk_khz * 1e-6;
    printf("device %s sms %d clock %.3f GHz\n", p.name, p.multiProcessorCount, ghz);
    uint32_t *out; CK(cudaMalloc(&out, 148 * 16 * 256 * 4 * 4));
    int n = p.multiProcessorCount;
    run<0>("vsad4 x64", 64, n, ghz, out);
    run<1>("vsad4 x64 + 16 imad + 16 min", 64, n, ghz, out);
    run<2>("vsad4 x64 + 4 lds128", 64, n, ghz, out);
    run<3>("vsad4 x64 + 4 lds128 + 4 bcast", 64, n, ghz, out);
    run<4>("(vabsdiff4 + dp4a) x64", 64, n, ghz, out);
    run<5>("(iadd+xor) x64", 128, n, ghz, out);
    run<6>("vsad4 x64 + 16 redux", 64, n, ghz, out);
    run<7>("vsad4 x64 + 16 shf", 64, n, ghz, out);
    run<8>("lds128 bcast x16", 16, n, ghz, out);
    run<9>("lds128 distinct x16", 16, n, ghz, out);
    return 0;
}
