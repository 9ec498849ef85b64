// tma_bench.cu - TMA box throughput for the small u8 windows k_hme fetches: cycles per box and bytes/clk/SM as a
// function of box width / rows, with 4 CTAs per SM each keeping `depth` windows in flight.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                             const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__global__ void __launch_bounds__(128) k_bench(const __grid_constant__ CUtensorMap map, int ib, int br, int nbox, int iters, int depth,
                                               int nslots, long long *cycles, unsigned *sink) {
    extern __shared__ __align__(1024) uint8_t buf[];
    __shared__ __align__(8) unsigned long long bar[8];
    const int win = ib * br * nbox;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 8; i++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar[i])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    unsigned rng = blockIdx.x * 2654435761u + 12345u, acc = 0;
    const long long t0 = clock64();
    int issued = 0;
    for (int it = 0; it < iters; it++) {
        if (threadIdx.x < 32) {
            while (issued < iters && issued - it < depth) {
                rng = rng * 1664525u + 1013904223u;
                const int x = ((rng >> 8) % 1700) & ~15, y = (rng >> 20) % 900, z = (rng >> 3) % nslots;
                const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar[issued % 8]);
                const uint32_t d = (uint32_t)__cvta_generic_to_shared(buf + (size_t)(issued % depth) * win);
                if (threadIdx.x == 0)
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"((uint32_t)win) : "memory");
                __syncwarp();
                for (int i = threadIdx.x; i < nbox; i += 32)
                    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                                 ::"r"(d + i * ib * br), "l"(&map), "r"(x), "r"(y + i * br), "r"(z), "r"(b) : "memory");
                issued++;
            }
        }
        const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar[it % 8]);
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(b), "r"((uint32_t)((it / 8) & 1)) : "memory");
        acc += buf[(size_t)(it % depth) * win + threadIdx.x];
        __syncthreads();
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    if (acc == 0xdeadbeef) *sink = acc;
}

int main() {
    const int stride = 2176, rows = 1216, slots = 40;
    const size_t slot_bytes = (size_t)stride * rows + 4096;
    uint8_t *d; long long *dc; unsigned *sink;
    CK(cudaMalloc(&d, slot_bytes * slots));
    CK(cudaMemset(d, 1, slot_bytes * slots));
    const int ctas = 148 * 4;
    CK(cudaMalloc(&dc, ctas * sizeof(long long)));
    CK(cudaMalloc(&sink, 4));
    EncodeFn encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &qres));
    CK(cudaFuncSetAttribute(k_bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 50 * 1024));
    struct V { int ib, br, nbox, depth; } vs[] = {{96, 8, 9, 1}, {96, 8, 9, 4}, {96, 24, 3, 4}, {96, 72, 1, 4}, {96, 72, 1, 1}, {48, 8, 9, 4}, {128, 8, 9, 4},
                                                   {224, 8, 3, 4}, {80, 8, 7, 4}, {80, 56, 1, 4}, {256, 8, 4, 4}, {256, 32, 1, 4}, {64, 8, 5, 4}, {64, 40, 1, 4}};
    for (auto &v : vs) {
        CUtensorMap map;
        cuuint64_t dims[3] = {(cuuint64_t)stride, (cuuint64_t)rows, (cuuint64_t)slots}, strides[2] = {(cuuint64_t)stride, (cuuint64_t)slot_bytes};
        cuuint32_t box[3] = {(cuuint32_t)v.ib, (cuuint32_t)v.br, 1}, estr[3] = {1, 1, 1};
        CUresult r = encode(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode %d\n", (int)r); return 1; }
        const int iters = 200, win = v.ib * v.br * v.nbox;
        for (int rep = 0; rep < 2; rep++) {
            k_bench<<<ctas, 128, win * v.depth>>>(map, v.ib, v.br, v.nbox, iters, v.depth, slots, dc, sink);
            CK(cudaDeviceSynchronize());
        }
        std::vector<long long> c(ctas);
        CK(cudaMemcpy(c.data(), dc, ctas * sizeof(long long), cudaMemcpyDeviceToHost));
        double avg = 0; for (auto x : c) avg += x; avg /= ctas;
        printf("box %3d x %2d rows, %d boxes/window (%5d B), depth %d: %7.0f cycles/window/CTA, %6.1f cycles/box/SM, %5.1f B/clk/SM\n", v.ib, v.br,
               v.nbox, win, v.depth, avg / iters, avg / iters / v.nbox / 4, (double)win * 4 * iters / avg);
    }
    return 0;
}
