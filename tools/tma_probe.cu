// tma_probe.cu - checks the TMA features k_hme relies on: a rank-3 u8 tensor map over (x, y, slot), boxes fetched at
// arbitrary BYTE x coordinates (the four byte-shifted window copies), several lanes issuing boxes onto one mbarrier.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o tma_probe tma_probe.cu   (no -lcuda: entry point via the runtime)
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

constexpr int IB = 32, BR = 8, NCOPY = 4, NBOX = 3; // 4 shifted copies of a 32-byte x 24-row window

__global__ void k_probe(const __grid_constant__ CUtensorMap map, const CUtensorMap *gmap, int mode, int x0, int y0, int slot, uint8_t *out) {
    __shared__ __align__(128) uint8_t buf[NCOPY * NBOX * BR * IB];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        if (threadIdx.x == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"((mode & 4) ? 0u : (uint32_t)sizeof(buf)) : "memory");
        __syncwarp();
        const CUtensorMap *mp = (mode & 2) ? gmap : &map;
        const int i0 = (mode & 1) ? threadIdx.x : (threadIdx.x == 0 ? 0 : NCOPY * NBOX), di = (mode & 1) ? 32 : 1;
        if (mode & 8) {
            if (threadIdx.x == 0)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"((uint32_t)__cvta_generic_to_shared(buf)), "l"(out + 4096), "r"((uint32_t)sizeof(buf)), "r"(bar_a) : "memory");
        } else if (!(mode & 4))
        for (int i = i0; i < NCOPY * NBOX; i += di) {
            const int      c = i / NBOX, bx = i % NBOX;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(buf + (size_t)i * BR * IB);
            asm volatile(
                "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                ::"r"(dst), "l"(mp), "r"(x0 + c), "r"(y0 + bx * BR), "r"(slot), "r"(bar_a) : "memory");
        }
    }
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(bar_a), "r"(0u) : "memory");
    for (int i = threadIdx.x; i < (int)sizeof(buf); i += blockDim.x) out[i] = buf[i];
}

int main(int argc, char **argv) {
    const int only = argc > 1 ? atoi(argv[1]) : -1, mode = argc > 2 ? atoi(argv[2]) : 1;
    const int stride = 2176, rows = 1216, slots = 3;
    const size_t slot_bytes = (size_t)stride * rows + 4096, off = 256;
    std::vector<uint8_t> h(slot_bytes * slots);
    srand(1);
    for (auto &v : h) v = (uint8_t)rand();
    uint8_t *d, *dout;
    CK(cudaMalloc(&d, h.size()));
    CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dout, 4096 + NCOPY * NBOX * BR * IB));
    EncodeFn encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &qres));
    if (!encode || qres != cudaDriverEntryPointSuccess) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    CUtensorMap map;
    cuuint64_t  dims[3] = {(cuuint64_t)stride, (cuuint64_t)rows, (cuuint64_t)slots};
    cuuint64_t  strides[2] = {(cuuint64_t)stride, (cuuint64_t)slot_bytes};
    cuuint32_t  box[3] = {IB, BR, 1}, estr[3] = {1, 1, 1};
    CUresult r = encode(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d + off, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
    CUtensorMap *gmap;
    CK(cudaMalloc(&gmap, sizeof map));
    CK(cudaMemcpy(gmap, &map, sizeof map, cudaMemcpyHostToDevice));
    int bad = 0;
    const int cases[4][3] = {{5, 7, 1}, {0, 0, 0}, {2141, 1190, 2}, {1023, 601, 1}}; // incl. a box hanging over the right/bottom edge
    int ci = 0;
    for (auto &cs : cases) {
        if (only >= 0 && ci++ != only) continue;
        k_probe<<<1, 128>>>(map, gmap, mode, cs[0], cs[1], cs[2], dout);
        CK(cudaDeviceSynchronize());
        std::vector<uint8_t> o(NCOPY * NBOX * BR * IB);
        CK(cudaMemcpy(o.data(), dout, o.size(), cudaMemcpyDeviceToHost));
        int nb = 0;
        for (int c = 0; c < NCOPY; c++)
            for (int y = 0; y < NBOX * BR; y++)
                for (int x = 0; x < IB; x++) {
                    const int gx = cs[0] + c + x, gy = cs[1] + y;
                    const uint8_t want = (gx < stride && gy < rows) ? h[off + slot_bytes * cs[2] + (size_t)gy * stride + gx] : 0;
                    nb += o[(c * NBOX * BR + y) * IB + x] != want;
                }
        printf("case x %d y %d slot %d: %d mismatches\n", cs[0], cs[1], cs[2], nb);
        bad += nb;
    }
    printf(bad ? "TMA PROBE FAILED\n" : "TMA PROBE OK\n");
    return bad != 0;
}
