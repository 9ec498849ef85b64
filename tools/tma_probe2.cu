// tma_probe2.cu - narrows down which tensor-map shapes the TMA unit accepts (one variant per process: argv[1])
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

__global__ void k2(const __grid_constant__ CUtensorMap map, int rank, int bytes, int x, int y, int z, uint8_t *out) {
    extern __shared__ __align__(1024) uint8_t buf[];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(&bar);
    const uint32_t dst   = (uint32_t)__cvta_generic_to_shared(buf);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"((uint32_t)bytes) : "memory");
        if (rank == 2)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(dst), "l"(&map), "r"(x), "r"(y), "r"(bar_a) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dst), "l"(&map), "r"(x), "r"(y), "r"(z), "r"(bar_a) : "memory");
    }
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(bar_a), "r"(0u) : "memory");
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) out[i] = buf[i];
}

int main(int argc, char **argv) {
    const int v = argc > 1 ? atoi(argv[1]) : 0;
    const int stride = 2176, rows = 1216, slots = 3;
    const size_t slot_bytes = (size_t)stride * rows + 4096;
    std::vector<uint8_t> h(slot_bytes * slots);
    srand(1);
    for (auto &b : h) b = (uint8_t)rand();
    uint8_t *d, *dout;
    CK(cudaMalloc(&d, h.size()));
    CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dout, 65536));
    EncodeFn encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &qres));
    // variants: rank, inner box bytes, box rows, dtype (0 u8, 1 u32), x, y
    struct V { int rank, ib, br, dt, x, y; } vs[] = {
        {2, 32, 8, 0, 0, 0}, {2, 64, 8, 0, 0, 0}, {2, 128, 16, 0, 16, 3}, {3, 32, 8, 0, 0, 0}, {2, 128, 16, 1, 0, 0},
        {2, 32, 8, 0, 5, 7}, {3, 64, 8, 0, 5, 7}, {2, 256, 8, 0, 1, 1}, {2, 16, 8, 0, 1, 1}, {2, 208, 8, 0, 3, 1}, {2, 80, 8, 0, 3, 1}, {2, 48, 8, 0, 3, 1}};
    const V &q = vs[v];
    const int es = q.dt ? 4 : 1;
    CUtensorMap map;
    cuuint64_t  dims[3] = {(cuuint64_t)stride / es, (cuuint64_t)rows, (cuuint64_t)slots};
    cuuint64_t  strides[2] = {(cuuint64_t)stride, (cuuint64_t)slot_bytes};
    cuuint32_t  box[3] = {(cuuint32_t)(q.ib / es), (cuuint32_t)q.br, 1}, estr[3] = {1, 1, 1};
    CUresult r = encode(&map, q.dt ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, q.rank, d, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("variant %d: rank %d box %dx%d dtype %s at (%d,%d): encode rc %d; ", v, q.rank, q.ib, q.br, q.dt ? "u32" : "u8", q.x, q.y, (int)r);
    if (r != CUDA_SUCCESS) { printf("\n"); return 1; }
    const int bytes = q.ib * q.br;
    k2<<<1, 128, 16384>>>(map, q.rank, bytes, q.x, q.y, 1, dout);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s\n", cudaGetErrorString(e)); return 1; }
    std::vector<uint8_t> o(bytes);
    CK(cudaMemcpy(o.data(), dout, bytes, cudaMemcpyDeviceToHost));
    int nb = 0;
    for (int y = 0; y < q.br; y++)
        for (int x = 0; x < q.ib; x++)
            nb += o[y * q.ib + x] != h[(q.rank == 3 ? slot_bytes : 0) + (size_t)(q.y + y) * stride + q.x * es + x];
    printf("%d mismatches\n", nb);
    return nb != 0;
}
