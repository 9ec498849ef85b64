// h2d_probe.cu - host->device copy rate of one padded 1080p luma plane: pitched 2D copy (source pitch 2056, destination
// pitch 2176) vs one contiguous copy; and the device->host copy of a result slab.
#include <cuda_runtime.h>
#include <stdio.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)
int main() {
    const size_t w = 2056, h = 1216, dp = 2176;
    unsigned char *hsrc, *d, *hdst;
    CK(cudaMallocHost(&hsrc, w * h));
    CK(cudaMallocHost(&hdst, 2300000));
    CK(cudaMalloc(&d, dp * h));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int mode = 0; mode < 3; mode++) {
        float best = 1e9f;
        for (int rep = 0; rep < 20; rep++) {
            CK(cudaEventRecord(e0, st));
            if (mode == 0) CK(cudaMemcpy2DAsync(d, dp, hsrc, w, w, h, cudaMemcpyHostToDevice, st));
            else if (mode == 1) CK(cudaMemcpyAsync(d, hsrc, w * h, cudaMemcpyHostToDevice, st));
            else CK(cudaMemcpyAsync(hdst, d, 2267460, cudaMemcpyDeviceToHost, st));
            CK(cudaEventRecord(e1, st));
            CK(cudaStreamSynchronize(st));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            if (ms < best) best = ms;
        }
        const double bytes = mode == 2 ? 2267460.0 : (double)w * h;
        printf("%s: %.1f us, %.1f GB/s\n", mode == 0 ? "H2D pitched 2D" : mode == 1 ? "H2D contiguous" : "D2H contiguous", best * 1e3, bytes / best / 1e6);
    }
    return 0;
}
