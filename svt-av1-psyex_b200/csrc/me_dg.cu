/*
 * me_dg.cu - k_dg_hme: the dynamic-GOP detector's HME level 0 (TASK_DG_DETECTOR_HME) for one
 * (source, reference) pair of registered pictures.
 *
 * Reference (Source/Lib/Codec): early_hme_b64 (pd_process.c:393-484) = clamp of a centred search area on the
 * sixteenth plane + one full-SAD svt_sad_loop_kernel of the b64's 16x16 sixteenth block over it;
 * dg_detector_hme_level0 (pd_process.c:492-560) = that for every b64 and the reduction to DGDetectorMetrics.
 *
 * One CTA per b64. The search window (at most 143 x 143 bytes) is staged once in shared memory, re-aligned so
 * that search position x = 0 sits at byte 0 of a row; a lane owns one byte alignment s = x & 3 and one search
 * row, and evaluates 8 positions x = 32g + s + 4k with three 16-byte loads, eleven funnel shifts and 32
 * VABSDIFF4 per block row. The first minimum in raster order (the reference's strict '<') is the minimum of
 * (sad << 14 | y << 7 | x). Metrics are integer sums, so the order of the atomics does not matter.
 */
#include <cuda_runtime.h>

#include "me_kernels.h"

#define DG_THREADS 256
#define DG_PW 36 /* words per staged window row: 128 + 15 bytes, 4 x odd -> 8 rows hit 8 distinct 16-byte bank groups */
#define DG_ROWS 143
#define DG_MAX_SA 128

__device__ __forceinline__ uint32_t dg_vsad4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__global__ void __launch_bounds__(DG_THREADS) k_dg_hme(const uint8_t *__restrict__ pool, MePicGeom g, uint32_t src_slot,
                                                       uint32_t ref_slot, int sa, MeGpuDgMetrics *acc) {
    __shared__ __align__(16) uint32_t win[DG_ROWS * DG_PW];
    __shared__ __align__(16) uint32_t srcw[16 * 4];
    __shared__ uint32_t s_best;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bx = (int)(blockIdx.x % (uint32_t)g.b64_cols), by = (int)(blockIdx.x / (uint32_t)g.b64_cols);
    const MePlaneGeom pl = g.plane[2];
    const int org_x = bx * 16, org_y = by * 16; /* ((int16_t)b64_origin) >> 2, pd_process.c:522-523 */

    /* pd_process.c:408-447: the left / top corrections move the origin only (the subtracted term is zero) */
    int sa_w = (sa + 7) & ~7, sa_h = sa;
    const int pad_w = pl.pad - 1, pad_h = pl.pad - 1;
    int ox = -(sa_w >> 1), oy = -(sa_h >> 1);
    if (org_x + ox < -pad_w)
        ox = -pad_w - org_x;
    if (org_x + ox > pl.width - 1)
        ox -= (org_x + ox) - (pl.width - 1);
    if (org_x + ox + sa_w > pl.width)
        sa_w = max(1, sa_w - ((org_x + ox + sa_w) - pl.width));
    sa_w = sa_w < 8 ? sa_w : sa_w & ~7;
    if (org_y + oy < -pad_h)
        oy = -pad_h - org_y;
    if (org_y + oy > pl.height - 1)
        oy -= (org_y + oy) - (pl.height - 1);
    if (org_y + oy + sa_h > pl.height)
        sa_h = max(1, sa_h - ((org_y + oy + sa_h) - pl.height));

    /* stage the 16x16 source block and the window (position (0,0) at byte 0 of row 0) */
    {
        const uint8_t *sp = pool + (size_t)src_slot * g.slot_bytes + pl.offset + (size_t)(pl.pad + org_y) * pl.stride + pl.pad + org_x;
        if (tid < 64)
            srcw[tid] = *(const uint32_t *)(sp + (size_t)(tid >> 2) * pl.stride + (tid & 3) * 4);
        if (tid == 0)
            s_best = 0xFFFFFFFFu;
        const uint8_t  *w00 = pool + (size_t)ref_slot * g.slot_bytes + pl.offset + (size_t)(pl.pad + org_y + oy) * pl.stride + pl.pad + org_x + ox;
        const int       a   = (int)((uintptr_t)w00 & 3);
        const uint32_t *g0  = (const uint32_t *)(w00 - a);
        const int       nw = (sa_w + 15 + 3) >> 2, nrows = sa_h + 15, sw = pl.stride >> 2;
        for (int i = tid; i < nrows * nw; i += DG_THREADS) {
            const int       r = i / nw, cw = i - r * nw;
            const uint32_t *p = g0 + (size_t)r * sw + cw;
            win[r * DG_PW + cw] = __funnelshift_r(p[0], p[1], 8 * a);
        }
    }
    __syncthreads();

    /* warp task = 32 positions x 8 search rows; lane = (byte alignment s, row) */
    const int ngx = (sa_w + 31) >> 5, nyb = (sa_h + 7) >> 3;
    const int s = lane & 3, ry = lane >> 2;
    uint32_t  best = 0xFFFFFFFFu;
    for (int t = warp; t < ngx * nyb; t += DG_THREADS / 32) {
        const int gx = t % ngx, yb = t / ngx;
        const int y = 8 * yb + ry, x0 = 32 * gx + s;
        uint32_t  sad[8];
#pragma unroll
        for (int k = 0; k < 8; k++) sad[k] = 0;
        const uint4 *rp = (const uint4 *)(win + y * DG_PW + 8 * gx);
        const uint4 *sp = (const uint4 *)srcw;
#pragma unroll 4
        for (int r = 0; r < 16; r++) {
            uint32_t w[12], v[11];
#pragma unroll
            for (int i = 0; i < 3; i++) {
                const uint4 q = rp[i];
                w[4 * i] = q.x; w[4 * i + 1] = q.y; w[4 * i + 2] = q.z; w[4 * i + 3] = q.w;
            }
#pragma unroll
            for (int i = 0; i < 11; i++) v[i] = __funnelshift_r(w[i], w[i + 1], 8 * s);
            const uint4 q = sp[r];
#pragma unroll
            for (int k = 0; k < 8; k++) {
                sad[k] = dg_vsad4(q.x, v[k], sad[k]);
                sad[k] = dg_vsad4(q.y, v[k + 1], sad[k]);
                sad[k] = dg_vsad4(q.z, v[k + 2], sad[k]);
                sad[k] = dg_vsad4(q.w, v[k + 3], sad[k]);
            }
            rp += DG_PW / 4;
        }
        if (y < sa_h) {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const int x = x0 + 4 * k;
                if (x < sa_w)
                    best = min(best, (sad[k] << 14) | (uint32_t)(y << 7 | x));
            }
        }
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (lane == 0)
        atomicMin(&s_best, best);
    __syncthreads();

    if (tid == 0) {
        /* pd_process.c:478-481 and :525-552 */
        const uint32_t key = s_best, sad = key >> 14;
        const int      col = ((int)(key & 127u) + ox) * 4, row = ((int)((key >> 7) & 127u) + oy) * 4;
        atomicAdd((unsigned long long *)&acc->tot_dist, (unsigned long long)sad);
        if (sad > 16 * 16 * 30)
            atomicAdd(&acc->tot_cplx, 1u);
        if (col != 0 || row != 0)
            atomicAdd(&acc->tot_active, 1u);
        int in = 0;
        if (by < g.b64_rows / 2)
            in += row > 0 ? -1 : row < 0 ? 1 : 0;
        else if (by > g.b64_rows / 2)
            in += row > 0 ? 1 : row < 0 ? -1 : 0;
        if (bx < g.b64_cols / 2)
            in += col > 0 ? -1 : col < 0 ? 1 : 0;
        else if (bx > g.b64_cols / 2)
            in += col > 0 ? 1 : col < 0 ? -1 : 0;
        if (in)
            atomicAdd(&acc->sum_in_vectors, in);
    }
}

static unsigned long long g_dg_launches = 0;
unsigned long long me_dg_launch_count(void) { return g_dg_launches; }

void me_launch_dg(const uint8_t *pool, const MePicGeom *g, uint32_t src_slot, uint32_t ref_slot, int sa, MeGpuDgMetrics *acc,
                  cudaStream_t st) {
    k_dg_hme<<<g->b64_count, DG_THREADS, 0, st>>>(pool, *g, src_slot, ref_slot, sa, acc);
    g_dg_launches++;
}

/* ------------------------------------------------------------------------------------------------
 * me_gpu_dg_detector_noalloc: the same kernel for callers that may not allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). No counterpart in the reference.
 *
 * Scratch (the two sixteenth planes at the pool's pitch + the metric accumulator) lives in ONE file-scope device array
 * of ME_GPU_DG_NOALLOC_MAX_BYTES that every GPU loading this library holds until exit. In host code `g_dg_scratch` is a
 * host-side stand-in for the symbol, not a device address: its device address is fetched once per device with
 * cudaGetSymbolAddress (that device current) and only that pointer is handed to the kernel.
 * All calls share the array and it keeps the last call's contents: the accumulator is cleared on every call, the planes
 * are overwritten in full. Calls that overlap on the GPU race; calls from one process are serialised here by a mutex.
 * ---------------------------------------------------------------------------------------------- */
#include <mutex>

static __device__ __align__(256) uint8_t g_dg_scratch[ME_GPU_DG_NOALLOC_MAX_BYTES];

#define DG_NOALLOC_ACC_BYTES 256 /* the accumulator's share at the head of the scratch (keeps the planes 256-byte aligned) */
#define DG_NOALLOC_MAX_DEVICES 64

static std::mutex  g_dg_noalloc_mu;
static uint8_t    *g_dg_scratch_ptr[DG_NOALLOC_MAX_DEVICES]; /* device address of g_dg_scratch, per device */
static const char *g_dg_noalloc_err = "";

static int dg_noalloc_fail(int code, const char *msg) {
    g_dg_noalloc_err = msg; /* string literals only: nothing is allocated on this path either */
    return code;
}

/* once per device, with that device current: the symbol's device address (and the module's load, which the first use of
 * any kernel of the library triggers). Call it outside a graph capture; the entry point calls it when needed. */
extern "C" int me_gpu_dg_detector_noalloc_setup(void) {
    std::lock_guard<std::mutex> lk(g_dg_noalloc_mu);
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0)
        return dg_noalloc_fail(ME_GPU_ERR_NO_DEVICE, "no CUDA device (this engine has no CPU fallback)");
    if (dev >= DG_NOALLOC_MAX_DEVICES)
        return dg_noalloc_fail(ME_GPU_ERR_BAD_ARG, "device ordinal beyond the table of scratch addresses");
    if (!g_dg_scratch_ptr[dev]) {
        void *p = nullptr;
        if (cudaGetSymbolAddress(&p, g_dg_scratch) != cudaSuccess || !p) {
            cudaGetLastError();
            return dg_noalloc_fail(ME_GPU_ERR_CUDA, "cudaGetSymbolAddress(g_dg_scratch) failed");
        }
        cudaFuncAttributes fa;
        if (cudaFuncGetAttributes(&fa, k_dg_hme) != cudaSuccess) {
            cudaGetLastError();
            return dg_noalloc_fail(ME_GPU_ERR_CUDA, "k_dg_hme is not loadable on this device (sm_100a required)");
        }
        g_dg_scratch_ptr[dev] = (uint8_t *)p;
    }
    return ME_GPU_OK;
}

extern "C" const char *me_gpu_dg_detector_noalloc_error(void) { return g_dg_noalloc_err; }

extern "C" int me_gpu_dg_detector_noalloc(const uint8_t *src_sixteenth, int src_stride, const uint8_t *ref_sixteenth, int ref_stride,
                                          int width, int height, uint8_t input_resolution, MeGpuDgMetrics *metrics) {
    if (!src_sixteenth || !ref_sixteenth || !metrics)
        return dg_noalloc_fail(ME_GPU_ERR_BAD_ARG, "null argument");
    if (width <= 0 || height <= 0 || (width & 7) || (height & 7))
        return dg_noalloc_fail(ME_GPU_ERR_BAD_ARG, "luma size must be a positive multiple of 8");
    MePicGeom g;
    me_make_geom(&g, width, height);
    const MePlaneGeom pl        = g.plane[2];
    const int         row_bytes = pl.width + 2 * pl.pad;
    const size_t      slot      = ((size_t)pl.stride * (size_t)pl.rows + 255) & ~(size_t)255;
    if (DG_NOALLOC_ACC_BYTES + 2 * slot > (size_t)ME_GPU_DG_NOALLOC_MAX_BYTES)
        return dg_noalloc_fail(ME_GPU_ERR_NO_MEMORY,
                               "picture too large for the static scratch of me_gpu_dg_detector_noalloc (ME_GPU_DG_NOALLOC_MAX_BYTES: "
                               "two sixteenth planes of up to 3840x2160 pictures)");
    if (src_stride < row_bytes || ref_stride < row_bytes)
        return dg_noalloc_fail(ME_GPU_ERR_BAD_ARG, "stride smaller than the padded sixteenth plane's row (width / 4 + 32)");
    int rc = me_gpu_dg_detector_noalloc_setup();
    if (rc != ME_GPU_OK)
        return rc;
    std::lock_guard<std::mutex> lk(g_dg_noalloc_mu);
    int dev = 0;
    cudaGetDevice(&dev);
    uint8_t        *base = g_dg_scratch_ptr[dev]; /* the DEVICE address; never g_dg_scratch itself */
    MeGpuDgMetrics *acc  = reinterpret_cast<MeGpuDgMetrics *>(base);
    uint8_t        *pool = base + DG_NOALLOC_ACC_BYTES;
    /* the kernel addresses planes as pool + slot * slot_bytes + plane[2].offset: slot 0 = source, slot 1 = reference */
    g.plane[2].offset = 0;
    g.slot_bytes      = slot;
    const int sa = input_resolution <= 1 ? 16 : input_resolution <= 2 ? 64 : 128; /* pd_process.c:497-498 */
    cudaStream_t st = 0; /* the default stream: no stream is created either */
    cudaError_t  e  = cudaMemsetAsync(acc, 0, DG_NOALLOC_ACC_BYTES, st); /* the array keeps the last call's sums */
    if (e == cudaSuccess)
        e = cudaMemcpy2DAsync(pool, pl.stride, src_sixteenth, src_stride, row_bytes, pl.rows, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess)
        e = cudaMemcpy2DAsync(pool + slot, pl.stride, ref_sixteenth, ref_stride, row_bytes, pl.rows, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        me_launch_dg(pool, &g, 0u, 1u, sa, acc, st);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(metrics, acc, sizeof(MeGpuDgMetrics), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return dg_noalloc_fail(ME_GPU_ERR_CUDA, "me_gpu_dg_detector_noalloc: a CUDA call failed");
    }
    return ME_GPU_OK;
}
