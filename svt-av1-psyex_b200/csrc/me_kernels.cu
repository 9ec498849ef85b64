/*
 * me_kernels.cu - hand-written sm_100a kernels of the open-loop ME engine.
 *
 *   k_downsample_pad   svt_aom_downsample_2d_c + svt_aom_generate_padding   (pic_analysis_process.c:130-160,
 *                      pic_operators.c:397-443): one pass writes a whole padded plane of the pyramid
 *   k_pad_full         edge replication of the full-resolution plane (svt_aom_pad_input_pictures)
 *   (me_hme.cu: k_hme, the per-b64 chain up to the small integer searches; me_dg.cu: k_dg_hme)
 *   k_fullsearch       open_loop_me_fullpel_search_sblock (motion_estimation.c:781-817) with the 85 square
 *                      PU SADs per position (motion_estimation.c:210-425): persistent CTAs pull 64x32-position
 *                      tiles, stage the reference window into four byte-shifted shared-memory copies, keep 8
 *                      positions per thread in registers, VABSDIFF4.ACC for the SADs, packed (sad, scan index)
 *                      keys reduced with redux.sync + shared/global atomicMin so that the lowest scan index
 *                      wins ties exactly as the reference's strict '<' does
 *   k_epilogue         me_prune_ref, candidate arrays, compute_distortion, GM detection per b64
 *
 * All arithmetic is integer; results are bit-exact with the reference C code (tests/test_gpu_parity.py).
 */
#include <cuda_runtime.h>
#include <stdint.h>

#include "me_ctrl.h"
#include "me_kernels.h"

/* ------------------------------------------------------------------------------------------------
 * small device helpers
 * ---------------------------------------------------------------------------------------------- */
__device__ __forceinline__ uint32_t vsad4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ const uint8_t *plane_ptr(const uint8_t *pool, const MePicGeom &g, uint32_t slot, int level) {
    return pool + (size_t)slot * g.slot_bytes + g.plane[level].offset;
}

/* ------------------------------------------------------------------------------------------------
 * pyramid
 * ---------------------------------------------------------------------------------------------- */

/* dst(px, py) over the whole padded plane = rounded 2x2 average of the source interior at the clamped
 * interior coordinate, i.e. downsample (step 2) followed by edge replication. 4 output bytes per thread. */
__global__ void __launch_bounds__(256) k_downsample_pad(uint8_t *pool, size_t slot_off, MePlaneGeom src, MePlaneGeom dst) {
    const int wx = blockIdx.x * blockDim.x + threadIdx.x; /* word index in the row */
    const int py = blockIdx.y;
    if (wx * 4 >= dst.stride)
        return;
    const uint8_t *s  = pool + slot_off + src.offset;
    uint8_t       *d  = pool + slot_off + dst.offset;
    int            iy = py - dst.pad;
    iy                = iy < 0 ? 0 : (iy > dst.height - 1 ? dst.height - 1 : iy);
    const uint8_t *r0 = s + (size_t)(src.pad + 2 * iy) * src.stride + src.pad;
    const uint8_t *r1 = r0 + src.stride;
    uint32_t       out = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int ix = wx * 4 + i - dst.pad;
        ix     = ix < 0 ? 0 : (ix > dst.width - 1 ? dst.width - 1 : ix);
        const uint32_t v = ((uint32_t)r0[2 * ix] + r0[2 * ix + 1] + r1[2 * ix] + r1[2 * ix + 1] + 2) >> 2;
        out |= v << (8 * i);
    }
    *(uint32_t *)(d + (size_t)py * dst.stride + wx * 4) = out;
}

/* rows of `row_bytes` bytes from a contiguous staging copy (pitch src_pitch) to their place in the pool (pitch dst_pitch);
 * all three are multiples of 4. A contiguous host->device copy followed by this runs at three times the rate of a
 * pitched cudaMemcpy2DAsync of 2 KB rows (tools/h2d_probe.cu: 48 vs 16 GB/s). */
__global__ void __launch_bounds__(256) k_repitch(const uint8_t *__restrict__ src, int src_pitch, uint8_t *__restrict__ dst, int dst_pitch,
                                                 int row_words) {
    const int wx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (wx < row_words)
        *(uint32_t *)(dst + (size_t)y * dst_pitch + 4 * wx) = *(const uint32_t *)(src + (size_t)y * src_pitch + 4 * wx);
}

/* replicate the interior of the full plane into its 68-pixel border (rows and columns) */
__global__ void __launch_bounds__(256) k_pad_full(uint8_t *pool, size_t slot_off, MePlaneGeom pl) {
    const int wx = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y;
    const int padded_w = pl.width + 2 * pl.pad;
    if (wx * 4 >= padded_w)
        return;
    const int px = wx * 4;
    const bool inside_y = py >= pl.pad && py < pl.pad + pl.height;
    const bool inside_x = px >= pl.pad && px < pl.pad + pl.width;
    if (inside_x && inside_y)
        return; /* pad and width are multiples of 4: a word is never half interior */
    uint8_t *base = pool + slot_off + pl.offset;
    int      iy   = py - pl.pad;
    iy            = iy < 0 ? 0 : (iy > pl.height - 1 ? pl.height - 1 : iy);
    const uint8_t *row = base + (size_t)(pl.pad + iy) * pl.stride + pl.pad;
    uint32_t       out = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int ix = px + i - pl.pad;
        ix     = ix < 0 ? 0 : (ix > pl.width - 1 ? pl.width - 1 : ix);
        out |= (uint32_t)row[ix] << (8 * i);
    }
    *(uint32_t *)(base + (size_t)py * pl.stride + px) = out;
}

/* ------------------------------------------------------------------------------------------------
 * k_fullsearch: integer full search, all 85 square PUs per position
 * ---------------------------------------------------------------------------------------------- */
#define FS_PW 28 /* words per window row and copy (32 + 63 + 3 bytes): 4 x odd, so that 8 rows hit 8 distinct 16-byte bank groups */
#define FS_ROWS (FS_TH + 64) /* FS_TH + 63 window rows (+1) */
#define FS_KEY_SHIFT 10 /* 32 x 32 tile positions */
#define FS_BIAS (1u << 17) /* added to the 8x8 SADs of positions outside the tile: they can never win */
#define FS_SBEST 96 /* words per warp: best (sad << 10 | tile position) of the 85 PUs over the warp's 8 search rows */

/* a tile visits every PU once per warp and sbest has a row per warp: no atomic is needed */
#ifndef FS_STORE_MODE
#define FS_STORE_MODE 2
#endif
#if FS_STORE_MODE == 0
#define FS_STORE(i, m) do { if (lane == 0) atomicMin(&sbest[i], m); } while (0)
#elif FS_STORE_MODE == 1
#define FS_STORE(i, m) do { if (lane == 0) sbest[i] = (m); } while (0)
#else
#define FS_STORE(i, m) do { sbest[i] = (m); } while (0)
#endif
/* a + b on the FMA pipe (IMAD with a multiplier the compiler cannot see is 1): the ALU pipe is the one VABSDIFF4 needs */
__device__ __forceinline__ uint32_t fs_add_fma(uint32_t a, uint32_t b, uint32_t one) {
#ifdef FS_ADD_ON_ALU
    (void)one;
    return a + b;
#else
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
    return d;
#endif
}
template <bool SUB, int KX>
__device__ __forceinline__ void fs_compute(const uint32_t *__restrict__ wc, const uint32_t *__restrict__ srcw,
                                           uint32_t *sbest, int xw0, int ly, int s, int wvalid, int hvalid, int lane, uint32_t one) {
    constexpr int RSTEP = SUB ? 2 : 1; /* window rows per summed block row */
    constexpr int NR    = SUB ? 4 : 8; /* summed rows per 8x8 block */
    uint32_t      bias[KX], idx[KX];
    uint32_t      vmask = 0;
#pragma unroll
    for (int k = 0; k < KX; k++) {
        const int  lx = xw0 + s + 4 * k;
        const bool v  = lx < wvalid && ly < hvalid;
        bias[k]       = v ? 0u : FS_BIAS;
        idx[k]        = (uint32_t)(ly * FS_TW + lx);
        vmask |= (v ? 1u : 0u) << k;
    }
    const int j0 = xw0 >> 2;
    uint32_t  s64[KX];
#pragma unroll
    for (int k = 0; k < KX; k++) s64[k] = 0;
#pragma unroll 1
    for (int q = 0; q < 4; q++) {
        const int qx = q & 1, qy = q >> 1;
        uint32_t  s32[KX];
#pragma unroll
        for (int k = 0; k < KX; k++) s32[k] = 0;
#pragma unroll 1
        for (int h = 0; h < 2; h++) {
            uint32_t s16[2][KX];
#pragma unroll
            for (int k = 0; k < KX; k++) s16[0][k] = s16[1][k] = 0;
#pragma unroll
            for (int br = 0; br < 2; br++) {
                uint32_t acc[4][KX];
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int k = 0; k < KX; k++) acc[i][k] = bias[k];
                const int brow = 32 * qy + 16 * h + 8 * br; /* first pixel row of this 8x8 block row */
#pragma unroll
                for (int r = 0; r < NR; r++) {
                    const uint4 *rp = (const uint4 *)(wc + (ly + brow + r * RSTEP) * FS_PW + j0 + 8 * qx);
                    const uint4 *sp = (const uint4 *)(srcw + (brow / RSTEP + r) * 16 + 8 * qx);
                    uint32_t     rw[16], sw[8];
#pragma unroll
                    for (int i = 0; i < (7 + KX + 3) / 4; i++) {
                        const uint4 v = rp[i];
                        rw[4 * i] = v.x; rw[4 * i + 1] = v.y; rw[4 * i + 2] = v.z; rw[4 * i + 3] = v.w;
                    }
#pragma unroll
                    for (int i = 0; i < 2; i++) {
                        const uint4 v = sp[i];
                        sw[4 * i] = v.x; sw[4 * i + 1] = v.y; sw[4 * i + 2] = v.z; sw[4 * i + 3] = v.w;
                    }
#pragma unroll
                    for (int cw = 0; cw < 8; cw++)
#pragma unroll
                        for (int k = 0; k < KX; k++) acc[cw >> 1][k] = vsad4(sw[cw], rw[cw + k], acc[cw >> 1][k]);
                }
                /* four 8x8 PUs done for 8 positions: z-order index 21 + 16q + 8h + 4(i>>1) + 2br + (i&1) */
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    uint32_t m = 0xFFFFFFFFu;
#pragma unroll
                    for (int k = 0; k < KX; k++) m = min(m, (acc[i][k] << FS_KEY_SHIFT) + idx[k]);
                    m = __reduce_min_sync(0xffffffffu, m);
                    FS_STORE(21 + 16 * q + 8 * h + 4 * (i >> 1) + 2 * br + (i & 1), m);
#pragma unroll
                    for (int k = 0; k < KX; k++) s16[i >> 1][k] = fs_add_fma(acc[i][k], s16[i >> 1][k], one);
                }
            }
            /* two 16x16 PUs: 5 + 4q + 2h + {0,1} */
#pragma unroll
            for (int i = 0; i < 2; i++) {
                uint32_t m = 0xFFFFFFFFu;
#pragma unroll
                for (int k = 0; k < KX; k++) m = min(m, (s16[i][k] << FS_KEY_SHIFT) + idx[k]);
                m = __reduce_min_sync(0xffffffffu, m);
                FS_STORE(5 + 4 * q + 2 * h + i, m);
#pragma unroll
                for (int k = 0; k < KX; k++) s32[k] = fs_add_fma(s16[i][k], s32[k], one);
            }
        }
        /* 32x32 PU 1 + q: positions outside the tile are masked explicitly (their bias overflows the key) */
        {
            uint32_t m = 0xFFFFFFFFu;
#pragma unroll
            for (int k = 0; k < KX; k++) {
                const uint32_t key = (s32[k] << FS_KEY_SHIFT) + idx[k];
                m                  = min(m, (vmask >> k) & 1u ? key : 0xFFFFFFFFu);
            }
            m = __reduce_min_sync(0xffffffffu, m);
            FS_STORE(1 + q, m);
#pragma unroll
            for (int k = 0; k < KX; k++) s64[k] = fs_add_fma(s32[k], s64[k], one);
        }
    }
    {
        uint32_t m = 0xFFFFFFFFu;
#pragma unroll
        for (int k = 0; k < KX; k++) {
            const uint32_t key = (s64[k] << FS_KEY_SHIFT) + idx[k];
            m                  = min(m, (vmask >> k) & 1u ? key : 0xFFFFFFFFu);
        }
        m = __reduce_min_sync(0xffffffffu, m);
        FS_STORE(0, m);
    }
}

/* ---- TMA / mbarrier / cp.async helpers of k_fullsearch ---- */
__device__ __forceinline__ uint32_t fs_smem(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fs_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fs_mbar_expect(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void fs_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done)
                     : "r"(bar), "r"(parity), "r"(1000u)
                     : "memory");
}
__device__ __forceinline__ void fs_tma_box(uint32_t dst, const void *map, int x, int y, int z, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void fs_cp_async4(uint32_t dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}

#define FS_RAW_PITCH 128 /* bytes per row of the TMA box: 15 (alignment) + 31 (positions) + 64 (block) <= 128 */
#define FS_RAW_ROWS 96 /* FS_TH + 63 window rows, in one box */
#define FS_RAW_BYTES (FS_RAW_PITCH * FS_RAW_ROWS)
/* shared memory: raw[2] (TMA double buffer) | win[4][FS_ROWS][FS_PW] (the four byte-shifted copies) | srcw[64][16] | sbest[4][FS_SBEST] */
#define FS_SMEM_BYTES (2 * FS_RAW_BYTES + (4 * FS_ROWS * FS_PW + 64 * 16 + 4 * FS_SBEST) * 4)

/* One tile per iteration, software-pipelined over a double buffer: while the CTA searches tile i, the window of tile i + 1
 * is in flight (ONE TMA box of 128 x 96 bytes on the (x, row, slot) tensor map of the pool, 16-byte-aligned column, completion
 * on an mbarrier) and the work-list index of tile i + 2 is being fetched. TMA cannot shift bytes, so the four byte-shifted
 * copies the SAD loop reads with LDS.128 are made shared -> shared from the raw window (3 LDS + 4 funnel shifts + 4 STS per
 * word, a uniform switch on the alignment instead of per-word selects and divisions); the source block comes in by
 * cp.async while that pass runs. */
__global__ void __launch_bounds__(FS_THREADS, FS_CTAS_PER_SM)
    k_fullsearch(MeJob *jobs, FsTileList *list, const uint8_t *pool, const __grid_constant__ HmeMaps maps) {
    extern __shared__ __align__(1024) uint8_t fs_dyn[];
    uint8_t  *raw   = fs_dyn; /* [2][FS_RAW_ROWS][FS_RAW_PITCH] */
    uint32_t *win   = reinterpret_cast<uint32_t *>(fs_dyn + 2 * FS_RAW_BYTES); /* [4][FS_ROWS][FS_PW] */
    uint32_t *srcw  = win + 4 * FS_ROWS * FS_PW; /* [64][16] (32 rows used for sub-sampled SAD) */
    uint32_t *sbest = srcw + 64 * 16; /* [4 warps][FS_SBEST] packed (sad << 10 | tile position): a tile visits every PU once per warp, so plain stores */
    __shared__ __align__(8) unsigned long long s_bar[2];
    __shared__ FsTile   s_desc[2];
    __shared__ uint32_t s_idx[2]; /* work-list index of the tile in s_desc[b]; >= total: none */
    const int      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t total = list->count > list->cap ? 0u : list->count;
    const uint32_t one = list->cap != 0u ? 1u : 0u; /* 1, but only at run time: see fs_add_fma */

    /* thread 0 runs the tile pipeline: index of tile i + 2 (atomic), descriptor of tile i + 1, TMA of tile i + 1 */
    uint32_t idx_next = 0xFFFFFFFFu, idx_next2 = 0xFFFFFFFFu;
    FsTile   d_next;
    if (tid == 0) {
        fs_mbar_init(fs_smem(&s_bar[0]), 1);
        fs_mbar_init(fs_smem(&s_bar[1]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        const uint32_t i0 = atomicAdd(&list->next, 1u);
        s_idx[0] = i0;
        s_idx[1] = 0xFFFFFFFFu;
        if (i0 < total) {
            const FsTile t0 = list->tiles[i0];
            s_desc[0]       = t0;
            const MeJob *j0 = &jobs[t0.job];
            const MePlaneGeom &pl = j0->g.plane[0];
            const int b = (int)(t0.task / ME_NREF);
            const int gx = (b % j0->g.b64_cols) * 64 + t0.x0 + (int)t0.tx, gy = (b / j0->g.b64_cols) * 64 + t0.y0 + (int)t0.ty;
            fs_mbar_expect(fs_smem(&s_bar[0]), FS_RAW_BYTES);
            fs_tma_box(fs_smem(raw), maps.m[FS_MAP], (pl.pad + gx) & ~15, pl.pad + gy, t0.ref_slot, fs_smem(&s_bar[0]));
            idx_next = atomicAdd(&list->next, 1u);
            if (idx_next < total)
                d_next = list->tiles[idx_next];
        }
    }
    for (uint32_t it = 0;; it++) {
        const int buf = (int)(it & 1);
        __syncthreads(); /* the previous tile is done with win / srcw / sbest and with raw[buf ^ 1]; s_idx / s_desc[buf] are visible */
        const uint32_t ti = s_idx[buf];
        if (ti >= total)
            break;
        const FsTile tile = s_desc[buf];
        if (tid == 0) { /* tile it + 1: its window goes into the other buffer while this one is searched */
            s_idx[buf ^ 1] = idx_next;
            if (idx_next < total) {
                s_desc[buf ^ 1] = d_next;
                const MeJob *jn = &jobs[d_next.job];
                const MePlaneGeom &pl = jn->g.plane[0];
                const int b = (int)(d_next.task / ME_NREF);
                const int gx = (b % jn->g.b64_cols) * 64 + d_next.x0 + (int)d_next.tx, gy = (b / jn->g.b64_cols) * 64 + d_next.y0 + (int)d_next.ty;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); /* the shift pass read raw[buf ^ 1] through the generic proxy */
                fs_mbar_expect(fs_smem(&s_bar[buf ^ 1]), FS_RAW_BYTES);
                fs_tma_box(fs_smem(raw + (buf ^ 1) * FS_RAW_BYTES), maps.m[FS_MAP], (pl.pad + gx) & ~15, pl.pad + gy, d_next.ref_slot,
                           fs_smem(&s_bar[buf ^ 1]));
                idx_next2 = atomicAdd(&list->next, 1u); /* consumed at the end of this iteration: the latency hides behind the search */
            } else
                idx_next2 = 0xFFFFFFFFu;
        }
        MeJob        *j    = &jobs[tile.job];
        FsTask        f;
        f.valid = 1; f.probe = tile.probe; f.ref_slot = tile.ref_slot;
        f.x0 = tile.x0; f.y0 = tile.y0; f.w = tile.w; f.h = tile.h;
        const uint32_t b   = tile.task / ME_NREF;
        const bool    sub  = j->p.me_search_method == ME_GPU_SUB_SAD_SEARCH;
        const MePlaneGeom pl = j->g.plane[0];
        const int org_x = (int)(b % (uint32_t)j->g.b64_cols) * 64, org_y = (int)(b / (uint32_t)j->g.b64_cols) * 64;
        const int kx = tile.kx;
        const int wvalid = min(4 * kx, f.w - (int)tile.tx), hvalid = min(FS_TH, f.h - (int)tile.ty);

        /* source block rows (every other row for sub-sampled SAD): asynchronous, lands while the shift pass runs */
        {
            const uint8_t *sp = plane_ptr(pool, j->g, j->src_slot, 0) + (size_t)(pl.pad + org_y) * pl.stride + pl.pad + org_x;
            const int      nrows = sub ? 32 : 64, rstep = sub ? 2 : 1;
            for (int i = tid; i < nrows * 16; i += FS_THREADS)
                fs_cp_async4(fs_smem(srcw + i), sp + (size_t)((i >> 4) * rstep) * pl.stride + (i & 15) * 4);
            asm volatile("cp.async.commit_group;" ::: "memory");
            for (int i = tid; i < 4 * FS_SBEST; i += FS_THREADS) sbest[i] = 0xFFFFFFFFu; /* warps without positions leave it so */
        }
        /* the four byte-shifted copies of the window: copy s, word cw of row r = raw bytes a + s + 4 cw .. of that row */
        {
            const int gx = org_x + f.x0 + (int)tile.tx;
            const int a  = (pl.pad + gx) & 15; /* byte of search position 0 in a raw row (the box starts on a 16-byte column) */
            const int nrows = hvalid + (sub ? 62 : 63) + 1;
            fs_mbar_wait(fs_smem(&s_bar[buf]), (it >> 1) & 1);
            const uint32_t *rw0 = reinterpret_cast<const uint32_t *>(raw + buf * FS_RAW_BYTES) + (a >> 2);
            const int a3 = a & 3;
            if (lane < FS_PW) {
#define FS_SHIFT_ROWS(E0, E1, E2, E3)                                                          \
    for (int r = warp; r < nrows; r += FS_THREADS / 32) {                                      \
        const uint32_t *q  = rw0 + r * (FS_RAW_PITCH / 4) + lane;                              \
        const uint32_t  w0 = q[0], w1 = q[1], w2 = q[2];                                       \
        (void)w2;                                                                              \
        uint32_t *o = win + r * FS_PW + lane;                                                  \
        o[0]                     = E0;                                                         \
        o[FS_ROWS * FS_PW]       = E1;                                                         \
        o[2 * FS_ROWS * FS_PW]   = E2;                                                         \
        o[3 * FS_ROWS * FS_PW]   = E3;                                                         \
    }
                switch (a3) { /* uniform over the CTA */
                case 0: FS_SHIFT_ROWS(w0, __funnelshift_r(w0, w1, 8), __funnelshift_r(w0, w1, 16), __funnelshift_r(w0, w1, 24)) break;
                case 1: FS_SHIFT_ROWS(__funnelshift_r(w0, w1, 8), __funnelshift_r(w0, w1, 16), __funnelshift_r(w0, w1, 24), w1) break;
                case 2: FS_SHIFT_ROWS(__funnelshift_r(w0, w1, 16), __funnelshift_r(w0, w1, 24), w1, __funnelshift_r(w1, w2, 8)) break;
                default: FS_SHIFT_ROWS(__funnelshift_r(w0, w1, 24), w1, __funnelshift_r(w1, w2, 8), __funnelshift_r(w1, w2, 16)) break;
                }
#undef FS_SHIFT_ROWS
            }
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        /* compute: warp -> 32 x 8 positions, lane -> (byte shift s, row), KX positions x = s + 4k per thread */
        {
            /* the warp -> quarter-tile assignment rotates with the tile number: partly filled tiles leave different
               SM sub-partitions idle. (Splitting the columns of a one-row-group tile over the four warps - KX 2 each - was
               measured: +1.5 points at 1080p, -0.9 at 4K: the other CTAs of the SM fill the ALU pipe anyway.) */
            const int wq  = (warp + (int)ti) & 3;
            const int xw0 = 0, yw0 = wq * 8;
            if (xw0 < wvalid && yw0 < hvalid) {
                const int       s = lane >> 3, ly = yw0 + (lane & 7);
                const uint32_t *wc = win + s * FS_ROWS * FS_PW;
#define FS_CALL(SUB, K) fs_compute<SUB, K>(wc, srcw, sbest + warp * FS_SBEST, xw0, ly, s, wvalid, hvalid, lane, one)
                if (sub) {
                    if (kx == 8) FS_CALL(true, 8); else if (kx == 7) FS_CALL(true, 7); else if (kx == 6) FS_CALL(true, 6); else FS_CALL(true, 5);
                } else {
                    if (kx == 8) FS_CALL(false, 8); else if (kx == 7) FS_CALL(false, 7); else if (kx == 6) FS_CALL(false, 6); else FS_CALL(false, 5);
                }
#undef FS_CALL
            }
        }
        if (tid == 0) { /* the descriptor of tile it + 2, for the top of the next iteration */
            idx_next = idx_next2;
            if (idx_next < total)
                d_next = list->tiles[idx_next];
        }
        __syncthreads();
        /* merge the tile's best keys into the task's 64-bit (sad, scan order) keys */
        if (tid < ME_NPU) {
            const uint32_t key = min(min(sbest[tid], sbest[FS_SBEST + tid]), min(sbest[2 * FS_SBEST + tid], sbest[3 * FS_SBEST + tid]));
            if (key != 0xFFFFFFFFu) {
                const uint32_t rawk = key >> FS_KEY_SHIFT, pos = key & ((1u << FS_KEY_SHIFT) - 1u);
                const uint32_t lx = pos & (FS_TW - 1), ly = pos / FS_TW;
                const uint32_t sad = sub ? rawk << 1 : rawk;
                const uint32_t ord = f.probe ? 0u : 1u + (tile.ty + ly) * (uint32_t)f.w + tile.tx + lx;
                atomicMin(&j->keys[(size_t)tile.task * ME_NPU + tid], ((unsigned long long)sad << 32) | ord);
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------------
 * k_epilogue: one CTA per b64
 * ---------------------------------------------------------------------------------------------- */
#define EP_THREADS 96
__global__ void __launch_bounds__(EP_THREADS) k_epilogue(MeJob *jobs) {
    __shared__ uint32_t best_sad[ME_NREF][ME_NPU], best_mv[ME_NREF][ME_NPU], me_distortion[ME_NPU];
    __shared__ B64State st; /* the block's state (HBM, written by k_hme): every later access is a shared-memory one */
    MeJob         *j = &jobs[blockIdx.y];
    const uint32_t b = blockIdx.x;
    const int      tid = threadIdx.x;
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(&j->state[b]);
        uint32_t       *dst = reinterpret_cast<uint32_t *>(&st);
        for (int i = tid; i < (int)(sizeof(B64State) / 4); i += EP_THREADS) dst[i] = src[i];
    }
    __syncthreads();
    EpiCtx         e;
    e.j    = j;
    e.p    = &j->p;
    e.s    = &st;
    e.keys = &j->keys[(size_t)b * ME_NREF * ME_NPU];
    e.b    = b;
    const int org_x = (int)(b % (uint32_t)j->g.b64_cols) * 64, org_y = (int)(b / (uint32_t)j->g.b64_cols) * 64;
    e.b64_w         = (uint32_t)min(64, j->g.width - org_x);
    e.b64_h         = (uint32_t)min(64, j->g.height - org_y);
    e.best_sad      = best_sad;
    e.best_mv       = best_mv;
    e.me_distortion = me_distortion;
    for (int i = tid; i < ME_NREF * ME_NPU; i += EP_THREADS) me_epi_decode(&e, i / ME_NPU, i % ME_NPU);
    __syncthreads();
    if (j->p.me_type == ME_GPU_ME_MCTF) { /* no pruning, candidates or statistics (ME:3103,3126): results of slot (0,0) */
        if (tid < ME_NPU)
            me_epi_tf_pu(&e, tid);
        if (tid == 0) {
            me_epi_tf_block(&e);
            me_epi_taps(&e);
            atomicAdd(&j->ops_total[0], (unsigned long long)e.s->ops);
            atomicAdd(&j->ops_total[1], (unsigned long long)e.s->ops_fs);
            atomicAdd(&j->ops_total[4], (unsigned long long)e.s->ops_fs_inline);
        }
    } else {
        const int lane = tid & 31, warp = tid >> 5;
        if (j->p.enable_hme_flag && j->p.enable_me_hme_ref_pruning) {
            /* me_prune_ref: the 64 best 8x8 SADs of a reference summed by a warp (two per lane), then the test on thread 0 */
            for (int k = warp; k < ME_NREF; k += EP_THREADS / 32) {
                const int li = k / ME_NR, ri = k % ME_NR;
                if (li < (int)j->p.num_of_list_to_search && ri < (int)j->p.num_of_ref_pic_to_search[li]) {
                    const bool     on = st.do_ref[li][ri] != 0;
                    const uint32_t v  = __reduce_add_sync(0xffffffffu, on ? best_sad[k][21 + lane] + best_sad[k][53 + lane] : 0u);
                    if (lane == 0)
                        st.hme_sad[li][ri] = on ? v : ME_MAX_SAD_VALUE * 64u;
                }
            }
            __syncthreads();
            if (tid == 0)
                me_epi_prune_decide(&e);
        }
        __syncthreads();
        if (tid < j->p.max_number_of_pus_per_sb)
            me_epi_candidates(&e, tid);
        __syncthreads();
        /* the three serial tails are independent of each other: one warp each */
        if (tid == 0)
            me_epi_distortion(&e);
        else if (tid == 32)
            me_epi_gm(&e);
        else if (tid == 64) {
            me_epi_taps(&e);
            atomicAdd(&j->ops_total[0], (unsigned long long)e.s->ops);
            atomicAdd(&j->ops_total[1], (unsigned long long)e.s->ops_fs);
            atomicAdd(&j->ops_total[4], (unsigned long long)e.s->ops_fs_inline);
        }
    }
    if (j->dbg_best_sad)
        for (int i = tid; i < ME_NREF * ME_NPU; i += EP_THREADS)
            j->dbg_best_sad[(size_t)b * ME_NREF * ME_NPU + i] = best_sad[i / ME_NPU][i % ME_NPU];
    if (j->dbg_best_mv)
        for (int i = tid; i < ME_NREF * ME_NPU; i += EP_THREADS)
            j->dbg_best_mv[(size_t)b * ME_NREF * ME_NPU + i] = best_mv[i / ME_NPU][i % ME_NPU];
}

/* ------------------------------------------------------------------------------------------------
 * integer-SAD issue-rate probe (roofline denominator): independent VABSDIFF4.ACC chains
 * ---------------------------------------------------------------------------------------------- */
__global__ void __launch_bounds__(256) k_sad_peak(uint32_t *out, int iters, uint32_t seed) {
    uint32_t acc[16], a[4], b[4];
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = i;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        a[i] = seed * (threadIdx.x + i);
        b[i] = seed ^ (threadIdx.x * 77 + i);
    }
    for (int it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int i = 0; i < 16; i++) {
                uint32_t d;
                asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a[i & 3]), "r"(b[r]), "r"(acc[i]));
                acc[i] = d;
            }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) s += acc[i];
    if (s == 0x12345678u)
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

/* ------------------------------------------------------------------------------------------------
 * launchers
 * ---------------------------------------------------------------------------------------------- */
static unsigned long long g_launches = 0;
unsigned long long me_launch_count(void) { return g_launches; }

cudaError_t me_kernels_setup(void) {
    return cudaFuncSetAttribute(k_fullsearch, cudaFuncAttributeMaxDynamicSharedMemorySize, FS_SMEM_BYTES);
}

void me_launch_pad_full(uint8_t *pool, size_t slot_off, const MePlaneGeom *pl, cudaStream_t st) {
    dim3 grid((pl->width + 2 * pl->pad + 1023) / 1024, pl->rows);
    k_pad_full<<<grid, 256, 0, st>>>(pool, slot_off, *pl);
    g_launches++;
}

void me_launch_downsample(uint8_t *pool, size_t slot_off, const MePlaneGeom *src, const MePlaneGeom *dst, cudaStream_t st) {
    dim3 grid((dst->stride / 4 + 255) / 256, dst->rows);
    k_downsample_pad<<<grid, 256, 0, st>>>(pool, slot_off, *src, *dst);
    g_launches++;
}

void me_launch_fullsearch(MeJob *jobs, FsTileList *list, const uint8_t *pool, const HmeMaps *maps, int sms, cudaStream_t st) {
    k_fullsearch<<<sms * FS_CTAS_PER_SM, FS_THREADS, FS_SMEM_BYTES, st>>>(jobs, list, pool, *maps);
    g_launches++;
}

void me_launch_repitch(const uint8_t *src, int src_pitch, uint8_t *dst, int dst_pitch, int row_bytes, int rows, cudaStream_t st) {
    dim3 grid((row_bytes / 4 + 255) / 256, rows);
    k_repitch<<<grid, 256, 0, st>>>(src, src_pitch, dst, dst_pitch, row_bytes / 4);
    g_launches++;
}

void me_launch_epilogue(MeJob *jobs, int njobs, int max_b64, cudaStream_t st) {
    dim3 grid(max_b64, njobs);
    k_epilogue<<<grid, EP_THREADS, 0, st>>>(jobs);
    g_launches++;
}

/* returns VABSDIFF4 lane-ops per second on the current device */
double me_measure_sad_peak(cudaStream_t st) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint32_t *out = nullptr;
    if (cudaMalloc(&out, (size_t)sms * 8 * 256 * 4) != cudaSuccess)
        return 0.0;
    const int iters = 8192;
    k_sad_peak<<<sms * 8, 256, 0, st>>>(out, 256, 1234567u);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0, st);
        k_sad_peak<<<sms * 8, 256, 0, st>>>(out, iters, 1234567u);
        cudaEventRecord(e1, st);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double rate = 64.0 * iters * (double)sms * 8 * 256 / (ms * 1e-3);
        if (rate > best)
            best = rate;
    }
    g_launches += 4;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return best;
}
