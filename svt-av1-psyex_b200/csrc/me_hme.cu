/*
 * me_hme.cu - k_hme: the whole per-b64 chain of open-loop ME up to (and including) the small integer
 * searches, in ONE launch per picture.
 *
 * One CTA (4 warps) per 64x64 block. The block's MeContext state (B64State), its search tasks and their
 * results live in shared memory for the whole chain; nothing goes through HBM between stages and there is
 * no launch boundary between them (b64s are independent of each other: motion_estimation.c keeps no
 * cross-b64 state):
 *
 *   for every stage of me_ctrl.h (init/zz, pre-HME, HME L0, L1, L2, centre, probe, main):
 *       thread 0   runs the integer control stage on the shared-memory state -> up to 32 tasks
 *       warp 0     cuts the tasks into work items (bands of search rows) of similar cost
 *       4 warps    pull items from a shared counter, execute them on their own 8 KB staging buffer and
 *                  merge (sad, scan index) keys with a shared-memory atomicMin per task
 *
 * Item executors (all SADs are VABSDIFF4.U8.ACC on 32-bit words; unaligned reference words are built with
 * funnel shifts from the aligned shared-memory words):
 *   lane = search position, whole block per lane: the 16x8 blocks of pre-HME / HME-L0
 *   lane = (block row, 16-byte chunk), positions in sequence, redux.sync.add per position: big blocks with
 *          few positions (HME-L1/L2, zero-MV and centre SADs)
 *   fullsearch_item: lane = two 8x8 blocks (z-order): the 85 square-PU SADs of a position by shuffle
 *          butterflies, 7 running (sad, position) bests per lane in scan order (8x8-variance probe and search
 *          areas of up to FS_WARP_MAX positions)
 * Larger search areas are cut into 64x32-position tiles and appended to the work list of k_fullsearch.
 */
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "me_ctrl.h"
#include "me_kernels.h"

#define HM_WARPS 4
#define HM_THREADS (HM_WARPS * 32) /* search threads */
#define HM_ALL_THREADS (HM_THREADS + 32) /* + the warp that requests the windows (TMA) */
#define HM_BUF_WORDS 2048 /* per-warp staging buffer: source block (<= 512 words) + reference window (15 rows x
                             131 words of the widest pre-HME strip fit; taller windows are banded) */
#ifndef FS_WARP_MAX
#define FS_WARP_MAX 256
#endif
/* FS_WARP_MAX: search areas up to this many positions (and <= FS_WARP_MAXW wide) stay in k_hme */
#ifndef FS_WARP_MAXW
#define FS_WARP_MAXW 48
#endif
#define HM_ITEM_POS_A 384 /* search positions per work item, lane = position */
#define HM_ITEM_POS_B 32 /* ... lane = block row */
#define HM_ITEM_POS_FS 16 /* ... full-search positions */

#define HM_RAW_BYTES 16384 /* largest window of a job: half of the ring */

/* source-block pitches in shared memory (words): padded so that lanes on different block rows hit different banks */
#define HM_SRC0_PITCH 24
#define HM_SRC1_PITCH 12
#define HM_SRC2_PITCH 4

struct HmPlan { /* how a task is cut into jobs (written by warp 0 at the start of a stage) */
    int16_t pwb, neb; /* positions per x band, evaluated rows per y band */
    uint8_t nxb, nyb, map, kx;
    uint8_t g_log2, new_path, team /* warps that search a job: 1 or 2 for jobs of few units, else 4 */, pad_;
};

struct HmJobD { /* one window job, 32 bytes */
    int16_t  px0, pw, e0, ne; /* band of search positions: columns [px0, px0 + pw), evaluated rows [e0, e0 + ne) */
    int16_t  xt, y0; /* TMA column (multiple of 16) and row of the window */
    uint16_t off128; /* window offset in the ring, in 128-byte units */
    uint16_t nrows; /* window rows (whole boxes) */
    uint16_t total; /* lanes' worth of work of the SAD loop: units x lanes per unit */
    uint8_t  task, map, aoff /* byte offset of position px0 in a window row */, ncls, nchunk;
    int8_t   dep; /* last earlier job of the stage whose window overlaps this one (-1: none) */
    uint32_t m_chunk, m_cls;
};
#define HM_MAXJOBS 32 /* window jobs per stage; tasks beyond go to the per-warp path */
#define HM_NBAR 8 /* window jobs in flight (ring of job descriptors and mbarriers) */
#define HM_RING_BYTES (2 * HM_RAW_BYTES)

#ifdef ME_HME_TRACE
__device__ unsigned g_trace_cnt;
#endif
/* planner thresholds (swept in round 1, +-1 % around these): search areas at least this wide / with at least this many
   8-position units take hm_job_search8; jobs of up to HM_TEAM1_UNITS units are searched by one warp each; 8-wide areas
   of at least HM_TALL_MIN_ROWS rows take hm_job_search_tall */
#define HM_K8_MIN_PW 32
#define HM_K8_MIN_UNITS 64
#define HM_TEAM1_UNITS 16
#define HM_LANES_TARGET 96
#define HM_TALL_MIN_ROWS 32

struct HmeShared {
    B64State           s;
    SearchTask         t[ME_SLOTS];
    SearchResult       r[ME_SLOTS];
    FsTask             f[ME_NREF];
    unsigned long long tkey[ME_SLOTS]; /* best (sad << 32 | scan index) of each search task */
    uint16_t           item_first[ME_SLOTS + 1]; /* prefix sums: items of task i are [first[i], first[i+1]) */
    uint16_t           item_rows[ME_SLOTS]; /* search rows per item */
    int                next_item;
    int                team; /* warps per window job in this round (4 / team jobs are searched at a time) */
    int                next_lo; /* first task of the next round of window jobs */
    HmPlan             plan[ME_SLOTS];
    uint16_t           job_first[ME_SLOTS + 1]; /* prefix sums of the window jobs */
    MeGpuParams        p; /* the picture's controls: the control stages read them hundreds of times, one dependent load each */
    alignas(8) unsigned long long mbar[HM_NBAR]; /* TMA completion of the windows in flight */
    HmJobD             jobs[HM_MAXJOBS];
    alignas(16) uint32_t src0[64 * HM_SRC0_PITCH], src1[32 * HM_SRC1_PITCH], src2[16 * HM_SRC2_PITCH]; /* the b64's source block on the three planes (padded pitches) */
    /* per-warp staging buffers of the item path; the two window buffers of the job path alias them
       (2 x HM_RAW_BYTES + 128 bytes of slack for masked over-reads) */
    alignas(128) uint32_t buf[HM_WARPS][HM_BUF_WORDS + 8];
};
static_assert(sizeof(uint32_t) * HM_WARPS * (HM_BUF_WORDS + 8) >= 2 * HM_RAW_BYTES + 128, "window buffers alias the warp buffers");

__device__ __forceinline__ uint32_t hm_vsad4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ const uint8_t *hm_plane(const uint8_t *pool, const MePicGeom &g, uint32_t slot, int level) {
    return pool + (size_t)slot * g.slot_bytes + g.plane[level].offset;
}

/* floor(i / d) for i < 2^20, d < 2^12: one IMAD.HI instead of a division sequence */
__device__ __forceinline__ uint32_t hm_magic(uint32_t d) { return d <= 1 ? 0u : (uint32_t)(0xFFFFFFFFu / d) + 1u; }
__device__ __forceinline__ uint32_t hm_div(uint32_t i, uint32_t d, uint32_t magic) { return d <= 1 ? i : __umulhi(i, magic); }

/* stage `nrows` window rows of `pitch` aligned words starting at g0 into win (one warp) */
__device__ __forceinline__ void hm_stage_window(uint32_t *win, const uint32_t *g0, int stride_words, int nrows, int pitch,
                                                int lane) {
    const int      n = nrows * pitch;
    const uint32_t m = hm_magic((uint32_t)pitch);
#pragma unroll 4
    for (int i = lane; i < n; i += 32) {
        const int r = (int)hm_div((uint32_t)i, (uint32_t)pitch, m), cw = i - r * pitch;
        win[i]      = __ldg(g0 + (size_t)r * stride_words + cw);
    }
}

/* ------------------------------------------------------------------------------------------------
 * svt_sad_loop_kernel_c over the evaluated search rows [e_lo, e_hi) of one task, executed by one warp.
 * Returns the warp's best (sad << 32 | scan index) key (the same value in every lane).
 * ---------------------------------------------------------------------------------------------- */
/* out of line, and the row loops of the window executors not unrolled (below): k_hme is bound by instruction fetch as much as
   by anything (342 KB of SASS, five CTAs per SM at different points of it): the rarely used per-warp path kept out of the hot
   code and a 4x smaller generic SAD loop are worth +8 % on the 1080p headline (4700 -> 5080 frames/s), unrolling by 8 costs 6 % */
__device__ __noinline__ unsigned long long search_item(const SearchTask &t, int e_lo, int e_hi, const MeJob *j, const uint8_t *pool,
                                          uint32_t *buf, unsigned int *overflow, int lane) {
    const MePlaneGeom &pl   = j->g.plane[t.level];
    const uint8_t     *srcp = hm_plane(pool, j->g, j->src_slot, t.level);
    const uint8_t     *refp = hm_plane(pool, j->g, t.ref_slot, t.level);
    const int          bw = t.bw, rows = t.rows, step = t.row_step;
    const int          nw  = (bw + 3) >> 2; /* words per block row */
    const int          nwp = (nw + 3) & ~3; /* padded to whole 16-byte chunks: 4, 8, 12 or 16 */
    const int          stride_words = pl.stride >> 2;
    const int          src_words = rows * nwp;
    uint32_t          *srcw = buf, *win = buf + src_words;
    const int          win_cap = HM_BUF_WORDS - src_words;
    const unsigned long long none = ((unsigned long long)ME_SAD_LOOP_INIT << 32) | 0xFFFFFFFFull;

    /* source block -> shared words, zero beyond the block width (source rows are 4-byte aligned) */
    {
        const uint8_t *sp = srcp + (size_t)(pl.pad + t.sy) * pl.stride + pl.pad + t.sx;
        const uint32_t m  = hm_magic((uint32_t)nwp);
        for (int i = lane; i < src_words; i += 32) {
            const int r = (int)hm_div((uint32_t)i, (uint32_t)nwp, m), cw = i - r * nwp;
            uint32_t  v = 0;
            if (cw < nw) {
                v = __ldg((const uint32_t *)(sp + (size_t)(r * step) * pl.stride) + cw);
                const int nb = bw - cw * 4;
                if (nb < 4)
                    v &= (1u << (8 * nb)) - 1u;
            }
            srcw[i] = v;
        }
    }
    const uint8_t *w00   = refp + (size_t)(pl.pad + t.ry) * pl.stride + pl.pad + t.rx; /* search point (0,0) */
    const int      a     = (int)((uintptr_t)w00 & 3);
    const int      pitch = (((a + t.sa_w - 1 + bw + 3) >> 2) + 1) | 1; /* odd: rows spread over the banks */
    const bool     skip  = bw == 16 && rows <= 16 && t.skip_line;
    const int      span  = (rows - 1) * step;
    const int      rows_avail = win_cap / pitch;
    const int      band  = skip ? ((rows_avail - span - 1) / 2 + 1) : (rows_avail - span);
    if (rows_avail < span + (skip ? 2 : 1)) { /* window wider than the staging buffer: reported, never silent */
        if (lane == 0)
            atomicExch(overflow, 2u);
        return none;
    }
    const int      chunks = nwp >> 2;
    const bool     modeB  = rows * chunks >= 32;
    const uint32_t saw = (uint32_t)t.sa_w, sam = hm_magic(saw);

    uint32_t best = ME_SAD_LOOP_INIT, bidx = 0xFFFFFFFFu;
    for (int e0 = e_lo; e0 < e_hi; e0 += band) {
        const int e1   = min(e_hi, e0 + band);
        const int ymin = skip ? 2 * e0 + 1 : e0;
        const int ymax = skip ? 2 * (e1 - 1) + 1 : e1 - 1;
        __syncwarp();
        hm_stage_window(win, (const uint32_t *)(w00 - a + (ptrdiff_t)ymin * pl.stride), stride_words, ymax - ymin + span + 1,
                        pitch, lane);
        __syncwarp();
        const int npos = (e1 - e0) * t.sa_w;
        if (!modeB && bw == 16) {
            /* lane = search position; 16-byte block rows: one LDS.128 of source, 5 words of window per row */
            const int rstep = pitch * step;
            for (int p = lane; p < npos; p += 32) {
                const int       ee = (int)hm_div((uint32_t)p, saw, sam), x = p - ee * t.sa_w;
                const int       o  = a + x, sh = (o & 3) * 8;
                const uint32_t *wp = win + (skip ? 2 * ee : ee) * pitch + (o >> 2);
                const uint4    *sp = (const uint4 *)srcw;
                uint32_t        s0 = 0, s1 = 0, s2 = 0, s3 = 0; /* independent chains: the SAD pipe has 4-cycle latency */
#pragma unroll 4
                for (int r = 0; r < rows; r++) {
                    const uint4    s4 = sp[r];
                    const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3], w4 = wp[4];
                    s0 = hm_vsad4(s4.x, __funnelshift_r(w0, w1, sh), s0);
                    s1 = hm_vsad4(s4.y, __funnelshift_r(w1, w2, sh), s1);
                    s2 = hm_vsad4(s4.z, __funnelshift_r(w2, w3, sh), s2);
                    s3 = hm_vsad4(s4.w, __funnelshift_r(w3, w4, sh), s3);
                    wp += rstep;
                }
                const uint32_t sad = (s0 + s1) + (s2 + s3);
                if (sad < best) { /* positions arrive in increasing scan order per lane: first minimum kept */
                    best = sad;
                    bidx = (uint32_t)((e0 + ee) * t.sa_w + x);
                }
            }
        } else if (!modeB) {
            /* lane = search position, any block width (picture-edge blocks) */
            for (int p = lane; p < npos; p += 32) {
                const int       ee = (int)hm_div((uint32_t)p, saw, sam), x = p - ee * t.sa_w;
                const int       o  = a + x, sh = (o & 3) * 8;
                const uint32_t *wp = win + (skip ? 2 * ee : ee) * pitch + (o >> 2);
                const uint32_t *sp = srcw;
                uint32_t        sad = 0;
                for (int r = 0; r < rows; r++) {
                    uint32_t prev = wp[0];
                    for (int cw = 0; cw < nw; cw++) {
                        const uint32_t next = wp[cw + 1];
                        uint32_t       v    = __funnelshift_r(prev, next, sh);
                        prev                = next;
                        const int nb        = bw - cw * 4;
                        if (nb < 4)
                            v &= (1u << (8 * nb)) - 1u;
                        sad = hm_vsad4(sp[cw], v, sad);
                    }
                    wp += pitch * step;
                    sp += nwp;
                }
                if (sad < best) {
                    best = sad;
                    bidx = (uint32_t)((e0 + ee) * t.sa_w + x);
                }
            }
        } else {
            /* lane = (block row, 16-byte chunk) items; positions one after the other, redux.add per position */
            const int  items = rows * chunks;
            const bool masked = (bw & 15) != 0 || (items & 31) != 0;
            uint32_t   sw[4][4], mk[4][4];
            int        off[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int  it = lane + 32 * k;
                const bool on = it < items;
                const int  r = on ? it / chunks : 0, ch = on ? it - r * chunks : 0;
                off[k]       = r * step * pitch + ch * 4;
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int nb = bw - (ch * 4 + q) * 4;
                    mk[k][q]     = !on || nb <= 0 ? 0u : (nb < 4 ? (1u << (8 * nb)) - 1u : 0xFFFFFFFFu);
                    sw[k][q]     = on ? srcw[r * nwp + ch * 4 + q] : 0u;
                }
            }
            const int npass = (items + 31) >> 5;
            for (int p = 0; p < npos; p++) {
                const int       ee = (int)hm_div((uint32_t)p, saw, sam), x = p - ee * t.sa_w;
                const int       o  = a + x, sh = (o & 3) * 8;
                const uint32_t *wp = win + (skip ? 2 * ee : ee) * pitch + (o >> 2);
                uint32_t        p0 = 0, p1 = 0, p2 = 0, p3 = 0;
                if (!masked) {
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        if (k < npass) {
                            const uint32_t *q  = wp + off[k];
                            const uint32_t  w0 = q[0], w1 = q[1], w2 = q[2], w3 = q[3], w4 = q[4];
                            p0 = hm_vsad4(sw[k][0], __funnelshift_r(w0, w1, sh), p0);
                            p1 = hm_vsad4(sw[k][1], __funnelshift_r(w1, w2, sh), p1);
                            p2 = hm_vsad4(sw[k][2], __funnelshift_r(w2, w3, sh), p2);
                            p3 = hm_vsad4(sw[k][3], __funnelshift_r(w3, w4, sh), p3);
                        }
                } else {
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        if (k < npass) {
                            const uint32_t *q  = wp + off[k];
                            const uint32_t  w0 = q[0], w1 = q[1], w2 = q[2], w3 = q[3], w4 = q[4];
                            p0 = hm_vsad4(sw[k][0], __funnelshift_r(w0, w1, sh) & mk[k][0], p0);
                            p1 = hm_vsad4(sw[k][1], __funnelshift_r(w1, w2, sh) & mk[k][1], p1);
                            p2 = hm_vsad4(sw[k][2], __funnelshift_r(w2, w3, sh) & mk[k][2], p2);
                            p3 = hm_vsad4(sw[k][3], __funnelshift_r(w3, w4, sh) & mk[k][3], p3);
                        }
                }
                const uint32_t sad = __reduce_add_sync(0xffffffffu, (p0 + p1) + (p2 + p3));
                if (sad < best) {
                    best = sad;
                    bidx = (uint32_t)((e0 + ee) * t.sa_w + x);
                }
            }
        }
    }
    unsigned long long key = ((unsigned long long)best << 32) | bidx;
    if (!modeB) /* warp minimum of (sad, scan index); in the row mode every lane already holds the same key */
        for (int o = 16; o; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
            key                            = other < key ? other : key;
        }
    return key;
}

/* ------------------------------------------------------------------------------------------------
 * Window jobs: the whole CTA works on one search window at a time.
 *
 * A job is a band of a task's search area (all of it when the window fits). Its reference window is
 * fetched by TMA (cp.async.bulk.tensor: one box, or a few stacked ones, form a dense row-major window whose
 * row pitch is the box width) one job ahead of the SAD loop, into one of two buffers, and completes on the
 * buffer's mbarrier. TMA boxes start on 16-byte columns, so search position x sits at byte
 * aoff + x of a window row; unaligned words are funnel-shifted in registers.
 *
 * SAD loop: a unit = KX positions of one search row that share their byte alignment (x = ca + 4k), so
 * that the NW + KX - 1 shifted reference words of a block row feed NW * KX VABSDIFF4; units of small
 * search areas are split over G lanes by block rows and joined by shuffles. Results merge as
 * (sad << 32 | scan index) keys: the minimum key is the first minimum of the reference's scan.
 * ---------------------------------------------------------------------------------------------- */
__device__ __forceinline__ uint32_t hm_smem(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void hm_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void hm_mbar_expect(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void hm_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done) /* suspend-time hint: a waiting warp should not take issue slots from the warps it waits for */
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done)
                     : "r"(bar), "r"(parity), "r"(1000u)
                     : "memory");
}
__device__ __forceinline__ void hm_tma_box(uint32_t dst, const void *map, int x, int y, int z, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ void hm_tma_prefetch_l2(const void *map, int x, int y, int z) {
    asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(map), "r"(x), "r"(y), "r"(z) : "memory");
}

/* box widths of the tensor maps, by plane: sixteenth {48, 80, 224}, quarter {80}, full {96, 128}; map 6 (full plane,
 * 128 x 96) is the window of a k_fullsearch tile (me_kernels.cu) */
__host__ __device__ __forceinline__ int hm_map_width(int m) {
    return m == 0 ? 48 : m == 1 ? 80 : m == 2 ? 224 : m == 3 ? 80 : m == 4 ? 96 : 128;
}
__host__ __device__ __forceinline__ int hm_map_level(int m) { return m < 3 ? 2 : m == 3 ? 1 : 0; }
/* rows per box: the usual window of the stage that uses the map in ONE box (the TMA unit takes a box every ~35 cycles
 * per SM whatever its size: tools/tma_bench.cu); taller windows stack boxes */
__host__ __device__ __forceinline__ int hm_map_rows(int m) {
    return m == 0 ? 64 : m == 1 ? 56 : m == 2 ? 24 : m == 3 ? 40 : m == 6 ? 96 : 72;
}

/* cut one task into window jobs (or leave it to the per-warp path: picture-edge blocks) */
__device__ void hm_plan_task(const SearchTask &t, const MeJob *j, int org_x, int org_y, HmPlan &pl) {
    pl.new_path = 0;
    pl.nxb = pl.nyb = 0;
    const int lv = t.level, full = 64 >> lv;
    if (t.bw != full || t.sx != (org_x >> lv) || t.sy != (org_y >> lv) || (int)t.rows * t.row_step > full || t.sa_w < 1)
        return;
    const MePlaneGeom &g    = j->g.plane[lv];
    const bool         skip = t.bw == 16 && t.rows <= 16 && t.skip_line;
    const int          ymul = skip ? 2 : 1, erows = skip ? (t.sa_h >> 1) : t.sa_h;
    if (erows < 1)
        return;
    const int span = (t.rows - 1) * t.row_step;
    const int a16  = (g.pad + t.rx) & 15;
    const int first = lv == 2 ? 0 : lv == 1 ? 3 : 4, last = lv == 2 ? 2 : lv == 1 ? 3 : 5;
    const int maxw  = hm_map_width(last);
    int       pwb = t.sa_w, nxb = 1;
    if (a16 + t.sa_w - 1 + t.bw > maxw) { /* too wide for one box: bands of a multiple of 16 positions (same alignment) */
        pwb = (maxw - 15 - t.bw + 1) & ~15;
        nxb = (t.sa_w + pwb - 1) / pwb;
    }
    const int need = a16 + pwb - 1 + t.bw;
    int       map  = first;
    while (map < last && hm_map_width(map) < need) map++;
    const int pitch    = hm_map_width(map);
    const int rows_cap = (HM_RAW_BYTES / pitch) / hm_map_rows(map) * hm_map_rows(map);
    const int ne_max   = (rows_cap - span - 1) / ymul + 1;
    if (rows_cap < span + 1 || ne_max < 1 || nxb > 255)
        return;
    int       kx      = (t.bw == 16 && pwb >= 16) ? 4 : 2;
    const int ncls    = min(4, pwb), per_cls = (pwb + 3) >> 2, nchunk = (per_cls + kx - 1) / kx;
    const int nyb = (erows + ne_max - 1) / ne_max, neb = (erows + nyb - 1) / nyb;
    if (nyb > 255)
        return;
    int       units   = ncls * nchunk * neb;
    /* wide areas of 16-byte block rows: units of 8 positions on 32-byte window segments (hm_job_search8), one lane each */
    const bool wide = t.bw == 16 && pwb >= HM_K8_MIN_PW && 4 * ((pwb + 31) >> 5) * neb >= HM_K8_MIN_UNITS;
    if (wide) {
        kx    = 8;
        units = 4 * (((15 + pwb - 1) >> 5) + 1) * neb;
    }
    /* tall thin areas (pre-HME region 0: 8 positions wide, hundreds of rows) of 16x8-row blocks: hm_job_search_tall */
    const bool tall = !wide && HM_TALL_MIN_ROWS > 0 && t.bw == 16 && t.rows == 8 && nxb == 1 && pwb <= 8 && neb >= HM_TALL_MIN_ROWS &&
                      t.row_step % ymul == 0;
    if (tall) {
        const int ys = t.row_step / ymul;
        kx    = 1;
        units = 4 * ys * ((neb + 4 * ys - 1) / (4 * ys));
    }
    const int win     = pitch * (((neb - 1) * ymul + span + 1 + hm_map_rows(map) - 1) / hm_map_rows(map) * hm_map_rows(map));
    /* jobs of few units are searched by one warp each, four at a time per CTA, when four windows fit the ring (measured:
       two warps per job with the next batch in flight is slower than one warp per job and no prefetch of the next batch) */
    pl.team = (uint8_t)(!wide && !tall && units <= HM_TEAM1_UNITS && win <= HM_RING_BYTES / 4 ? 1 : 4);
    int g_log2 = 2 - lv; /* lanes per unit: the 16-byte column segments of the block row ... */
    if (tall)
        g_log2 = 0;
    else if (!wide && pl.team < 4) {
        /* one warp per job: the split with the fewest (passes over the warp) x (block rows per lane); ties keep the
           smaller group (fewer shuffles). HME level 2 (12 units of 32 rows x 4 segments): 8 lanes per unit, 3 x 16
           row iterations instead of 2 x 32 */
        int best_cost = 1 << 30;
        for (int gl = 2 - lv; gl <= 5 && ((1 << gl) >> (2 - lv)) <= (int)t.rows; gl++) {
            const int rg = (1 << gl) >> (2 - lv);
            const int cost = (((units << gl) + 31) >> 5) * (((int)t.rows + rg - 1) / rg);
            if (cost < best_cost) {
                best_cost = cost;
                g_log2    = gl;
            }
        }
    } else
        while (!wide && g_log2 < 5 && (units << g_log2) < HM_LANES_TARGET && ((2 << g_log2) >> (2 - lv)) <= (int)t.rows)
            g_log2++; /* ... x row groups: up to the team's worth of lanes */
    pl.pwb = (int16_t)pwb;
    pl.neb = (int16_t)neb;
    pl.nxb = (uint8_t)nxb;
    pl.nyb = (uint8_t)nyb;
    pl.map = (uint8_t)map;
    pl.kx  = (uint8_t)kx;
    pl.g_log2   = (uint8_t)g_log2;
    pl.new_path = 1;
}

/* job number -> job descriptor (one thread per job) */
__device__ void hm_job_decode(const SearchTask *tasks, const HmPlan *plans, const uint16_t *job_first, const MeJob *j, int jb, HmJobD &J) {
    int task = 0;
    for (int step = 16; step; step >>= 1) /* last task with job_first[task] <= jb */
        if (job_first[task + step] <= jb)
            task += step;
    const SearchTask  &t  = tasks[task];
    const HmPlan      &pl = plans[task];
    const MePlaneGeom &g  = j->g.plane[t.level];
    const int  k = jb - job_first[task], yb = k / pl.nxb, xb = k - yb * pl.nxb;
    const bool skip = t.bw == 16 && t.rows <= 16 && t.skip_line;
    const int  ymul = skip ? 2 : 1, erows = skip ? (t.sa_h >> 1) : t.sa_h;
    const int  px0 = xb * pl.pwb, pw = min((int)pl.pwb, t.sa_w - px0);
    const int  e0 = yb * pl.neb, ne = min((int)pl.neb, erows - e0);
    const int  x0 = g.pad + t.rx + px0, br = hm_map_rows(pl.map);
    const int  ncls = min(4, pw), nchunk = (((pw + 3) >> 2) + pl.kx - 1) / pl.kx;
    J.task   = (uint8_t)task;
    J.map    = pl.map;
    J.px0    = (int16_t)px0;
    J.pw     = (int16_t)pw;
    J.e0     = (int16_t)e0;
    J.ne     = (int16_t)ne;
    J.xt     = (int16_t)(x0 & ~15);
    J.aoff   = (uint8_t)(x0 & 15);
    J.y0     = (int16_t)(g.pad + t.ry + ymul * e0 + (skip ? 1 : 0));
    J.nrows  = (uint16_t)(((ne - 1) * ymul + (t.rows - 1) * t.row_step + 1 + br - 1) / br * br); /* whole boxes */
    J.off128 = 0;
    J.ncls   = (uint8_t)ncls;
    J.nchunk = (uint8_t)nchunk;
    J.total  = (uint16_t)((ncls * nchunk * ne) << pl.g_log2);
    J.dep    = -1;
    J.m_chunk = hm_magic((uint32_t)nchunk);
    J.m_cls   = hm_magic((uint32_t)ncls);
    if (pl.kx == 1) { /* hm_job_search_tall: (byte class, row parity, group of 4 * ys search rows) */
        const int ys = t.row_step / ymul, ng = (ne + 4 * ys - 1) / (4 * ys);
        J.ncls    = (uint8_t)ys;
        J.nchunk  = (uint8_t)ng;
        J.total   = (uint16_t)(4 * ys * ng);
    }
    if (pl.kx == 8) { /* hm_job_search8: (byte class, 32-byte segment of the window row, search row) */
        const int nseg = (((int)J.aoff + pw - 1) >> 5) + 1;
        J.ncls    = 4;
        J.nchunk  = (uint8_t)nseg;
        J.total   = (uint16_t)(4 * nseg * ne);
        J.m_chunk = hm_magic((uint32_t)nseg);
    }
}

/* Thread 0, before the CTA searches job `jb`: request the windows of the following jobs while the ring has room.
 * Jobs are searched in order with a CTA barrier after each, so the jobs before `jb` are done: job n can go as soon as
 * the last earlier job on its ring bytes (dep) is one of them. Job n (N = seq0 + n over the whole kernel) completes
 * on mbarrier N % HM_NBAR; at most HM_NBAR jobs are in flight, so the previous user of the mbarrier is done too. */
__device__ __forceinline__ void hm_fetch_ahead(const HmeMaps &maps, const SearchTask *tasks, const HmJobD *jobs, unsigned long long *full,
                                               uint32_t ring, int njobs, uint32_t seq0, int jb, int &fetched) {
    while (fetched < njobs && fetched - jb < HM_NBAR && jobs[fetched].dep < jb) {
        const HmJobD  &Jn  = jobs[fetched];
        const uint32_t bar = hm_smem(&full[(seq0 + (uint32_t)fetched) % HM_NBAR]);
        const int      pitch = hm_map_width(Jn.map), off = (int)Jn.off128 * 128, br = hm_map_rows(Jn.map);
        hm_mbar_expect(bar, (uint32_t)((int)Jn.nrows * pitch));
        for (int y = 0; y < (int)Jn.nrows; y += br)
            hm_tma_box(ring + (uint32_t)(off + y * pitch), maps.m[Jn.map], Jn.xt, Jn.y0 + y, tasks[Jn.task].ref_slot, bar);
        fetched++;
    }
}

/* the SAD loop of one job over the CTA; returns this warp's best key (same value in every lane).
 * A unit (KX positions) is spread over G = 2^g_log2 lanes: the 16-byte column segments of the block row (bw / 16 of
 * them, adjacent lanes -> adjacent banks) times row groups; partial sums are joined by shuffles. */
template <int KX>
__device__ __forceinline__ unsigned long long hm_job_search(const HmJobD &J, const SearchTask &t, int g_log2, const uint32_t *raw,
                                                            const uint32_t *src, int src_pitch4, int warp, int nthr, int lane) {
    const bool skip = t.bw == 16 && t.rows <= 16 && t.skip_line;
    const int  ymul = skip ? 2 : 1, rows = t.rows, step = t.row_step;
    const int  pitchw = hm_map_width(J.map) >> 2, G = 1 << g_log2;
    const int  ncs_log2 = 2 - t.level, RG = G >> ncs_log2; /* column segments: 1 / 2 / 4 on the sixteenth / quarter / full plane */
    const int  ncls = J.ncls, nchunk = J.nchunk, total = J.total;
    const uint32_t m_chunk = J.m_chunk, m_cls = J.m_cls;
    unsigned long long key = ((unsigned long long)ME_SAD_LOOP_INIT << 32) | 0xFFFFFFFFull;
    for (int base = warp * 32; base < total; base += nthr) {
        const int  u  = base + lane;
        const bool on = u < total;
        const int  g = u & (G - 1), v = u >> g_log2;
        const int  cs = g & ((1 << ncs_log2) - 1), rg = g >> ncs_log2;
        const int  v1 = (int)hm_div((uint32_t)v, (uint32_t)nchunk, m_chunk), chunk = v - v1 * nchunk;
        const int  e = (int)hm_div((uint32_t)v1, (uint32_t)ncls, m_cls), ca = v1 - e * ncls;
        const int  o = J.aoff + ca + 4 * KX * chunk, sh = (o & 3) * 8;
        uint32_t   acc[KX];
#pragma unroll
        for (int k = 0; k < KX; k++) acc[k] = 0;
        if (on) {
            const uint32_t *wp = raw + (e * ymul + rg * step) * pitchw + (o >> 2) + 4 * cs;
            const uint4    *sp = reinterpret_cast<const uint4 *>(src) + rg * step * src_pitch4 + cs;
            const int       dw = step * pitchw * RG, ds = step * src_pitch4 * RG;
#pragma unroll 1
            for (int r = rg; r < rows; r += RG) {
                const uint4 s4 = *sp;
                uint32_t    w[4 + KX];
#pragma unroll
                for (int q = 0; q < 4 + KX; q++) w[q] = wp[q];
#pragma unroll
                for (int q = 0; q < 3 + KX; q++) w[q] = __funnelshift_r(w[q], w[q + 1], sh);
#pragma unroll
                for (int k = 0; k < KX; k++) {
                    acc[k] = hm_vsad4(s4.x, w[k], acc[k]);
                    acc[k] = hm_vsad4(s4.y, w[k + 1], acc[k]);
                    acc[k] = hm_vsad4(s4.z, w[k + 2], acc[k]);
                    acc[k] = hm_vsad4(s4.w, w[k + 3], acc[k]);
                }
                wp += dw;
                sp += ds;
            }
        }
        for (int d = G >> 1; d; d >>= 1)
#pragma unroll
            for (int k = 0; k < KX; k++) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], d);
        if (on && g == 0) {
            const uint32_t idx0 = (uint32_t)((J.e0 + e) * t.sa_w + J.px0);
#pragma unroll
            for (int k = 0; k < KX; k++) {
                const int x = ca + 4 * (KX * chunk + k);
                if (x < J.pw && acc[k] < ME_SAD_LOOP_INIT) {
                    const unsigned long long cand = ((unsigned long long)acc[k] << 32) | (idx0 + (uint32_t)x);
                    key                           = cand < key ? cand : key;
                }
            }
        }
    }
    for (int d = 16; d; d >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, d);
        key                            = other < key ? other : key;
    }
    return key;
}

/* Wide search areas of 16-byte block rows (pre-HME / HME level 0 on the sixteenth plane). A lane owns the 8 positions
 * whose block rows start at bytes 32m + c + 4k (k = 0..7) of window row e: the 12 aligned words of the segment come in
 * with three LDS.128 (adjacent lanes = the four byte classes c of one segment: same words, broadcast; a quarter warp
 * reads 64 contiguous bytes: no bank conflicts, where the 4-word lane stride of hm_job_search<4> costs four wavefronts
 * per load), 11 funnel shifts build the shifted words, and a block row is 32 VABSDIFF4. Positions left of the job's
 * first column (byte aoff) or beyond its width are computed and dropped. */
__device__ __forceinline__ unsigned long long hm_job_search8(const HmJobD &J, const SearchTask &t, const uint32_t *raw,
                                                             const uint32_t *src, int src_pitch4, int warp, int nthr, int lane) {
    const bool skip = t.bw == 16 && t.rows <= 16 && t.skip_line;
    const int  ymul = skip ? 2 : 1, rows = t.rows, step = t.row_step;
    const int  pitch4 = hm_map_width(J.map) >> 4; /* window row pitch in 16-byte units */
    const int  nseg = J.nchunk, total = J.total;
    const uint32_t m_seg = J.m_chunk, pw = (uint32_t)J.pw;
    /* inside the job a 32-bit key is enough: SAD of a 16-byte-wide block of <= 16 rows < 2^16, and a job has fewer
       positions than its window has bytes (<= 16384): (sad << 16 | e * pw + x) orders like the 64-bit key */
    uint32_t best = 0xFFFFFFFFu;
    for (int base = warp * 32; base < total; base += nthr) {
        const int  u  = base + lane;
        const bool on = u < total;
        const int  c = u & 3, v = u >> 2;
        const int  e = (int)hm_div((uint32_t)v, (uint32_t)nseg, m_seg), m = v - e * nseg;
        const int  sh = 8 * c;
        uint32_t   acc[8];
#pragma unroll
        for (int k = 0; k < 8; k++) acc[k] = 0;
        if (on) {
            const uint4 *wp = reinterpret_cast<const uint4 *>(raw) + e * ymul * pitch4 + 2 * m;
            const uint4 *sp = reinterpret_cast<const uint4 *>(src);
            const int    dw = step * pitch4, ds = step * src_pitch4;
#pragma unroll 2
            for (int r = 0; r < rows; r++) {
                const uint4 s4 = *sp;
                const uint4 a = wp[0], b = wp[1], d = wp[2];
                uint32_t    w[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, d.x, d.y, d.z, d.w};
#pragma unroll
                for (int q = 0; q < 11; q++) w[q] = __funnelshift_r(w[q], w[q + 1], sh);
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    acc[k] = hm_vsad4(s4.x, w[k], acc[k]);
                    acc[k] = hm_vsad4(s4.y, w[k + 1], acc[k]);
                    acc[k] = hm_vsad4(s4.z, w[k + 2], acc[k]);
                    acc[k] = hm_vsad4(s4.w, w[k + 3], acc[k]);
                }
                wp += dw;
                sp += ds;
            }
            const int      x0 = 32 * m + c - (int)J.aoff;
            const uint32_t l0 = (uint32_t)(e * (int)pw + x0); /* wraps for x0 < 0: only used where the position is valid */
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint32_t cand = (acc[k] << 16) + l0 + 4u * k;
                best                = min(best, (uint32_t)(x0 + 4 * k) < pw ? cand : 0xFFFFFFFFu);
            }
        }
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (best == 0xFFFFFFFFu)
        return ((unsigned long long)ME_SAD_LOOP_INIT << 32) | 0xFFFFFFFFull;
    const uint32_t lidx = best & 0xFFFFu, e = lidx / pw, x = lidx - e * pw;
    return ((unsigned long long)(best >> 16) << 32) | (uint32_t)((J.e0 + (int)e) * t.sa_w + J.px0 + (int)x);
}

/* Tall thin search areas (<= 8 positions wide) of 16-byte x 8-row blocks: pre-HME region 0 (8 x 100..400). A lane owns
 * the two positions x = c, c + 4 of FOUR search rows whose block rows fall on the same window rows (e, e + ys, e + 2 ys,
 * e + 3 ys with ys = row_step / ymul: block row r of search row e + i ys is window row (e ymul) + (i + r) row_step), so
 * 11 window rows - 6 words and 5 funnel shifts each - feed the 8 x 32 VABSDIFF4 of eight positions, where one lane per
 * search row loads 8 window rows for two. 32-bit keys as in hm_job_search8. */
__device__ __forceinline__ unsigned long long hm_job_search_tall(const HmJobD &J, const SearchTask &t, const uint32_t *raw,
                                                                 const uint32_t *src, int src_pitch4, int warp, int nthr, int lane) {
    const bool skip = t.bw == 16 && t.rows <= 16 && t.skip_line;
    const int  ymul = skip ? 2 : 1, step = t.row_step, ys = J.ncls /* row_step / ymul: 1 or 2 */;
    const int  pitchw = hm_map_width(J.map) >> 2, total = J.total, ne = J.ne, maxrow = (int)J.nrows - 1;
    const uint32_t pw = (uint32_t)J.pw;
    uint32_t   best = 0xFFFFFFFFu;
    for (int base = warp * 32; base < total; base += nthr) {
        const int u = base + lane;
        if (u < total) {
            const int c = u & 3, v = u >> 2;
            const int par = v & (ys - 1), g = ys == 2 ? v >> 1 : v;
            const int e0 = 4 * ys * g + par, o = (int)J.aoff + c, sh = (o & 3) * 8;
            const uint32_t *wbase = raw + (o >> 2);
            const uint4    *sp = reinterpret_cast<const uint4 *>(src);
            uint32_t acc[4][2];
#pragma unroll
            for (int i = 0; i < 4; i++) acc[i][0] = acc[i][1] = 0;
            uint4 S[8];
#pragma unroll
            for (int k = 0; k < 11; k++) {
                const uint32_t *q = wbase + min(e0 * ymul + k * step, maxrow) * pitchw; /* rows past the window: dropped positions only */
                uint32_t w[6];
#pragma unroll
                for (int i = 0; i < 6; i++) w[i] = q[i];
#pragma unroll
                for (int i = 0; i < 5; i++) w[i] = __funnelshift_r(w[i], w[i + 1], sh);
                if (k < 8)
                    S[k] = sp[k * step * src_pitch4];
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const int r = k - i;
                    if (r >= 0 && r < 8) {
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            acc[i][h] = hm_vsad4(S[r].x, w[h], acc[i][h]);
                            acc[i][h] = hm_vsad4(S[r].y, w[h + 1], acc[i][h]);
                            acc[i][h] = hm_vsad4(S[r].z, w[h + 2], acc[i][h]);
                            acc[i][h] = hm_vsad4(S[r].w, w[h + 3], acc[i][h]);
                        }
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int e = e0 + i * ys;
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const uint32_t x = (uint32_t)(c + 4 * h), cand = (acc[i][h] << 16) + (uint32_t)e * pw + x;
                    best = min(best, x < pw && e < ne ? cand : 0xFFFFFFFFu);
                }
            }
        }
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (best == 0xFFFFFFFFu)
        return ((unsigned long long)ME_SAD_LOOP_INIT << 32) | 0xFFFFFFFFull;
    const uint32_t lidx = best & 0xFFFFu, e = lidx / pw, x = lidx - e * pw;
    return ((unsigned long long)(best >> 16) << 32) | (uint32_t)((J.e0 + (int)e) * t.sa_w + J.px0 + (int)x);
}

/* ------------------------------------------------------------------------------------------------
 * open_loop_me_fullpel_search_sblock over the search rows [y_lo, y_hi) of a small area (or the single
 * probe position), one warp. Lane z owns the 8x8 blocks with z-order index z and z + 32; the 16x16 /
 * 32x32 / 64x64 sums are butterflies over lanes (z-order: 4 consecutive lanes = one 16x16, 16 = one 32x32).
 * ---------------------------------------------------------------------------------------------- */
template <bool SUB>
__device__ void fullsearch_item(const FsTask &f, int y_lo, int y_hi, const MeJob *j, const uint8_t *pool, uint32_t *win,
                                unsigned long long *keys, int org_x, int org_y, int lane, const uint32_t *src0) {
    constexpr int      NR = SUB ? 4 : 8, RSTEP = SUB ? 2 : 1;
    const MePlaneGeom &pl = j->g.plane[0];
    const int          stride_words = pl.stride >> 2;
    /* z-order -> block coordinates */
    const int bx = (lane & 1) | ((lane >> 1) & 2) | ((lane >> 2) & 4);
    const int by = ((lane >> 1) & 1) | ((lane >> 2) & 2); /* + 4 for the second block */
    uint32_t  sw[2][NR][2];
    {
        /* the b64's source block is resident in shared memory (HmeShared.src0, pitch HM_SRC0_PITCH words) */
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int r = 0; r < NR; r++) {
                const uint2 v = *reinterpret_cast<const uint2 *>(src0 + (8 * (by + 4 * b) + r * RSTEP) * HM_SRC0_PITCH + 2 * bx);
                sw[b][r][0]   = v.x;
                sw[b][r][1]   = v.y;
            }
    }
    const uint8_t *w00   = hm_plane(pool, j->g, f.ref_slot, 0) + (size_t)(pl.pad + org_y + f.y0) * pl.stride + pl.pad + org_x + f.x0;
    const int      a     = (int)((uintptr_t)w00 & 3);
    const int      pitch = (((a + f.w - 1 + 64 + 3) >> 2) + 1) | 1; /* <= 31 words for w <= FS_WARP_MAXW */
    const int      span  = SUB ? 62 : 63;
    const int      band  = HM_BUF_WORDS / pitch - span; /* >= 5 rows of positions */
    uint32_t bs8[2] = {0xFFFFFFFFu, 0xFFFFFFFFu}, bs16[2] = {0xFFFFFFFFu, 0xFFFFFFFFu}, bs32[2] = {0xFFFFFFFFu, 0xFFFFFFFFu},
             bs64 = 0xFFFFFFFFu;
    uint32_t bi8[2] = {0, 0}, bi16[2] = {0, 0}, bi32[2] = {0, 0}, bi64 = 0;
    for (int y0 = y_lo; y0 < y_hi; y0 += band) {
        const int y1 = min(y_hi, y0 + band);
        __syncwarp();
        hm_stage_window(win, (const uint32_t *)(w00 - a + (ptrdiff_t)y0 * pl.stride), stride_words, y1 - y0 + span, pitch, lane);
        __syncwarp();
        for (int y = y0; y < y1; y++)
            for (int x = 0; x < f.w; x++) {
                const int o = a + x + 8 * bx, sh = (o & 3) * 8;
                uint32_t  a8[2];
#pragma unroll
                for (int b = 0; b < 2; b++) {
                    const uint32_t *wp = win + (y - y0 + 8 * (by + 4 * b)) * pitch + (o >> 2);
                    uint32_t        acc0 = 0, acc1 = 0;
#pragma unroll
                    for (int r = 0; r < NR; r++) {
                        const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2];
                        acc0 = hm_vsad4(sw[b][r][0], __funnelshift_r(w0, w1, sh), acc0);
                        acc1 = hm_vsad4(sw[b][r][1], __funnelshift_r(w1, w2, sh), acc1);
                        wp += pitch * RSTEP;
                    }
                    a8[b] = acc0 + acc1;
                }
                uint32_t s16[2], s32[2];
#pragma unroll
                for (int b = 0; b < 2; b++) {
                    uint32_t v = a8[b];
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    s16[b] = v;
                    v += __shfl_xor_sync(0xffffffffu, v, 4);
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    s32[b] = v;
                }
                uint32_t s64 = s32[0] + s32[1];
                s64 += __shfl_xor_sync(0xffffffffu, s64, 16);
                const uint32_t idx = (uint32_t)(y * f.w + x);
#pragma unroll
                for (int b = 0; b < 2; b++) {
                    if (a8[b] < bs8[b]) { bs8[b] = a8[b]; bi8[b] = idx; }
                    if (s16[b] < bs16[b]) { bs16[b] = s16[b]; bi16[b] = idx; }
                    if (s32[b] < bs32[b]) { bs32[b] = s32[b]; bi32[b] = idx; }
                }
                if (s64 < bs64) { bs64 = s64; bi64 = idx; }
            }
    }
    /* merge into the task's (sad, scan order) keys: order 0 = probe position, 1 + scan index otherwise */
    const uint32_t shl = SUB ? 1 : 0;
#define HM_MERGE(pu, sad, idx) \
    atomicMin(&keys[pu], ((unsigned long long)((sad) << shl) << 32) | (f.probe ? 0u : 1u + (idx)))
#pragma unroll
    for (int b = 0; b < 2; b++) {
        HM_MERGE(21 + 32 * b + lane, bs8[b], bi8[b]);
        if ((lane & 3) == 0)
            HM_MERGE(5 + 8 * b + (lane >> 2), bs16[b], bi16[b]);
        if ((lane & 15) == 0)
            HM_MERGE(1 + 2 * b + (lane >> 4), bs32[b], bi32[b]);
    }
    if (lane == 0)
        HM_MERGE(0, bs64, bi64);
#undef HM_MERGE
}

/* ------------------------------------------------------------------------------------------------
 * k_hme
 * ---------------------------------------------------------------------------------------------- */
__global__ void __launch_bounds__(HM_ALL_THREADS, 5)
    k_hme(MeJob *jobs, FsTileList *list, const uint8_t *pool, const __grid_constant__ HmeMaps maps, int phase) {
    extern __shared__ __align__(1024) unsigned char hm_raw[];
    HmeShared     &sm  = *reinterpret_cast<HmeShared *>(hm_raw);
    MeJob         *j   = &jobs[blockIdx.y];
    const int      b   = blockIdx.x;
    const int      lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool     producer = warp == HM_WARPS; /* warp 4 only requests windows; it takes no part in strided loops */
    const int      tid = producer ? (1 << 28) : (int)threadIdx.x;
    unsigned long long *keys = &j->keys[(size_t)b * ME_NREF * ME_NPU];
    const unsigned long long none = ((unsigned long long)ME_SAD_LOOP_INIT << 32) | 0xFFFFFFFFull;

    CtrlCtx c;
    c.j     = j;
    c.p     = &sm.p;
    c.s     = &sm.s;
    c.t     = sm.t;
    c.r     = sm.r;
    c.f     = sm.f;
    c.keys  = keys;
    c.org_x = (b % j->g.b64_cols) * 64;
    c.org_y = (b / j->g.b64_cols) * 64;
    c.b64_w = (uint32_t)min(64, j->g.width - c.org_x);
    c.b64_h = (uint32_t)min(64, j->g.height - c.org_y);
    c.pic_w = j->g.width;
    c.pic_h = j->g.height;
    c.only = min(lane, ME_NREF); /* warp 0 runs the control stages: lane = (list, reference) slot, lanes >= ME_NREF only help clearing */
    c.lane = lane;
    c.nlanes = 32;
    /* phase 0: the whole chain. With enable_me_sr_adjustment == 2 the picture takes two launches around the tiled searches of
       reference (0, 0): phase 1 = the whole chain with the integer search of (0, 0) only, phase 2 = the integer-search stages
       of the other references, continuing from the state phase 1 left in HBM (me_ctrl.h, CtrlCtx.fs_which) */
    c.fs_which = phase;
    const bool sub = j->p.me_search_method == ME_GPU_SUB_SAD_SEARCH;

    for (int i = tid; i < (int)(sizeof(MeGpuParams) / 4); i += HM_THREADS)
        reinterpret_cast<uint32_t *>(&sm.p)[i] = reinterpret_cast<const uint32_t *>(&j->p)[i];
    if (phase != 2) {
        mc_init_keys(keys, tid, HM_THREADS);
        for (int i = tid; i < ME_SLOTS; i += HM_THREADS) {
            sm.t[i].valid = 0;
            sm.r[i].sad   = 0;
            sm.r[i].found = 0;
            sm.r[i].x = sm.r[i].y = 0;
        }
    } else { /* the block's state, and the check_00_center tasks and results the probe stage reads, as phase 1 left them */
        const uint32_t *src = reinterpret_cast<const uint32_t *>(&j->state[b]);
        uint32_t       *dst = reinterpret_cast<uint32_t *>(&sm.s);
        for (int i = tid; i < (int)(sizeof(B64State) / 4); i += HM_THREADS) dst[i] = src[i];
        for (int i = tid; i < ME_SLOTS; i += HM_THREADS) {
            sm.t[i] = j->tasks[(size_t)b * ME_SLOTS + i];
            sm.r[i] = j->results[(size_t)b * ME_SLOTS + i];
        }
    }
    if (tid == 0) {
        for (int i = 0; i < HM_NBAR; i++) hm_mbar_init(hm_smem(&sm.mbar[i]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    /* the b64's source block on the three planes stays in shared memory for the whole chain (rows below a short
       picture-edge block are padding: loaded, never summed) */
    if (!producer)
        for (int base = 0; base < 64 * 16 + 32 * 8 + 16 * 4; base += 4 * HM_THREADS) { /* four loads in flight per thread: HBM latency */
            uint32_t  v[4];
            uint32_t *dst[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int i = base + u * HM_THREADS + tid;
                dst[u]      = nullptr;
                if (i < 64 * 16 + 32 * 8 + 16 * 4) {
                    const int lv = i < 64 * 16 ? 0 : i < 64 * 16 + 32 * 8 ? 1 : 2;
                    const int q  = i - (lv == 0 ? 0 : lv == 1 ? 64 * 16 : 64 * 16 + 32 * 8);
                    const int r = q >> (4 - lv), cw = q & ((16 >> lv) - 1);
                    const MePlaneGeom &pl = j->g.plane[lv];
                    const uint8_t     *sp = hm_plane(pool, j->g, j->src_slot, lv) + (size_t)(pl.pad + (c.org_y >> lv) + r) * pl.stride + pl.pad +
                                        (c.org_x >> lv);
                    const int sp_w = lv == 0 ? HM_SRC0_PITCH : lv == 1 ? HM_SRC1_PITCH : HM_SRC2_PITCH;
                    dst[u]         = (lv == 0 ? sm.src0 : lv == 1 ? sm.src1 : sm.src2) + r * sp_w + cw;
                    v[u]           = __ldg((const uint32_t *)sp + cw);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (dst[u])
                    *dst[u] = v[u];
        }
    __threadfence();
    __syncthreads();
    uint32_t  seq  = 0; /* window jobs so far: job n uses mbarrier n % HM_NBAR, parity (n / HM_NBAR) & 1 */
    uint8_t  *raw0 = reinterpret_cast<uint8_t *>(&sm.buf[0][0]);
    int       fetched = 0; /* producer: jobs of the stage whose windows have been requested */

#ifdef ME_HME_TRACE
    long long tr[ST_COUNT][12];
    long long t_begin = clock64();
    unsigned long long gt_begin, gt_stage[ST_COUNT];
    unsigned smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_begin));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
#endif
    for (int st = phase == 2 ? ST_PROBE : 0; st < ST_COUNT; st++) {
        { /* stages without work for this picture's controls cost a round of CTA barriers: skipped (uniform over the grid).
             What they would have consumed stays pending for the next stage that runs (mc_consume_pending). */
            const MeGpuParams &pp = sm.p;
            const bool hme = pp.enable_hme_flag != 0;
            if ((st == ST_PREHME1 && !(pp.prehme_enable && pp.prehme_l1_early_exit)) || (st == ST_HME_L0B && !mc_l0_two_phase(&pp)) ||
                (st == ST_HME_L1 && !(hme && pp.enable_hme_level1_flag)) || (st == ST_HME_L2 && !(hme && pp.enable_hme_level2_flag)) ||
                (st == ST_MAIN && !pp.me_8x8_var_enabled))
                continue;
        }
#ifdef ME_HME_TRACE
        long long tr0 = clock64();
        long long tr_wait = 0, tr_search = 0, tr_bar = 0, tr_round = 0, tr_decode = 0, tr_first = 0, tr_mark = 0;
        int       tr_jobs = 0;
#endif
        if (warp == 0) {
            mc_run_stage(&c, st);
            if (lane == 0)
                sm.next_item = 0;
        }
        __syncthreads();
#ifdef ME_HME_TRACE
        long long tr1 = clock64();
#endif
        const bool search_stage = st <= ST_CENTRE;
        /* warp 0: cut the tasks into work items (bands of search rows) */
        if (warp == 0) {
            int n = 0, rows_per = 1, nj = 0;
            if (search_stage) {
                const SearchTask t = sm.t[lane];
                HmPlan           pl;
                pl.new_path = 0;
                if (t.valid)
                    hm_plan_task(t, j, c.org_x, c.org_y, pl);
                if (pl.new_path)
                    nj = (int)pl.nxb * (int)pl.nyb;
                if (t.valid) { /* the per-warp split, used when the task cannot be a window job */
                    const bool skip  = t.bw == 16 && t.rows <= 16 && t.skip_line;
                    const int  erows = skip ? (t.sa_h >> 1) : t.sa_h;
                    const bool modeB = t.rows * ((((t.bw + 3) >> 2) + 3) >> 2) >= 32;
                    rows_per         = max(1, (modeB ? HM_ITEM_POS_B : HM_ITEM_POS_A) / max(1, (int)t.sa_w));
                    n                = (erows + rows_per - 1) / rows_per;
                }
                if (nj > HM_MAXJOBS) { /* a task that overflows the job table on its own goes to the per-warp path */
                    pl.new_path = 0;
                    nj          = 0;
                }
                if (pl.new_path)
                    n = 0;
                sm.plan[lane] = pl;
                sm.tkey[lane] = none;
            } else if (lane < ME_NREF) {
                const FsTask f = sm.f[lane];
                if (f.valid) {
                    if (f.w <= FS_WARP_MAXW && (int)f.w * (int)f.h <= FS_WARP_MAX) {
                        atomicAdd(reinterpret_cast<unsigned long long *>(&sm.s.ops_fs_inline), (unsigned long long)f.w * f.h * (sub ? 64 * 32 : 64 * 64));
                        rows_per = max(1, HM_ITEM_POS_FS / max(1, (int)f.w));
                        n        = (f.h + rows_per - 1) / rows_per;
                    } else
                        fs_emit_tiles(f, list, (uint32_t)b * ME_NREF + (uint32_t)lane, (uint16_t)blockIdx.y);
                }
            }
            int incl = n;
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o)
                    incl += v;
            }
            sm.item_first[lane + 1] = (uint16_t)incl;
            sm.item_rows[lane]      = (uint16_t)rows_per;
            if (lane == 0)
                sm.item_first[0] = sm.job_first[0] = 0;
        }
#ifdef ME_HME_TRACE
        const long long tr_pa = clock64();
#endif
        __syncthreads();
#ifdef ME_HME_TRACE
        const long long tr_pb = clock64();
#endif
        /* window jobs, in rounds of up to HM_MAXJOBS jobs (whole tasks; the 4K presets' HME level 0 cuts a b64's 28
           quadrant searches into ~280 jobs) */
        for (int task_lo = 0; search_stage && task_lo < ME_SLOTS;) {
#ifdef ME_HME_TRACE
            tr_mark = clock64();
#endif
            if (warp == 0) {
                const HmPlan pl = sm.plan[lane];
                const int    nj = lane >= task_lo && pl.new_path ? (int)pl.nxb * (int)pl.nyb : 0;
                int          inclj = nj;
                for (int o = 1; o < 32; o <<= 1) {
                    const int vj = __shfl_up_sync(0xffffffffu, inclj, o);
                    if (lane >= o)
                        inclj += vj;
                }
                const bool     over = inclj > HM_MAXJOBS; /* job table full: this task and the ones after it wait for the next round */
                const unsigned ovm  = __ballot_sync(0xffffffffu, over);
                const int      kept = __reduce_max_sync(0xffffffffu, over ? 0 : inclj);
                /* warps per job in this round: the largest team any of its window jobs needs */
                const int team = __reduce_max_sync(0xffffffffu, nj > 0 && !over ? (int)pl.team : 1);
                sm.job_first[lane + 1] = (uint16_t)(over ? kept : inclj);
                if (lane == 0) {
                    sm.job_first[0] = 0;
                    sm.team         = team;
                    sm.next_lo      = ovm ? __ffs(ovm) - 1 : ME_SLOTS;
                }
            }
            __syncthreads();
            const int njobs = sm.job_first[ME_SLOTS];
            task_lo         = sm.next_lo;
#ifdef ME_HME_TRACE
            tr_round += clock64() - tr_mark;
            tr_mark = clock64();
#endif
            if (njobs == 0)
                break;
            for (int jb = tid; jb < njobs; jb += HM_THREADS) hm_job_decode(sm.t, sm.plan, sm.job_first, j, jb, sm.jobs[jb]);
            __syncthreads();
#ifdef ME_HME_TRACE
            tr_decode += clock64() - tr_mark;
            tr_mark = clock64();
#endif
            fetched = 0;
            if (producer) {
                if (lane == 0) { /* ring placement, in job order (FIFO): depends on the sizes only, never on timing */
                    { /* job 0 sits at the start of the ring whatever follows: its window is requested before the others are placed */
                        HmJobD &J0 = sm.jobs[0];
                        J0.off128  = 0;
                        J0.dep     = -1;
                        const uint32_t bar = hm_smem(&sm.mbar[seq % HM_NBAR]);
                        const int      pitch = hm_map_width(J0.map), br = hm_map_rows(J0.map);
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); /* generic-proxy accesses to the ring come first */
                        hm_mbar_expect(bar, (uint32_t)((int)J0.nrows * pitch));
                        for (int y = 0; y < (int)J0.nrows; y += br)
                            hm_tma_box(hm_smem(raw0) + (uint32_t)(y * pitch), maps.m[J0.map], J0.xt, J0.y0 + y, sm.t[J0.task].ref_slot, bar);
                        fetched = 1;
                    }
                    int cur = 0;
                    if (sm.team < HM_WARPS) {
                        /* the jobs of a batch are searched together: batches alternate between the ring halves when they
                           fit one, else every batch starts at 0 (then the next batch is fetched after this one) */
                        const int per = HM_WARPS / sm.team;
                        for (int jb0 = 0; jb0 < njobs; jb0 += per) {
                            int bytes = 0;
                            for (int jb = jb0; jb < min(njobs, jb0 + per); jb++)
                                bytes += (int)sm.jobs[jb].nrows * hm_map_width(sm.jobs[jb].map);
                            cur = (bytes <= HM_RING_BYTES / 2 && ((jb0 / per) & 1)) ? HM_RING_BYTES / 2 : 0;
                            for (int jb = jb0; jb < min(njobs, jb0 + per); jb++) {
                                sm.jobs[jb].off128 = (uint16_t)(cur >> 7);
                                cur += (int)sm.jobs[jb].nrows * hm_map_width(sm.jobs[jb].map);
                            }
                        }
                    } else
                        for (int jb = 0; jb < njobs; jb++) {
                            const int size = (int)sm.jobs[jb].nrows * hm_map_width(sm.jobs[jb].map);
                            if (cur + size > HM_RING_BYTES)
                                cur = 0;
                            sm.jobs[jb].off128 = (uint16_t)(cur >> 7);
                            cur += size;
                        }
                }
                __syncwarp();
                for (int n = lane; n < njobs; n += 32) { /* ring dependencies (at most HM_NBAR jobs are in flight) */
                    const int off = (int)sm.jobs[n].off128 * 128, size = (int)sm.jobs[n].nrows * hm_map_width(sm.jobs[n].map);
                    int       dep = -1;
                    for (int k = max(0, n - (HM_NBAR - 1)); k < n; k++) {
                        const int offk = (int)sm.jobs[k].off128 * 128, sizek = (int)sm.jobs[k].nrows * hm_map_width(sm.jobs[k].map);
                        if (off < offk + sizek && offk < off + size)
                            dep = k;
                    }
                    sm.jobs[n].dep = (int8_t)dep;
                }
                __syncwarp();
            }
            /* searched while the TMA transfers of the following jobs are in flight: large jobs by the whole CTA one after
               the other, small ones by one warp each, four at a time */
            const uint32_t seq0 = seq;
            const int      team = sm.team, per_iter = HM_WARPS / team;
            for (int jb0 = 0; jb0 < njobs; jb0 += per_iter) {
                if (producer) {
                    if (lane == 0) {
                        if (jb0 == 0) /* generic-proxy accesses to the ring (per-warp path of earlier stages) come first */
                            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        hm_fetch_ahead(maps, sm.t, sm.jobs, sm.mbar, hm_smem(raw0), njobs, seq0, jb0, fetched);
                    }
                    __syncthreads();
                    continue;
                }
                const int jb = jb0 + warp / team;
                if (jb < njobs) {
                    const uint32_t    slot = (seq0 + (uint32_t)jb) % HM_NBAR, par = ((seq0 + (uint32_t)jb) / HM_NBAR) & 1;
                    __syncwarp();
#ifdef ME_HME_TRACE
                    const long long tw0 = clock64();
#endif
                    hm_mbar_wait(hm_smem(&sm.mbar[slot]), par); /* also orders the producer's ring placement before the read */
#ifdef ME_HME_TRACE
                    const long long tw1 = clock64();
                    if (jb0 == 0)
                        tr_first += tw1 - tr_mark;
#endif
                    const HmJobD      J    = sm.jobs[jb];
                    const SearchTask &t    = sm.t[J.task];
                    const HmPlan     &pl   = sm.plan[J.task];
                    const uint32_t   *raw  = reinterpret_cast<const uint32_t *>(raw0 + (size_t)J.off128 * 128);
                    const uint32_t   *srcb = t.level == 2 ? sm.src2 : t.level == 1 ? sm.src1 : sm.src0;
                    const int sp4 = (t.level == 2 ? HM_SRC2_PITCH : t.level == 1 ? HM_SRC1_PITCH : HM_SRC0_PITCH) / 4;
                    const int widx = team == HM_WARPS ? (warp + jb) & (HM_WARPS - 1) : warp % team, nthr = 32 * team;
                    unsigned long long key;
                    if (pl.kx == 8)
                        key = hm_job_search8(J, t, raw, srcb, sp4, widx, nthr, lane);
                    else if (pl.kx == 1)
                        key = hm_job_search_tall(J, t, raw, srcb, sp4, widx, nthr, lane);
                    else if (pl.kx == 4)
                        key = hm_job_search<4>(J, t, pl.g_log2, raw, srcb, sp4, widx, nthr, lane);
                    else
                        key = hm_job_search<2>(J, t, pl.g_log2, raw, srcb, sp4, widx, nthr, lane);
                    if (lane == 0)
                        atomicMin(&sm.tkey[J.task], key);
#ifdef ME_HME_TRACE
                    tr_wait += tw1 - tw0;
                    tr_search += clock64() - tw1;
                    tr_jobs++;
#endif
                }
#ifdef ME_HME_TRACE
                const long long tb0 = clock64();
#endif
                __syncthreads(); /* the windows' ring bytes are free again */
#ifdef ME_HME_TRACE
                tr_bar += clock64() - tb0;
#endif
            }
            seq += (uint32_t)njobs;
        }
#ifdef ME_HME_TRACE
        long long tr2 = clock64();
#endif
        const int total = sm.item_first[ME_SLOTS];
        for (; !producer;) {
            int it = 0;
            if (lane == 0)
                it = atomicAdd(&sm.next_item, 1);
            it = __shfl_sync(0xffffffffu, it, 0);
            if (it >= total)
                break;
            const unsigned hit  = __ballot_sync(0xffffffffu, sm.item_first[lane] <= it && it < sm.item_first[lane + 1]);
            const int      task = __ffs(hit) - 1;
            const int      rows_per = sm.item_rows[task];
            const int      lo = (it - sm.item_first[task]) * rows_per;
            if (search_stage) {
                const SearchTask &t = sm.t[task];
                const bool skip  = t.bw == 16 && t.rows <= 16 && t.skip_line;
                const int  erows = skip ? (t.sa_h >> 1) : t.sa_h;
                const unsigned long long key =
                    search_item(t, lo, min(erows, lo + rows_per), j, pool, sm.buf[warp], &list->overflow, lane);
                if (lane == 0)
                    atomicMin(&sm.tkey[task], key);
            } else {
                const FsTask &f = sm.f[task];
                if (sub)
                    fullsearch_item<true>(f, lo, min((int)f.h, lo + rows_per), j, pool, sm.buf[warp], keys + task * ME_NPU,
                                          c.org_x, c.org_y, lane, sm.src0);
                else
                    fullsearch_item<false>(f, lo, min((int)f.h, lo + rows_per), j, pool, sm.buf[warp], keys + task * ME_NPU,
                                           c.org_x, c.org_y, lane, sm.src0);
            }
        }
        if (!search_stage)
            __threadfence(); /* the probe keys are read (at L2) by thread 0 in the next stage */
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); /* warp buffers written here are TMA targets later */
#ifdef ME_HME_TRACE
        long long tr3w = clock64();
#endif
        __syncthreads();
#ifdef ME_HME_TRACE
        long long tr3 = clock64();
        tr[st][0] = tr1 - tr0; tr[st][1] = tr2 - tr1; tr[st][2] = tr3 - tr2; tr[st][3] = tr3w - tr2;
        tr[st][4] = tr_wait; tr[st][5] = tr_search; tr[st][6] = tr_bar; tr[st][7] = tr_jobs;
        tr[st][8] = tr_round; tr[st][9] = tr_decode; tr[st][10] = tr_first; tr[st][11] = tr_pa - tr1; tr[st][3] = tr_pb - tr_pa;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_stage[st]));
#endif
        if (search_stage && tid < ME_SLOTS && sm.t[tid].valid) {
            /* key -> what the callers of svt_sad_loop_kernel see */
            const SearchTask &t   = sm.t[tid];
            const unsigned long long key = sm.tkey[tid];
            SearchResult      r;
            r.sad   = (uint32_t)(key >> 32);
            r.found = r.sad < ME_SAD_LOOP_INIT;
            r.x = r.y = 0;
            r.pad_[0] = r.pad_[1] = r.pad_[2] = 0;
            if (r.found) {
                const bool     skip = t.bw == 16 && t.rows <= 16 && t.skip_line;
                const uint32_t idx = (uint32_t)key, ee = idx / (uint32_t)t.sa_w;
                r.x                = (int16_t)(idx - ee * (uint32_t)t.sa_w);
                r.y                = (int16_t)(skip ? 2 * ee + 1 : ee);
            }
            sm.r[tid] = r;
        }
        __syncthreads();
    }
#ifdef ME_HME_TRACE
    if (blockIdx.y == 0 && b == 100 && tid == 0) {
        for (int st = 0; st < ST_COUNT; st++)
            printf("b %d tid %d stage %d: ctrl %lld plan+jobs %lld items+barrier %lld (own items %lld) cycles; warp 0's window jobs: %lld "
                   "of them, TMA wait %lld, SAD loop %lld, barrier %lld; round planning %lld, decode %lld, decode -> first window there %lld; task planning %lld (+ barrier %lld)\n", b, tid, st, tr[st][0], tr[st][1], tr[st][2], tr[st][3],
                   tr[st][7], tr[st][4], tr[st][5], tr[st][6], tr[st][8], tr[st][9], tr[st][10], tr[st][11], tr[st][3]);
        printf("b %d tid %d total %lld cycles\n", b, tid, clock64() - t_begin);
    }
    if (tid == 0) { /* print only once every CTA of the launch is past its last stage (bounded wait): printf stalls the SM */
        const unsigned old = atomicAdd(&g_trace_cnt, 1u), target = (old / gridDim.x + 1u) * gridDim.x;
        for (int spin = 0; spin < 4000000 && *(volatile unsigned *)&g_trace_cnt < target; spin++) __nanosleep(100);
    }
    if (tid == 0)
        printf("CTA %d sm %u t0 %llu st %llu %llu %llu %llu %llu %llu %llu %llu %llu\n", b, smid, gt_begin, gt_stage[0] - gt_begin,
               gt_stage[1] - gt_begin, gt_stage[2] - gt_begin, gt_stage[3] - gt_begin, gt_stage[4] - gt_begin, gt_stage[5] - gt_begin,
               gt_stage[6] - gt_begin, gt_stage[7] - gt_begin, gt_stage[8] - gt_begin);
#endif
    /* the epilogue (after the tiled searches) continues from the state in HBM */
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(&sm.s);
        uint32_t       *dst = reinterpret_cast<uint32_t *>(&j->state[b]);
        for (int i = tid; i < (int)(sizeof(B64State) / 4); i += HM_THREADS) dst[i] = src[i];
        if (phase == 1) /* ... and so does phase 2 */
            for (int i = tid; i < ME_SLOTS; i += HM_THREADS) {
                j->tasks[(size_t)b * ME_SLOTS + i]   = sm.t[i];
                j->results[(size_t)b * ME_SLOTS + i] = sm.r[i];
            }
    }
}

static unsigned long long g_hme_launches = 0;
unsigned long long me_hme_launch_count(void) { return g_hme_launches; }

#ifndef HM_EXTRA_SMEM /* experiments only (tools/exp/mkvariant.sh): unused shared memory that lowers the CTAs per SM */
#define HM_EXTRA_SMEM 0
#endif
cudaError_t me_hme_setup(void) {
    return cudaFuncSetAttribute(k_hme, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(HmeShared) + HM_EXTRA_SMEM);
}

/* tensor maps of the pool's planes for pictures of geometry g (TMA: rank 3 = byte column, row, pool slot) */
int me_hme_make_maps(HmeMaps *maps, const uint8_t *pool, const MePicGeom *g, int nslots) {
    typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static_assert(sizeof(CUtensorMap) == 128, "HmeMaps holds CUtensorMap images");
    EncodeFn                        encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &qres) != cudaSuccess || !encode ||
        qres != cudaDriverEntryPointSuccess)
        return -1;
    for (int m = 0; m < HM_NMAPS; m++) {
        const MePlaneGeom &pl = g->plane[hm_map_level(m)];
        cuuint64_t dims[3]    = {(cuuint64_t)pl.stride, (cuuint64_t)pl.rows, (cuuint64_t)nslots};
        cuuint64_t strides[2] = {(cuuint64_t)pl.stride, (cuuint64_t)g->slot_bytes};
        cuuint32_t box[3] = {(cuuint32_t)hm_map_width(m), (cuuint32_t)hm_map_rows(m), 1}, estr[3] = {1, 1, 1};
        CUresult   r = encode(reinterpret_cast<CUtensorMap *>(maps->m[m]), CU_TENSOR_MAP_DATA_TYPE_UINT8, 3,
                              const_cast<uint8_t *>(pool) + pl.offset, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS)
            return (int)r;
    }
    return 0;
}

void me_launch_hme(MeJob *jobs, int njobs, int max_b64, int phase, FsTileList *list, const uint8_t *pool, const HmeMaps *maps,
                   cudaStream_t st) {
    dim3 grid(max_b64, njobs);
    k_hme<<<grid, HM_ALL_THREADS, sizeof(HmeShared) + HM_EXTRA_SMEM, st>>>(jobs, list, pool, *maps, phase);
    g_hme_launches++;
}
