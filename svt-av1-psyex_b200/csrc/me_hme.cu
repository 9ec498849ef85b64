/*
 * me_hme.cu - k_hme: the whole per-b64 chain of open-loop ME up to (and including) the small integer
 * searches, in ONE launch per picture.
 *
 * One CTA per 64x64 block, ONE WARP PER SEARCHED REFERENCE. The reference's per-b64 algorithm
 * (motion_estimation.c:3076-3153) is a chain of searches per (list, reference) - zero-MV SAD, pre-HME, HME level
 * 0 / 1 / 2, check_00_center, 8x8-variance probe, integer search - in which every search window depends on the
 * previous search of the SAME reference, and only four steps look across references (zero-MV pruning, pre-HME
 * pruning, final centre + HME pruning, and the epilogue). So a warp owns the chain of its reference from end to
 * end: it runs the control code of me_ctrl.h for its reference (lane 0), fetches its own search windows by TMA
 * into its own double-buffered slice of shared memory, searches them with all 32 lanes, and meets the other
 * warps of the block only at the (at most four) CTA barriers where the reference's code looks across references.
 * No warp ever waits for another reference's search, for a planning step, or for a producer warp.
 *
 * The block's MeContext state (B64State), the search tasks and results and the block's source pixels on the three
 * planes live in shared memory for the whole chain; nothing goes through HBM between stages.
 *
 * Search executors (all SADs are VABSDIFF4.U8.ACC on 32-bit words; the window arrives 16-byte aligned from TMA and
 * unaligned reference words are built with funnel shifts):
 *   hm_job_search8      wide areas of 16-byte block rows (pre-HME region 1, HME level 0): lane = 8 positions of one
 *                       32-byte window segment
 *   hm_job_search_tall  8-wide, hundreds of rows (pre-HME region 0): lane = 2 positions x 4 search rows
 *   hm_job_search<KX>   everything else (HME level 1 / 2, zero-MV and centre SADs): unit = KX positions, spread
 *                       over column segments x row groups of lanes
 *   search_item         generic fallback (picture-edge blocks of odd widths): window staged with LDG
 *   fullsearch_item     open_loop_me_fullpel_search_sblock for the probe and for search areas of up to
 *                       FS_WARP_MAX positions: lane = two 8x8 blocks (z-order), the 85 square-PU SADs by butterflies
 * Larger integer-search areas are cut into tiles and appended to the work list of k_fullsearch.
 */
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "me_ctrl.h"
#include "me_kernels.h"

#define HM_MAXW ME_NREF /* warps per CTA: one per searched (list, reference) */
#ifndef FS_WARP_MAX
#define FS_WARP_MAX 256
#endif
/* FS_WARP_MAX: search areas up to this many positions (and <= FS_WARP_MAXW wide) stay in k_hme */
#ifndef FS_WARP_MAXW
#define FS_WARP_MAXW 48
#endif

/* source-block pitches in shared memory (words): padded so that lanes on different block rows hit different banks */
#define HM_SRC0_PITCH 24
#define HM_SRC1_PITCH 12
#define HM_SRC2_PITCH 4

struct HmPlan { /* how a task is cut into window jobs */
    int16_t pwb, neb; /* positions per x band, evaluated rows per y band */
    uint8_t nxb, nyb, map, kx;
    uint8_t g_log2, new_path, pad_[2];
};

struct HmJobD { /* one window job, 32 bytes */
    int16_t  px0, pw, e0, ne; /* band of search positions: columns [px0, px0 + pw), evaluated rows [e0, e0 + ne) */
    int16_t  xt, y0; /* TMA column (multiple of 16) and row of the window */
    uint16_t pitch; /* bytes per window row = box width of the job's tensor map */
    uint16_t nrows; /* window rows (whole boxes) */
    uint16_t total; /* lanes' worth of work of the SAD loop: units x lanes per unit */
    uint8_t  task, map, aoff /* byte offset of position px0 in a window row */, ncls, nchunk;
    uint8_t  brows; /* rows per TMA box */
    uint32_t m_chunk, m_cls;
};
#define HM_ROUND 16 /* window jobs decoded at a time (one lane each) */
#define HM_TASKS 4 /* search tasks of one reference in one stage: 4 quadrants / 2 regions / 2 centre SADs */

/* planner thresholds (measured in round 1): areas at least this wide / with at least this many 8-position units take
 * hm_job_search8; 8-wide areas of at least this many rows take hm_job_search_tall */
#define HM_K8_MIN_PW 32
#define HM_K8_MIN_UNITS 64
#define HM_TALL_MIN_ROWS 32

struct HmWarp { /* private to one warp */
    alignas(8) unsigned long long mbar[2]; /* TMA completion of the two windows in flight */
    HmJobD   jobs[HM_ROUND];
    HmPlan   plan[HM_TASKS];
    uint16_t job_first[HM_TASKS + 1]; /* prefix sums of the window jobs over the stage's tasks */
    uint16_t pad_[3];
};

struct HmeShared {
    B64State           s;
    SearchTask         t[ME_SLOTS];
    SearchResult       r[ME_SLOTS];
    FsTask             f[ME_NREF];
    unsigned long long tkey[ME_SLOTS]; /* best (sad << 32 | scan index) of each search task */
    MeGpuParams        p; /* the picture's controls: the control code reads them hundreds of times */
    alignas(16) uint32_t src0[64 * HM_SRC0_PITCH], src1[32 * HM_SRC1_PITCH], src2[16 * HM_SRC2_PITCH]; /* the b64's source block on the three planes (padded pitches) */
    /* HmWarp[nwarps] and then the warps' window rings follow (ring_bytes each, 128-byte aligned) */
};
#define HM_FIXED_BYTES ((sizeof(HmeShared) + 127) & ~(size_t)127)
#define HM_WARP_BYTES(nw) ((sizeof(HmWarp) * (size_t)(nw) + 127) & ~(size_t)127)
/* registers per thread. Measured (1080p preset 4, frames/s): 128 -> 4530, 96 -> 4260, 80 -> 4250, 64 -> 3990: the spills cost more than the extra resident CTAs bring */
#ifndef HM_MAXREG
#define HM_MAXREG 128
#endif

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
__device__ unsigned long long search_item(const SearchTask &t, int e_lo, int e_hi, const MeJob *j, const uint8_t *pool,
                                          uint32_t *buf, int buf_words, unsigned int *overflow, int lane) {
    const MePlaneGeom &pl   = j->g.plane[t.level];
    const uint8_t     *srcp = hm_plane(pool, j->g, j->src_slot, t.level);
    const uint8_t     *refp = hm_plane(pool, j->g, t.ref_slot, t.level);
    const int          bw = t.bw, rows = t.rows, step = t.row_step;
    const int          nw  = (bw + 3) >> 2; /* words per block row */
    const int          nwp = (nw + 3) & ~3; /* padded to whole 16-byte chunks: 4, 8, 12 or 16 */
    const int          stride_words = pl.stride >> 2;
    const int          src_words = rows * nwp;
    uint32_t          *srcw = buf, *win = buf + src_words;
    const int          win_cap = buf_words - src_words;
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

/* box widths of the tensor maps, by plane: sixteenth {48, 80, 224}, quarter {80}, full {96, 128} */
__host__ __device__ __forceinline__ int hm_map_width(int m) {
    return m == 0 ? 48 : m == 1 ? 80 : m == 2 ? 224 : m == 3 ? 80 : m == 4 ? 96 : 128;
}
__host__ __device__ __forceinline__ int hm_map_level(int m) { return m < 3 ? 2 : m == 3 ? 1 : 0; }
/* rows per box: the usual window of the stage that uses the map in ONE box (the TMA unit takes a box every ~35 cycles
 * per SM whatever its size: tools/tma_bench.cu); taller windows stack boxes */
__host__ __device__ __forceinline__ int hm_map_rows(int m) {
    return m == 0 ? 64 : m == 1 ? 56 : m == 2 ? 24 : m == 3 ? 40 : 72;
}

/* cut one task into window jobs of at most `cap` bytes each; false when it cannot be done (picture-edge blocks of odd
 * widths, windows taller than the budget): the task then takes the generic path */
__device__ bool hm_plan_try(const SearchTask &t, const MeJob *j, int org_x, int org_y, int cap, HmPlan &pl) {
    pl.new_path = 0;
    pl.nxb = pl.nyb = 0;
    const int lv = t.level, full = 64 >> lv;
    if (t.bw != full || t.sx != (org_x >> lv) || t.sy != (org_y >> lv) || (int)t.rows * t.row_step > full || t.sa_w < 1)
        return false;
    const MePlaneGeom &g    = j->g.plane[lv];
    const bool         skip = t.bw == 16 && t.rows <= 16 && t.skip_line;
    const int          ymul = skip ? 2 : 1, erows = skip ? (t.sa_h >> 1) : t.sa_h;
    if (erows < 1)
        return false;
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
    const int rows_cap = (cap / pitch) / hm_map_rows(map) * hm_map_rows(map);
    const int ne_max   = (rows_cap - span - 1) / ymul + 1;
    if (rows_cap < span + 1 || ne_max < 1 || nxb > 255)
        return false;
    int       kx      = (t.bw == 16 && pwb >= 16) ? 4 : 2;
    const int ncls    = min(4, pwb), per_cls = (pwb + 3) >> 2, nchunk = (per_cls + kx - 1) / kx;
    const int nyb = (erows + ne_max - 1) / ne_max, neb = (erows + nyb - 1) / nyb;
    if (nyb > 255)
        return false;
    int       units   = ncls * nchunk * neb;
    /* wide areas of 16-byte block rows: units of 8 positions on 32-byte window segments (hm_job_search8), one lane each */
    const bool wide = t.bw == 16 && pwb >= HM_K8_MIN_PW && 4 * ((pwb + 31) >> 5) * neb >= HM_K8_MIN_UNITS;
    if (wide)
        kx = 8;
    /* tall thin areas (pre-HME region 0: 8 positions wide, hundreds of rows) of 16x8-row blocks: hm_job_search_tall */
    const bool tall = !wide && t.bw == 16 && t.rows == 8 && nxb == 1 && pwb <= 8 && neb >= HM_TALL_MIN_ROWS && t.row_step % ymul == 0;
    if (tall)
        kx = 1;
    int g_log2 = 2 - lv; /* lanes per unit: the 16-byte column segments of the block row ... */
    if (tall)
        g_log2 = 0;
    else if (!wide) {
        /* ... x row groups: the split with the fewest (passes over the warp) x (block rows per lane); ties keep the
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
    }
    pl.pwb = (int16_t)pwb;
    pl.neb = (int16_t)neb;
    pl.nxb = (uint8_t)nxb;
    pl.nyb = (uint8_t)nyb;
    pl.map = (uint8_t)map;
    pl.kx  = (uint8_t)kx;
    pl.g_log2   = (uint8_t)g_log2;
    pl.new_path = 1;
    return true;
}

/* A warp's ring holds two windows so that the next one arrives while this one is searched: jobs are planned for half
 * the ring; a task whose whole search area fits the ring in ONE window takes that instead (HME level 2 and the centre
 * SADs: 64-pixel blocks, a window of 96 x 72 bytes) */
__device__ void hm_plan_task(const SearchTask &t, const MeJob *j, int org_x, int org_y, int ring_bytes, HmPlan &pl) {
    HmPlan whole;
    const bool ok_half  = hm_plan_try(t, j, org_x, org_y, (ring_bytes / 2) & ~127, pl);
    if (ok_half && pl.nyb == 1)
        return;
    const bool ok_whole = hm_plan_try(t, j, org_x, org_y, ring_bytes, whole);
    if (ok_whole && (whole.nyb == 1 || !ok_half))
        pl = whole;
    else if (!ok_half)
        pl.new_path = 0;
}

/* job number (over the stage's tasks of this warp) -> job descriptor (one lane per job) */
__device__ void hm_job_decode(const SearchTask *tasks, const HmPlan *plans, const uint16_t *job_first, int ntasks, const MeJob *j, int jb,
                              HmJobD &J) {
    int task = 0;
    for (int i = 1; i < ntasks; i++) /* last task with job_first[task] <= jb */
        if (job_first[i] <= jb)
            task = i;
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
    J.pitch  = (uint16_t)hm_map_width(pl.map);
    J.brows  = (uint8_t)br;
    J.ncls   = (uint8_t)ncls;
    J.nchunk = (uint8_t)nchunk;
    J.total  = (uint16_t)((ncls * nchunk * ne) << pl.g_log2);
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
/* the SAD loop of one job over the CTA; returns this warp's best key (same value in every lane).
 * A unit (KX positions) is spread over G = 2^g_log2 lanes: the 16-byte column segments of the block row (bw / 16 of
 * them, adjacent lanes -> adjacent banks) times row groups; partial sums are joined by shuffles. */
template <int KX>
__device__ __forceinline__ unsigned long long hm_job_search(const HmJobD &J, const SearchTask &t, int g_log2, const uint32_t *raw,
                                                            const uint32_t *src, int src_pitch4, int warp, int nthr, int lane) {
    const bool skip = t.bw == 16 && t.rows <= 16 && t.skip_line;
    const int  ymul = skip ? 2 : 1, rows = t.rows, step = t.row_step;
    const int  pitchw = (int)J.pitch >> 2, G = 1 << g_log2;
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
#pragma unroll 4
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
    const int  pitch4 = (int)J.pitch >> 4; /* window row pitch in 16-byte units */
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
    const int  pitchw = (int)J.pitch >> 2, total = J.total, ne = J.ne, maxrow = (int)J.nrows - 1;
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
__device__ void fullsearch_item(const FsTask &f, int y_lo, int y_hi, const MeJob *j, const uint8_t *pool, uint32_t *win, int win_words,
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
    const int      band  = win_words / pitch - span; /* >= 5 rows of positions (the ring is at least 9 KB) */
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
 * one warp, one stage: the (up to HM_TASKS) search tasks sm.t[first .. first + n) of this warp's reference
 * ---------------------------------------------------------------------------------------------- */
#ifdef ME_HME_TRACE
/* per-phase cycle sums over all warps of all launches (lane 0's clock64), read back by me_hme_trace_read(): experiments only */
enum { TR_SETUP, TR_ZZ, TR_PHME, TR_CROSS, TR_L0, TR_L1, TR_L2, TR_CENTRE, TR_PROBE, TR_MAIN, TR_END, TR_PLAN, TR_DECODE, TR_WAIT, TR_SEARCH,
       TR_GENERIC, TR_CTRL, TR_JOBS, TR_WARPS, TR_TOTAL, TR_N };
__device__ unsigned long long g_trace[TR_N];
#define TR_ADD(i, v) do { if ((threadIdx.x & 31) == 0) atomicAdd(&g_trace[i], (unsigned long long)(v)); } while (0)
#define TR_T() clock64()
#else
#define TR_ADD(i, v) ((void)0)
#define TR_T() 0ll
#endif

struct HmWarpCtx {
    HmWarp  *w;
    uint8_t *ring; /* this warp's window ring */
    int      ring_bytes;
    uint32_t seq; /* window jobs of this warp so far: job q lands at the ring's low (even q) or high (odd q) end and completes on
                     mbarrier q & 1, phase (q >> 1) & 1 */
    int      lane;
};

__device__ __forceinline__ int hm_job_bytes(const HmJobD &J) { return (int)J.nrows * (int)J.pitch; }
__device__ __forceinline__ int hm_job_offset(const HmJobD &J, uint32_t q, int ring_bytes) {
    return (q & 1) ? (ring_bytes - hm_job_bytes(J)) & ~127 : 0;
}

/* lane 0: request the window of job J (sequence number q) */
__device__ __forceinline__ void hm_issue(const HmWarpCtx &wc, const HmeMaps &maps, const HmJobD &J, uint32_t q, uint32_t ref_slot) {
    const uint32_t bar   = hm_smem(&wc.w->mbar[q & 1]);
    const int      pitch = J.pitch, br = J.brows, off = hm_job_offset(J, q, wc.ring_bytes);
    const uint32_t dst   = hm_smem(wc.ring) + (uint32_t)off;
    hm_mbar_expect(bar, (uint32_t)hm_job_bytes(J));
    for (int y = 0; y < (int)J.nrows; y += br) hm_tma_box(dst + (uint32_t)(y * pitch), maps.m[J.map], J.xt, J.y0 + y, (int)ref_slot, bar);
}

__device__ __noinline__ void hm_run_tasks(HmeShared &sm, HmWarpCtx &wc, int first, int n, const MeJob *j, const uint8_t *pool, const HmeMaps &maps,
                             int org_x, int org_y, FsTileList *list) {
    const int                lane = wc.lane;
    const unsigned long long none = ((unsigned long long)ME_SAD_LOOP_INIT << 32) | 0xFFFFFFFFull;
    HmWarp                  &W    = *wc.w;
    /* plan: lane i cuts task i into window jobs */
    long long tr0 = TR_T();
    int  nj = 0;
    bool generic = false;
    if (lane < n) {
        const SearchTask t = sm.t[first + lane];
        HmPlan           pl;
        pl.new_path = 0;
        pl.nxb = pl.nyb = 0;
        if (t.valid) {
            hm_plan_task(t, j, org_x, org_y, wc.ring_bytes, pl);
            if (pl.new_path)
                nj = (int)pl.nxb * (int)pl.nyb;
            else
                generic = true;
        }
        W.plan[lane]         = pl;
        sm.tkey[first + lane] = none;
    }
    int incl = nj;
#pragma unroll
    for (int o = 1; o < HM_TASKS; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o)
            incl += v;
    }
    if (lane < HM_TASKS)
        W.job_first[lane + 1] = (uint16_t)incl;
    if (lane == 0)
        W.job_first[0] = 0;
    const int      total = __shfl_sync(0xffffffffu, incl, HM_TASKS - 1);
    const unsigned gmask = __ballot_sync(0xffffffffu, generic);
    __syncwarp();
    TR_ADD(TR_PLAN, TR_T() - tr0);
    TR_ADD(TR_JOBS, total);
    for (int base = 0; base < total; base += HM_ROUND) {
        const int nr = min(HM_ROUND, total - base);
        tr0 = TR_T();
        if (lane < nr)
            hm_job_decode(sm.t + first, W.plan, W.job_first, n, j, base + lane, W.jobs[lane]);
        __syncwarp();
        TR_ADD(TR_DECODE, TR_T() - tr0);
        if (lane == 0)
            hm_issue(wc, maps, W.jobs[0], wc.seq, sm.t[first + W.jobs[0].task].ref_slot);
        for (int i = 0; i < nr; i++) {
            const uint32_t q = wc.seq + (uint32_t)i;
            const HmJobD   J = W.jobs[i];
            /* the next window can come in while this one is searched when the two do not overlap in the ring */
            bool ahead = false;
            if (i + 1 < nr) {
                const HmJobD &Jn   = W.jobs[i + 1];
                const int     off  = hm_job_offset(J, q, wc.ring_bytes), offn = hm_job_offset(Jn, q + 1, wc.ring_bytes);
                ahead              = (q & 1) ? hm_job_bytes(Jn) <= off : hm_job_bytes(J) <= offn;
                if (ahead && lane == 0)
                    hm_issue(wc, maps, Jn, q + 1, sm.t[first + Jn.task].ref_slot);
            }
            tr0 = TR_T();
            hm_mbar_wait(hm_smem(&W.mbar[q & 1]), (q >> 1) & 1);
            TR_ADD(TR_WAIT, TR_T() - tr0);
            tr0 = TR_T();
            const SearchTask &t    = sm.t[first + J.task];
            const HmPlan     &pl   = W.plan[J.task];
            const uint32_t   *raw  = reinterpret_cast<const uint32_t *>(wc.ring + hm_job_offset(J, q, wc.ring_bytes));
            const uint32_t   *srcb = t.level == 2 ? sm.src2 : t.level == 1 ? sm.src1 : sm.src0;
            const int sp4 = (t.level == 2 ? HM_SRC2_PITCH : t.level == 1 ? HM_SRC1_PITCH : HM_SRC0_PITCH) / 4;
            unsigned long long key;
            if (pl.kx == 8)
                key = hm_job_search8(J, t, raw, srcb, sp4, 0, 32, lane);
            else if (pl.kx == 1)
                key = hm_job_search_tall(J, t, raw, srcb, sp4, 0, 32, lane);
            else if (pl.kx == 4)
                key = hm_job_search<4>(J, t, pl.g_log2, raw, srcb, sp4, 0, 32, lane);
            else
                key = hm_job_search<2>(J, t, pl.g_log2, raw, srcb, sp4, 0, 32, lane);
            if (lane == 0 && key < sm.tkey[first + J.task]) /* the task's keys are this warp's alone */
                sm.tkey[first + J.task] = key;
            __syncwarp(); /* every lane is done with the window before its bytes are requested again */
            TR_ADD(TR_SEARCH, TR_T() - tr0);
            if (i + 1 < nr && !ahead && lane == 0)
                hm_issue(wc, maps, W.jobs[i + 1], q + 1, sm.t[first + W.jobs[i + 1].task].ref_slot);
        }
        wc.seq += (uint32_t)nr;
        __syncwarp();
    }
    /* tasks that cannot be window jobs: the generic path, staged with plain loads into the ring */
    tr0 = TR_T();
    for (unsigned m = gmask; m; m &= m - 1) {
        const int         i = __ffs(m) - 1;
        const SearchTask &t = sm.t[first + i];
        const bool skip  = t.bw == 16 && t.rows <= 16 && t.skip_line;
        const int  erows = skip ? (t.sa_h >> 1) : t.sa_h;
        const unsigned long long key = search_item(t, 0, erows, j, pool, reinterpret_cast<uint32_t *>(wc.ring), wc.ring_bytes / 4 - 8,
                                                   &list->overflow, lane);
        if (lane == 0)
            sm.tkey[first + i] = key;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); /* ring bytes written here are TMA targets later */
        __syncwarp();
    }
    __syncwarp();
    TR_ADD(TR_GENERIC, TR_T() - tr0);
    if (lane < n && sm.t[first + lane].valid) {
        /* key -> what the callers of svt_sad_loop_kernel see */
        const SearchTask        &t   = sm.t[first + lane];
        const unsigned long long key = sm.tkey[first + lane];
        SearchResult             r;
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
        sm.r[first + lane] = r;
    }
    __syncwarp();
}

/* the integer search task sm.f[k] of this warp's reference: searched here when small, else cut into tiles for k_fullsearch */
__device__ __noinline__ void hm_run_fs(HmeShared &sm, HmWarpCtx &wc, int k, const MeJob *j, const uint8_t *pool, unsigned long long *keys, int org_x,
                          int org_y, FsTileList *list, uint32_t b, bool sub) {
    const FsTask f = sm.f[k];
    if (!f.valid)
        return;
    if (f.w <= FS_WARP_MAXW && (int)f.w * (int)f.h <= FS_WARP_MAX) {
        if (wc.lane == 0)
            atomicAdd(reinterpret_cast<unsigned long long *>(&sm.s.ops_fs_inline), (unsigned long long)f.w * f.h * (sub ? 64 * 32 : 64 * 64));
        if (sub)
            fullsearch_item<true>(f, 0, f.h, j, pool, reinterpret_cast<uint32_t *>(wc.ring), wc.ring_bytes / 4, keys + k * ME_NPU, org_x, org_y,
                                  wc.lane, sm.src0);
        else
            fullsearch_item<false>(f, 0, f.h, j, pool, reinterpret_cast<uint32_t *>(wc.ring), wc.ring_bytes / 4, keys + k * ME_NPU, org_x, org_y,
                                   wc.lane, sm.src0);
        __threadfence(); /* the probe keys are read back (at L2) by this warp's lane 0 */
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); /* ring bytes written here are TMA targets later */
    } else if (wc.lane == 0)
        fs_emit_tiles(f, list, b * ME_NREF + (uint32_t)k, (uint16_t)blockIdx.y);
    __syncwarp();
}

/* ------------------------------------------------------------------------------------------------
 * k_hme
 * ---------------------------------------------------------------------------------------------- */
#define HM_LANE0(stmt)                    \
    do {                                  \
        const long long tc_ = TR_T();     \
        if (lane == 0) {                  \
            stmt;                         \
        }                                 \
        __syncwarp();                     \
        TR_ADD(TR_CTRL, TR_T() - tc_);    \
    } while (0)
/* a step of the reference's code that looks across references: one thread, between two CTA barriers */
#define HM_CROSS(stmt)                                            \
    do {                                                          \
        const long long tx_ = TR_T();                             \
        __syncthreads();                                          \
        if (threadIdx.x == 0) {                                   \
            c.only = -1;                                          \
            stmt;                                                 \
            c.only = k;                                           \
            for (int i_ = 0; i_ < ME_SLOTS; i_++) sm.t[i_].valid = 0; /* the next stage numbers the task slots its own way */ \
        }                                                         \
        __syncthreads();                                          \
        TR_ADD(TR_CROSS, TR_T() - tx_);                           \
    } while (0)

__device__ __forceinline__ void hm_clear_slots(SearchTask *t, int first, int n) {
    for (int i = 0; i < n; i++) t[first + i].valid = 0;
}

__global__ void __maxnreg__(HM_MAXREG)
    k_hme(MeJob *jobs, FsTileList *list, const uint8_t *pool, const __grid_constant__ HmeMaps maps, int ring_bytes) {
    extern __shared__ __align__(1024) unsigned char hm_raw[];
    long long      tp  = TR_T();
    const long long t_begin = tp;
#define TR_PHASE(i) do { const long long tn_ = TR_T(); TR_ADD(i, tn_ - tp); tp = tn_; } while (0)
    HmeShared     &sm  = *reinterpret_cast<HmeShared *>(hm_raw);
    MeJob         *j   = &jobs[blockIdx.y];
    const uint32_t b   = blockIdx.x;
    const int      lane = threadIdx.x & 31, warp = threadIdx.x >> 5, tid = (int)threadIdx.x, nthr = (int)blockDim.x;
    unsigned long long *keys = &j->keys[(size_t)b * ME_NREF * ME_NPU];

    /* warp -> (list, reference): the searched references in the reference's loop order */
    const int n0 = j->p.num_of_ref_pic_to_search[0];
    const int li = warp < n0 ? 0 : 1, ri = warp < n0 ? warp : warp - n0;
    const int k  = li * ME_NR + ri;

    CtrlCtx c;
    c.j     = j;
    c.p     = &sm.p;
    c.s     = &sm.s;
    c.t     = sm.t;
    c.r     = sm.r;
    c.f     = sm.f;
    c.keys  = keys;
    c.org_x = (int)(b % (uint32_t)j->g.b64_cols) * 64;
    c.org_y = (int)(b / (uint32_t)j->g.b64_cols) * 64;
    c.b64_w = (uint32_t)min(64, j->g.width - c.org_x);
    c.b64_h = (uint32_t)min(64, j->g.height - c.org_y);
    c.pic_w = j->g.width;
    c.pic_h = j->g.height;
    c.only   = k; /* this warp's lane 0 runs the control code of its own reference */
    c.lane   = 0;
    c.nlanes = 1;

    HmWarpCtx wc;
    wc.w          = reinterpret_cast<HmWarp *>(hm_raw + HM_FIXED_BYTES) + warp;
    wc.ring       = hm_raw + HM_FIXED_BYTES + HM_WARP_BYTES(blockDim.x >> 5) + (size_t)warp * (size_t)ring_bytes;
    wc.ring_bytes = ring_bytes;
    wc.seq        = 0;
    wc.lane       = lane;

    for (int i = tid; i < (int)(sizeof(MeGpuParams) / 4); i += nthr)
        reinterpret_cast<uint32_t *>(&sm.p)[i] = reinterpret_cast<const uint32_t *>(&j->p)[i];
    mc_init_keys(keys, tid, nthr);
    for (int i = tid; i < ME_SLOTS; i += nthr) {
        sm.t[i].valid = 0;
        sm.r[i].sad   = 0;
        sm.r[i].found = 0;
        sm.r[i].x = sm.r[i].y = 0;
    }
    if (lane == 0) {
        hm_mbar_init(hm_smem(&wc.w->mbar[0]), 1);
        hm_mbar_init(hm_smem(&wc.w->mbar[1]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    /* the b64's source block on the three planes stays in shared memory for the whole chain (rows below a short
       picture-edge block are padding: loaded, never summed) */
    for (int base = 0; base < 64 * 16 + 32 * 8 + 16 * 4; base += 4 * nthr) { /* four loads in flight per thread: the set-up is HBM latency */
        uint32_t  v[4];
        uint32_t *dst[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i = base + u * nthr + tid;
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
    const MeGpuParams &p   = sm.p;
    const bool         sub = p.me_search_method == ME_GPU_SUB_SAD_SEARCH;
    const bool         hme = p.enable_hme_flag != 0;
#define HM_RUN(first, n) hm_run_tasks(sm, wc, (first), (n), j, pool, maps, c.org_x, c.org_y, list)

    /* init_me_hme_data: all 2 x 4 (list, reference) slots, searched or not (the pruning steps loop over all of them) */
    if (tid < ME_NREF) {
        CtrlCtx ci = c;
        ci.only    = tid;
        mc_stage_init(&ci);
    }
    __syncthreads();
    TR_PHASE(TR_SETUP);

    /* zero-MV SADs (ME:2382-2437); the pruning that follows looks across references */
    if (p.me_early_exit_th || p.me_safe_limit_zz_th) {
        HM_LANE0(mc_zz_emit(&c));
        HM_RUN(k, 1);
        HM_CROSS(mc_zz_consume(&c));
    }
    TR_PHASE(TR_ZZ);
    /* pre-HME (ME:1722-1796): with l1_early_exit list 1 reads list 0's results, so the lists go one after the other */
    if (p.prehme_enable) {
        const bool two = p.prehme_l1_early_exit && p.num_of_list_to_search > 1;
        for (int pass = 0; pass < (two ? 2 : 1); pass++) {
            if (!two || li == pass) {
                HM_LANE0(hm_clear_slots(sm.t, 2 * k, 2); mc_prehme_emit_list(&c, li));
                HM_RUN(2 * k, 2);
                HM_LANE0(mc_prehme_consume_tasks(&c));
            }
            if (two && pass == 0)
                __syncthreads();
        }
        HM_CROSS(mc_prehme_finish(&c));
    }
    TR_PHASE(TR_PHME);
    /* HME level 0 / 1 / 2 (ME:1906-2177): a chain per reference */
    if (hme && p.enable_hme_level0_flag) {
        HM_LANE0(hm_clear_slots(sm.t, 4 * k, 4); mc_l0_emit(&c));
        HM_RUN(4 * k, 4);
        HM_LANE0(mc_l0_consume(&c));
    }
    TR_PHASE(TR_L0);
    if (hme && p.enable_hme_level1_flag) {
        HM_LANE0(hm_clear_slots(sm.t, 4 * k, 4); mc_l1_emit(&c));
        HM_RUN(4 * k, 4);
        HM_LANE0(mc_refine_consume(&c, 1));
    }
    TR_PHASE(TR_L1);
    if (hme && p.enable_hme_level2_flag) {
        HM_LANE0(hm_clear_slots(sm.t, 4 * k, 4); mc_l2_emit(&c));
        HM_RUN(4 * k, 4);
        HM_LANE0(mc_refine_consume(&c, 2));
    }
    TR_PHASE(TR_L2);
    /* final centre + HME-based pruning (ME:2182-2380, 2477-2518): across references */
    HM_CROSS(mc_final_centre(&c); if (p.me_type == ME_GPU_ME_MCTF) sm.s.tf_exit = sm.s.hme_sad[0][0] < p.tf_me_exit_th; /* ME:3109-3113 */
             else if (hme) mc_hme_prune(&c) /* prune_ref = enable_hme_flag && me_type != ME_MCTF, ME:3103 */);
    /* integer search (ME:1249-1516): check_00_center SADs, 8x8-variance probe, main search; per reference */
    tp = TR_T();
    if (!sm.s.tf_exit) {
        HM_LANE0(mc_centre_emit(&c));
        HM_RUN(2 * k, 2);
        TR_PHASE(TR_CENTRE);
        HM_LANE0(mc_probe_emit(&c));
        hm_run_fs(sm, wc, k, j, pool, keys, c.org_x, c.org_y, list, b, sub);
        TR_PHASE(TR_PROBE);
        if (sm.s.probed[li][ri]) {
            HM_LANE0(mc_main_emit_after_probe(&c));
            hm_run_fs(sm, wc, k, j, pool, keys, c.org_x, c.org_y, list, b, sub);
        }
    }
#undef HM_RUN
    TR_PHASE(TR_MAIN);
    __syncthreads();
    TR_PHASE(TR_END);
    TR_ADD(TR_WARPS, 1);
    TR_ADD(TR_TOTAL, TR_T() - t_begin);
    /* the epilogue (after the tiled searches) continues from the state in HBM */
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(&sm.s);
        uint32_t       *dst = reinterpret_cast<uint32_t *>(&j->state[b]);
        for (int i = tid; i < (int)(sizeof(B64State) / 4); i += nthr) dst[i] = src[i];
    }
}

#ifdef ME_HME_TRACE
extern "C" int me_hme_trace_read(unsigned long long *out, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_trace, sizeof(unsigned long long) * TR_N);
    if (reset) {
        unsigned long long z[TR_N] = {0};
        cudaMemcpyToSymbol(g_trace, z, sizeof(z));
    }
    return TR_N;
}
#endif
static unsigned long long g_hme_launches = 0;
unsigned long long me_hme_launch_count(void) { return g_hme_launches; }

cudaError_t me_hme_setup(void) { return cudaFuncSetAttribute(k_hme, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024); }

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

/* Shared memory of a CTA of `nwarps` search warps: the fixed part + a window ring per warp. The ring is what is left of
 * the SM's 227 KB when `target` CTAs are resident - as many as the registers allow (HM_MAXREG) - never above 32 KB (the
 * in-job keys of hm_job_search8 index positions with 16 bits) nor below the largest single window (128 x 72 bytes). */
int me_hme_ring_bytes(int nwarps) {
    int target = (65536 / HM_MAXREG) / (32 * nwarps);
    if (target > 8)
        target = 8;
    if (target < 1)
        target = 1;
    for (;; target--) {
        const int budget = (227 * 1024) / target - 1024 /* per-CTA reservation */ - (int)HM_FIXED_BYTES - (int)HM_WARP_BYTES(nwarps) - 128 /* over-read slack */;
        int       ring   = (budget / nwarps) & ~127;
        if (ring > 32768)
            ring = 32768;
        if (ring >= 7168 || target == 1) /* 96 x 72: the whole window of an HME level 2 / centre search */
            return ring < 9216 && target == 1 ? 9216 : ring;
    }
}

void me_launch_hme(MeJob *jobs, int njobs, int max_b64, int nwarps, FsTileList *list, const uint8_t *pool, const HmeMaps *maps,
                   cudaStream_t st) {
    dim3         grid(max_b64, njobs);
    const int    ring = me_hme_ring_bytes(nwarps);
    const size_t smem = HM_FIXED_BYTES + HM_WARP_BYTES(nwarps) + (size_t)nwarps * (size_t)ring + 128;
    k_hme<<<grid, 32 * nwarps, smem, st>>>(jobs, list, pool, *maps, ring);
    g_hme_launches++;
}
