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
#include <cuda_runtime.h>
#include <stdint.h>

#include "me_ctrl.h"
#include "me_kernels.h"

#define HM_WARPS 4
#define HM_THREADS (HM_WARPS * 32)
#define HM_BUF_WORDS 2112 /* per-warp staging buffer: source block (<= 512 words) + reference window (15 rows x
                             131 words of the widest pre-HME strip fit; taller windows are banded) */
#define FS_WARP_MAX 256 /* search areas up to this many positions (and <= FS_WARP_MAXW wide) stay in k_hme */
#define FS_WARP_MAXW 48
#define HM_ITEM_POS_A 384 /* search positions per work item, lane = position */
#define HM_ITEM_POS_B 32 /* ... lane = block row */
#define HM_ITEM_POS_FS 16 /* ... full-search positions */

struct HmeShared {
    B64State           s;
    SearchTask         t[ME_SLOTS];
    SearchResult       r[ME_SLOTS];
    FsTask             f[ME_NREF];
    unsigned long long tkey[ME_SLOTS]; /* best (sad << 32 | scan index) of each search task */
    int                item_first[ME_SLOTS + 1]; /* prefix sums: items of task i are [first[i], first[i+1]) */
    int                item_rows[ME_SLOTS]; /* search rows per item */
    int                next_item;
    alignas(16) uint32_t buf[HM_WARPS][HM_BUF_WORDS + 8]; /* +8: masked over-reads of the last window row stay inside */
};

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
                uint32_t        sad = 0;
#pragma unroll 4
                for (int r = 0; r < rows; r++) {
                    const uint4    s4 = sp[r];
                    const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3], w4 = wp[4];
                    sad = hm_vsad4(s4.x, __funnelshift_r(w0, w1, sh), sad);
                    sad = hm_vsad4(s4.y, __funnelshift_r(w1, w2, sh), sad);
                    sad = hm_vsad4(s4.z, __funnelshift_r(w2, w3, sh), sad);
                    sad = hm_vsad4(s4.w, __funnelshift_r(w3, w4, sh), sad);
                    wp += rstep;
                }
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
                uint32_t        part = 0;
                if (!masked) {
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        if (k < npass) {
                            const uint32_t *q  = wp + off[k];
                            const uint32_t  w0 = q[0], w1 = q[1], w2 = q[2], w3 = q[3], w4 = q[4];
                            part = hm_vsad4(sw[k][0], __funnelshift_r(w0, w1, sh), part);
                            part = hm_vsad4(sw[k][1], __funnelshift_r(w1, w2, sh), part);
                            part = hm_vsad4(sw[k][2], __funnelshift_r(w2, w3, sh), part);
                            part = hm_vsad4(sw[k][3], __funnelshift_r(w3, w4, sh), part);
                        }
                } else {
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        if (k < npass) {
                            const uint32_t *q  = wp + off[k];
                            const uint32_t  w0 = q[0], w1 = q[1], w2 = q[2], w3 = q[3], w4 = q[4];
                            part = hm_vsad4(sw[k][0], __funnelshift_r(w0, w1, sh) & mk[k][0], part);
                            part = hm_vsad4(sw[k][1], __funnelshift_r(w1, w2, sh) & mk[k][1], part);
                            part = hm_vsad4(sw[k][2], __funnelshift_r(w2, w3, sh) & mk[k][2], part);
                            part = hm_vsad4(sw[k][3], __funnelshift_r(w3, w4, sh) & mk[k][3], part);
                        }
                }
                const uint32_t sad = __reduce_add_sync(0xffffffffu, part);
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
 * open_loop_me_fullpel_search_sblock over the search rows [y_lo, y_hi) of a small area (or the single
 * probe position), one warp. Lane z owns the 8x8 blocks with z-order index z and z + 32; the 16x16 /
 * 32x32 / 64x64 sums are butterflies over lanes (z-order: 4 consecutive lanes = one 16x16, 16 = one 32x32).
 * ---------------------------------------------------------------------------------------------- */
template <bool SUB>
__device__ void fullsearch_item(const FsTask &f, int y_lo, int y_hi, const MeJob *j, const uint8_t *pool, uint32_t *win,
                                unsigned long long *keys, int org_x, int org_y, int lane) {
    constexpr int      NR = SUB ? 4 : 8, RSTEP = SUB ? 2 : 1;
    const MePlaneGeom &pl = j->g.plane[0];
    const int          stride_words = pl.stride >> 2;
    /* z-order -> block coordinates */
    const int bx = (lane & 1) | ((lane >> 1) & 2) | ((lane >> 2) & 4);
    const int by = ((lane >> 1) & 1) | ((lane >> 2) & 2); /* + 4 for the second block */
    uint32_t  sw[2][NR][2];
    {
        const uint8_t *sp = hm_plane(pool, j->g, j->src_slot, 0) + (size_t)(pl.pad + org_y) * pl.stride + pl.pad + org_x;
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int r = 0; r < NR; r++) {
                const uint32_t *q = (const uint32_t *)(sp + (size_t)(8 * (by + 4 * b) + r * RSTEP) * pl.stride + 8 * bx);
                sw[b][r][0]       = __ldg(q);
                sw[b][r][1]       = __ldg(q + 1);
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
                    uint32_t        acc = 0;
#pragma unroll
                    for (int r = 0; r < NR; r++) {
                        const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2];
                        acc = hm_vsad4(sw[b][r][0], __funnelshift_r(w0, w1, sh), acc);
                        acc = hm_vsad4(sw[b][r][1], __funnelshift_r(w1, w2, sh), acc);
                        wp += pitch * RSTEP;
                    }
                    a8[b] = acc;
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

/* append the 64x32-position tiles of a large search area to the work list of k_fullsearch */
__device__ void hm_emit_tiles(const FsTask &f, FsTileList *list, uint32_t task, uint16_t job) {
    const uint32_t ntx = (uint32_t)(f.w + FS_TW - 1) / FS_TW, nty = (uint32_t)(f.h + FS_TH - 1) / FS_TH;
    const uint32_t n = ntx * nty, base = atomicAdd(&list->count, n);
    if (base + n > list->cap) {
        atomicExch(&list->overflow, 1u);
        return;
    }
    FsTile *t = list->tiles + base;
    for (uint32_t ty = 0; ty < nty; ty++)
        for (uint32_t tx = 0; tx < ntx; tx++, t++) {
            t->task = task;
            t->job  = job;
            t->tx   = (uint16_t)(tx * FS_TW);
            t->ty   = (uint16_t)(ty * FS_TH);
            t->ref_slot = f.ref_slot;
            t->x0 = f.x0; t->y0 = f.y0; t->w = f.w; t->h = f.h;
            t->probe = f.probe;
        }
}

/* ------------------------------------------------------------------------------------------------
 * k_hme
 * ---------------------------------------------------------------------------------------------- */
__global__ void __launch_bounds__(HM_THREADS, 4) k_hme(MeJob *jobs, FsTileList *list, const uint8_t *pool) {
    extern __shared__ __align__(16) unsigned char hm_raw[];
    HmeShared     &sm  = *reinterpret_cast<HmeShared *>(hm_raw);
    MeJob         *j   = &jobs[blockIdx.y];
    const int      b   = blockIdx.x;
    const int      tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long *keys = &j->keys[(size_t)b * ME_NREF * ME_NPU];
    const unsigned long long none = ((unsigned long long)ME_SAD_LOOP_INIT << 32) | 0xFFFFFFFFull;

    CtrlCtx c;
    c.j     = j;
    c.p     = &j->p;
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
    const bool sub = j->p.me_search_method == ME_GPU_SUB_SAD_SEARCH;

    mc_init_keys(keys, tid, HM_THREADS);
    for (int i = tid; i < ME_SLOTS; i += HM_THREADS) {
        sm.t[i].valid = 0;
        sm.r[i].sad   = 0;
        sm.r[i].found = 0;
        sm.r[i].x = sm.r[i].y = 0;
    }
    __threadfence();
    __syncthreads();

    for (int st = 0; st < ST_COUNT; st++) {
        if (tid == 0) {
            mc_run_stage(&c, st);
            sm.next_item = 0;
        }
        __syncthreads();
        const bool search_stage = st <= ST_CENTRE;
        /* warp 0: cut the tasks into work items (bands of search rows) */
        if (warp == 0) {
            int n = 0, rows_per = 1;
            if (search_stage) {
                const SearchTask t = sm.t[lane];
                if (t.valid) {
                    const bool skip  = t.bw == 16 && t.rows <= 16 && t.skip_line;
                    const int  erows = skip ? (t.sa_h >> 1) : t.sa_h;
                    const bool modeB = t.rows * ((((t.bw + 3) >> 2) + 3) >> 2) >= 32;
                    rows_per         = max(1, (modeB ? HM_ITEM_POS_B : HM_ITEM_POS_A) / max(1, (int)t.sa_w));
                    n                = (erows + rows_per - 1) / rows_per;
                }
                sm.tkey[lane] = none;
            } else if (lane < ME_NREF) {
                const FsTask f = sm.f[lane];
                if (f.valid) {
                    if (f.w <= FS_WARP_MAXW && (int)f.w * (int)f.h <= FS_WARP_MAX) {
                        rows_per = max(1, HM_ITEM_POS_FS / max(1, (int)f.w));
                        n        = (f.h + rows_per - 1) / rows_per;
                    } else
                        hm_emit_tiles(f, list, (uint32_t)b * ME_NREF + (uint32_t)lane, (uint16_t)blockIdx.y);
                }
            }
            int incl = n;
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o)
                    incl += v;
            }
            sm.item_first[lane + 1] = incl;
            sm.item_rows[lane]      = rows_per;
            if (lane == 0)
                sm.item_first[0] = 0;
        }
        __syncthreads();
        const int total = sm.item_first[ME_SLOTS];
        for (;;) {
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
                                          c.org_x, c.org_y, lane);
                else
                    fullsearch_item<false>(f, lo, min((int)f.h, lo + rows_per), j, pool, sm.buf[warp], keys + task * ME_NPU,
                                           c.org_x, c.org_y, lane);
            }
        }
        if (!search_stage)
            __threadfence(); /* the probe keys are read (at L2) by thread 0 in the next stage */
        __syncthreads();
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
    /* the epilogue (after the tiled searches) continues from the state in HBM */
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(&sm.s);
        uint32_t       *dst = reinterpret_cast<uint32_t *>(&j->state[b]);
        for (int i = tid; i < (int)(sizeof(B64State) / 4); i += HM_THREADS) dst[i] = src[i];
    }
}

static unsigned long long g_hme_launches = 0;
unsigned long long me_hme_launch_count(void) { return g_hme_launches; }

cudaError_t me_hme_setup(void) {
    return cudaFuncSetAttribute(k_hme, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(HmeShared));
}

void me_launch_hme(MeJob *jobs, int njobs, int max_b64, FsTileList *list, const uint8_t *pool, cudaStream_t st) {
    dim3 grid(max_b64, njobs);
    k_hme<<<grid, HM_THREADS, sizeof(HmeShared), st>>>(jobs, list, pool);
    g_hme_launches++;
}
