/*
 * me_oracle_b64.c - per-64x64-block driver of the CPU restatement (TEST INFRASTRUCTURE ONLY, see
 * me_oracle.h). ME = Codec/motion_estimation.c of the reference.
 */
#include "me_oracle.h"

#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#define OR_MIN(a, b) ((a) < (b) ? (a) : (b))
#define OR_MAX(a, b) ((a) > (b) ? (a) : (b))
#define OR_ABS(a) ((a) < 0 ? -(a) : (a))
#define OR_MAX_SAD_VALUE (128 * 128 * 255) /* Codec/motion_estimation.h:85 */
#define OR_MAX_U32 0xFFFFFFFFu
#define NL ME_GPU_MAX_LISTS
#define NR ME_GPU_MAX_REFS
#define NPU ME_GPU_SQUARE_PU_COUNT

typedef struct PreHme { /* SearchInfo, Codec/me_context.h:329 */
    uint16_t sa_w, sa_h;
    int16_t  col, row;
    uint64_t sad;
    uint8_t  valid;
} PreHme;

typedef struct B64Ctx {
    const MeGpuParams   *p;
    const OraclePicture *src;
    const OraclePicture *ref[NL][NR];
    int32_t              org_x, org_y; /* b64 origin, full resolution */
    uint32_t             b64_w, b64_h;
    int32_t              pic_w, pic_h; /* aligned picture size */
    /* HME state: [list][ref][sr_w][sr_h] */
    int16_t  l0x[NL][NR][2][2], l0y[NL][NR][2][2];
    int16_t  l1x[NL][NR][2][2], l1y[NL][NR][2][2];
    int16_t  l2x[NL][NR][2][2], l2y[NL][NR][2][2];
    uint64_t l0sad[NL][NR][2][2], l1sad[NL][NR][2][2], l2sad[NL][NR][2][2];
    /* SearchResults, Codec/me_context.h:348 */
    int16_t  sc_x[NL][NR], sc_y[NL][NR];
    uint64_t hme_sad[NL][NR];
    uint8_t  do_ref[NL][NR];
    uint32_t sr_div[NL][NR];
    uint32_t zz_sad[NL][NR];
    PreHme   phme[NL][NR][2];
    uint8_t  performed_phme[NL][NR][2];
    uint32_t best_sad[NL][NR][NPU];
    uint32_t best_mv[NL][NR][NPU];
    int16_t  sa_final[NL][NR][4];
    MeGpuSearchAreaMinMax hme_l0_sa; /* mutable: distance-based resizing edits and restores it */
    uint32_t me_distortion[NPU];
    uint64_t ops; /* algorithmic byte-SAD count, SURVEY 8(d) */
} B64Ctx;

static const uint8_t *plane_at(const OraclePlane *pl, int x, int y) {
    return pl->buf + (ptrdiff_t)(pl->org_y + y) * pl->stride + pl->org_x + x;
}

/* ME:1239-1243 svt_aom_get_scaled_picture_distance */
static uint16_t scaled_distance(uint16_t dist) { return (uint16_t)(((dist * 5) / 8) + ((dist % 8) ? 1 : 0)); }

/* ME:1232-1234 distance between the picture and a reference */
static uint16_t ref_distance(const B64Ctx *c, int li, int ri) {
    int64_t d = (int64_t)c->p->picture_number - (int64_t)c->ref[li][ri]->picture_number;
    return (uint16_t)(int16_t)OR_ABS(d);
}

/* Search-window clamp shared by ME:848-880 (HME L0), :949-982 (L1), :1050-1078 (L2), :1585-1629 (pre-HME)
 * and :1445-1491 (integer search). The left/top clamp moves the origin but - as in the reference, which
 * recomputes the overshoot with the already-moved origin - leaves the size unchanged. */
static void clamp_axis(int org, int16_t *origin, int16_t *size, int pad, int dim) {
    if (org + *origin < -pad)
        *origin = (int16_t)(-pad - org);
    if (org + *origin > dim - 1)
        *origin = (int16_t)(*origin - ((org + *origin) - (dim - 1)));
    if (org + *origin + *size > dim)
        *size = (int16_t)OR_MAX(1, *size - ((org + *origin + *size) - dim));
}

/* One svt_sad_loop_kernel call as issued by the HME stages: `sub` halves the rows and doubles the
 * strides (ME:891-909). Counts algorithmic byte-SADs. */
static void search_window(B64Ctx *c, const uint8_t *src, int src_stride, const OraclePlane *rp, int x0, int y0,
                          uint32_t bw, uint32_t bh, int sub, uint8_t skip_line, int16_t sa_w, int16_t sa_h,
                          uint64_t *sad, int16_t *bx, int16_t *by) {
    const uint8_t *ref = rp->buf + (ptrdiff_t)y0 * rp->stride + x0; /* x0,y0 are absolute buffer coordinates */
    uint32_t       rows = sub ? bh >> 1 : bh;
    oracle_sad_loop(src, sub ? src_stride * 2 : src_stride, ref, sub ? rp->stride * 2 : rp->stride, rows, bw, sad, bx,
                    by, rp->stride, skip_line, sa_w, sa_h);
    uint64_t lines = sa_h;
    if (bw == 16 && rows <= 16 && skip_line)
        lines = sa_h / 2;
    c->ops += lines * (uint64_t)sa_w * bw * rows;
    if (sub)
        *sad *= 2;
}

/* ME:3010-3070 init_me_hme_data (+ canonical reset of p_sb_best_sad, which the reference leaves stale
 * for pruned references; stale values are never read for an output). */
static void init_b64(B64Ctx *c) {
    memset(c->l0x, 0, sizeof c->l0x); memset(c->l0y, 0, sizeof c->l0y);
    memset(c->l1x, 0, sizeof c->l1x); memset(c->l1y, 0, sizeof c->l1y);
    memset(c->l2x, 0, sizeof c->l2x); memset(c->l2y, 0, sizeof c->l2y);
    memset(c->l0sad, 0, sizeof c->l0sad); memset(c->l1sad, 0, sizeof c->l1sad); memset(c->l2sad, 0, sizeof c->l2sad);
    memset(c->best_mv, 0, sizeof c->best_mv);
    memset(c->phme, 0, sizeof c->phme);
    memset(c->performed_phme, 0, sizeof c->performed_phme);
    memset(c->sa_final, 0, sizeof c->sa_final);
    for (int li = 0; li < NL; li++)
        for (int ri = 0; ri < NR; ri++) {
            c->do_ref[li][ri]  = 1;
            c->hme_sad[li][ri] = OR_MAX_U32;
            c->sc_x[li][ri] = c->sc_y[li][ri] = 0;
            c->sr_div[li][ri]  = 1;
            c->zz_sad[li][ri]  = OR_MAX_U32;
            for (int k = 0; k < NPU; k++) c->best_sad[li][ri][k] = OR_MAX_SAD_VALUE;
        }
    c->hme_l0_sa = c->p->hme_l0_sa;
}

/* ME:1667-1689 get_zz_sad + ME:2382-2437 init_zz_sad */
static void stage_zz(B64Ctx *c) {
    const MeGpuParams *p    = c->p;
    const uint8_t     *src  = plane_at(&c->src->full, c->org_x, c->org_y);
    uint32_t           best = OR_MAX_U32;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (!(p->temporal_layer_index > 0 || li == 0))
                continue;
            const OraclePlane *rp = &c->ref[li][ri]->full;
            uint32_t zz = oracle_nxm_sad(src, c->src->full.stride << 1, plane_at(rp, c->org_x, c->org_y),
                                         rp->stride << 1, c->b64_h >> 1, c->b64_w)
                << 1;
            c->ops += (uint64_t)(c->b64_h >> 1) * c->b64_w;
            zz                = (zz * 64 * 64) / (c->b64_w * c->b64_h);
            c->zz_sad[li][ri] = zz;
            best              = OR_MIN(best, zz);
        }
    if (p->temporal_layer_index > 0 && best < p->zz_sad_th)
        for (int li = 0; li < p->num_of_list_to_search; li++)
            for (int ri = 1; ri < p->num_of_ref_pic_to_search[li]; ri++)
                if ((uint32_t)((c->zz_sad[li][ri] - best) * 100) > (uint32_t)(p->zz_sad_pct * best))
                    c->do_ref[li][ri] = 0;
    if (p->me_safe_limit_zz_th) {
        int limit = p->hierarchical_levels > 0 && p->num_of_list_to_search == 2 &&
            p->pcs_temporal_layer_index >= p->hierarchical_levels && p->similar_brightness_refs &&
            c->zz_sad[0][0] < p->me_safe_limit_zz_th && c->zz_sad[1][0] < p->me_safe_limit_zz_th;
        if (limit)
            for (int li = 0; li < p->num_of_list_to_search; li++)
                for (int ri = 1; ri < p->num_of_ref_pic_to_search[li]; ri++) c->do_ref[li][ri] = 0;
    }
}

/* ME:1693-1719 check_prehme_early_exit */
static int prehme_early_exit(B64Ctx *c, int li, int ri, int sr) {
    const MeGpuParams *p = c->p;
    PreHme            *d = &c->phme[li][ri][sr];
    if (p->me_early_exit_th && c->zz_sad[li][ri] < p->me_early_exit_th) {
        d->col = d->row = 0;
        d->sad   = 0;
        d->valid = 1;
        return 1;
    }
    if (p->prehme_l1_early_exit) {
        const PreHme *l0 = &c->phme[0][ri][sr];
        if (li == 1 && l0->valid && (l0->sad < 32 * 32 || (OR_ABS(l0->col) < 16 && OR_ABS(l0->row) < 16))) {
            d->col   = (int16_t)-l0->col;
            d->row   = (int16_t)-l0->row;
            d->sad   = l0->sad;
            d->valid = 1;
            return 1;
        }
    }
    return 0;
}

/* ME:1568-1666 prehme_core: one long thin search on the 1/16 plane centred on (0,0) */
static void prehme_core(B64Ctx *c, int li, int ri, int sr) {
    const OraclePlane *rp  = &c->ref[li][ri]->sixteenth;
    PreHme            *d   = &c->phme[li][ri][sr];
    int                ox  = (int16_t)c->org_x >> 2, oy = (int16_t)c->org_y >> 2;
    int16_t            w   = (int16_t)d->sa_w, h = (int16_t)d->sa_h;
    int16_t            x0  = (int16_t) - (w >> 1), y0 = (int16_t) - (h >> 1);
    clamp_axis(ox, &x0, &w, rp->org_x - 1, rp->width);
    clamp_axis(oy, &y0, &h, rp->org_y - 1, rp->height);
    search_window(c, plane_at(&c->src->sixteenth, ox, oy), c->src->sixteenth.stride, rp, rp->org_x + ox + x0,
                  rp->org_y + oy + y0, c->b64_w >> 2, c->b64_h >> 2, c->p->hme_search_method == ME_GPU_SUB_SAD_SEARCH,
                  c->p->prehme_skip_search_line, w, h, &d->sad, &d->col, &d->row);
    d->col   = (int16_t)((d->col + x0) * 4);
    d->row   = (int16_t)((d->row + y0) * 4);
    d->valid = 1;
}

/* ME:1722-1796 prehme_b64 */
static void stage_prehme(B64Ctx *c) {
    const MeGpuParams *p    = c->p;
    uint32_t           best = OR_MAX_U32;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (p->temporal_layer_index > 0 || li == 0) {
                uint32_t f = scaled_distance(ref_distance(c, li, ri));
                for (int sr = 0; sr < 2; sr++) {
                    if (prehme_early_exit(c, li, ri, sr))
                        continue;
                    PreHme *d = &c->phme[li][ri][sr];
                    if (!c->do_ref[li][ri]) {
                        d->col = d->row = 0;
                        d->sad = OR_MAX_U32;
                        continue;
                    }
                    const MeGpuSearchAreaMinMax *cfg = &p->prehme_sa_cfg[sr];
                    d->sa_w = (uint16_t)OR_MIN(cfg->sa_min.width * f, (uint32_t)cfg->sa_max.width);
                    d->sa_h = (uint16_t)OR_MIN(cfg->sa_min.height * f, (uint32_t)cfg->sa_max.height);
                    prehme_core(c, li, ri, sr);
                    c->performed_phme[li][ri][sr] = 1;
                }
                uint32_t m = (uint32_t)OR_MIN(c->phme[li][ri][0].sad, c->phme[li][ri][1].sad);
                best       = OR_MIN(best, m);
            } else {
                for (int sr = 0; sr < 2; sr++) {
                    c->phme[1][ri][sr].col = (int16_t)-c->phme[0][ri][sr].col;
                    c->phme[1][ri][sr].row = (int16_t)-c->phme[0][ri][sr].row;
                    c->phme[1][ri][sr].sad = c->phme[0][ri][sr].sad;
                }
            }
        }
    if (p->temporal_layer_index > 0 && best < p->phme_sad_th)
        for (int li = 0; li < p->num_of_list_to_search; li++)
            for (int ri = 1; ri < p->num_of_ref_pic_to_search[li]; ri++) {
                if (!c->do_ref[li][ri])
                    continue;
                uint32_t s = (uint32_t)OR_MIN(c->phme[li][ri][0].sad, c->phme[li][ri][1].sad);
                if ((uint32_t)((s - best) * 100) > (uint32_t)(p->phme_sad_pct * best))
                    c->do_ref[li][ri] = 0;
            }
}

/* ME:1800-1867 get_hme_l0_search_area (edits c->hme_l0_sa when distance-based resizing is on; the
 * caller restores it, ME:1998-2003) */
static void hme_l0_area(B64Ctx *c, int li, int ri, uint16_t dist, int16_t *sa_w, int16_t *sa_h) {
    const MeGpuParams *p = c->p;
    if (p->enable_me_sr_adjustment && p->distance_based_hme_resizing) {
        uint8_t is_hor = 1, is_ver = 1, is_still = 0;
        if (p->reduce_hme_l0_sr_th_min && p->reduce_hme_l0_sr_th_max && (li || ri)) {
            int mvx = c->l0x[0][0][0][0], mvy = c->l0y[0][0][0][0];
            is_ver   = OR_ABS(mvx) < p->reduce_hme_l0_sr_th_min && OR_ABS(mvy) > p->reduce_hme_l0_sr_th_max;
            is_hor   = OR_ABS(mvx) > p->reduce_hme_l0_sr_th_max && OR_ABS(mvy) < p->reduce_hme_l0_sr_th_min;
            is_still = OR_ABS(mvx) < p->reduce_hme_l0_sr_th_min * 3 && OR_ABS(mvy) < p->reduce_hme_l0_sr_th_min * 3;
        }
        uint8_t x_off = is_hor ? 1 : 2, y_off = is_ver ? 1 : 2;
        if (p->enable_me_sr_adjustment == 2 && is_still)
            x_off = y_off = 4;
        c->hme_l0_sa.sa_min.width  = (uint16_t)(c->hme_l0_sa.sa_min.width / (x_off + ri));
        c->hme_l0_sa.sa_min.height = (uint16_t)(c->hme_l0_sa.sa_min.height / (y_off + ri));
        c->hme_l0_sa.sa_max.width  = (uint16_t)(c->hme_l0_sa.sa_max.width / (x_off + ri));
        c->hme_l0_sa.sa_max.height = (uint16_t)(c->hme_l0_sa.sa_max.height / (y_off + ri));
    }
    int32_t f = scaled_distance(dist);
    int16_t w = (int16_t)(c->hme_l0_sa.sa_min.width / p->num_hme_sa_w);
    w         = (int16_t)OR_MIN(((w * f) + 15) & ~0x0F, ((c->hme_l0_sa.sa_max.width / p->num_hme_sa_w) + 15) & ~0x0F);
    int16_t h = (int16_t)(c->hme_l0_sa.sa_min.height / p->num_hme_sa_h);
    h         = (int16_t)OR_MIN(h * f, c->hme_l0_sa.sa_max.height / p->num_hme_sa_h);
    *sa_w     = w;
    *sa_h     = h;
}

/* ME:820-920 hme_level_0: one quadrant of the 2x2 HME-L0 grid on the 1/16 plane */
static void hme_level_0(B64Ctx *c, int li, int ri, int16_t sa_w, int16_t sa_h, int sr_w, int sr_h) {
    const MeGpuParams *p  = c->p;
    const OraclePlane *rp = &c->ref[li][ri]->sixteenth;
    int                ox = (int16_t)c->org_x >> 2, oy = (int16_t)c->org_y >> 2;
    sa_w                  = (int16_t)((sa_w + 7) & ~0x07);
    int16_t x0            = (int16_t)(-(int16_t)((sa_w * p->num_hme_sa_w) >> 1) + (int16_t)(sa_w * sr_w));
    int16_t y0            = (int16_t)(-(int16_t)((sa_h * p->num_hme_sa_h) >> 1) + (int16_t)(sa_h * sr_h));
    clamp_axis(ox, &x0, &sa_w, rp->org_x - 1, rp->width);
    sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
    clamp_axis(oy, &y0, &sa_h, rp->org_y - 1, rp->height);
    search_window(c, plane_at(&c->src->sixteenth, ox, oy), c->src->sixteenth.stride, rp, rp->org_x + ox + x0,
                  rp->org_y + oy + y0, c->b64_w >> 2, c->b64_h >> 2, p->hme_search_method == ME_GPU_SUB_SAD_SEARCH, 0, sa_w,
                  sa_h, &c->l0sad[li][ri][sr_w][sr_h], &c->l0x[li][ri][sr_w][sr_h], &c->l0y[li][ri][sr_w][sr_h]);
    c->l0x[li][ri][sr_w][sr_h] = (int16_t)((c->l0x[li][ri][sr_w][sr_h] + x0) * 4);
    c->l0y[li][ri][sr_w][sr_h] = (int16_t)((c->l0y[li][ri][sr_w][sr_h] + y0) * 4);
}

/* ME:1872-1901 get_worst_quadrant: the last comparison does not update max_sad (kept as written) */
static void worst_quadrant(const B64Ctx *c, int li, int ri, uint8_t *w, uint8_t *h) {
    uint64_t m = 0;
    if (c->l0sad[li][ri][0][0] > m) { m = c->l0sad[li][ri][0][0]; *w = 0; *h = 0; }
    if (c->l0sad[li][ri][1][0] > m) { m = c->l0sad[li][ri][1][0]; *w = 1; *h = 0; }
    if (c->l0sad[li][ri][0][1] > m) { m = c->l0sad[li][ri][0][1]; *w = 0; *h = 1; }
    if (c->l0sad[li][ri][1][1] > m) { *w = 1; *h = 1; }
}

static void fill_quadrants(int16_t x[2][2], int16_t y[2][2], uint64_t s[2][2], int16_t vx, int16_t vy, uint64_t vs) {
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++) { x[a][b] = vx; y[a][b] = vy; s[a][b] = vs; }
}

/* ME:1906-2036 hme_level0_b64 */
static void stage_hme_l0(B64Ctx *c) {
    const MeGpuParams *p = c->p;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (p->me_early_exit_th && c->zz_sad[li][ri] < (p->me_early_exit_th >> 2)) {
                fill_quadrants(c->l0x[li][ri], c->l0y[li][ri], c->l0sad[li][ri], 0, 0, 0);
                continue;
            }
            if (p->prev_me_stage_based_exit_th) {
                int sr = c->phme[li][ri][0].sad <= c->phme[li][ri][1].sad ? 0 : 1;
                if (c->performed_phme[li][ri][sr] &&
                    c->phme[li][ri][sr].sad < (p->prev_me_stage_based_exit_th >> 4)) {
                    fill_quadrants(c->l0x[li][ri], c->l0y[li][ri], c->l0sad[li][ri], c->phme[li][ri][sr].col,
                                   c->phme[li][ri][sr].row, c->phme[li][ri][sr].sad);
                    continue;
                }
            }
            if (!c->do_ref[li][ri]) {
                fill_quadrants(c->l0x[li][ri], c->l0y[li][ri], c->l0sad[li][ri], 0, 0, OR_MAX_U32);
                continue;
            }
            if (!(p->temporal_layer_index > 0 || li == 0))
                continue;
            MeGpuSearchAreaMinMax base = c->hme_l0_sa;
            int16_t               sa_w = 0, sa_h = 0;
            hme_l0_area(c, li, ri, ref_distance(c, li, ri), &sa_w, &sa_h);
            for (int sr_h = 0; sr_h < p->num_hme_sa_h; sr_h++)
                for (int sr_w = 0; sr_w < p->num_hme_sa_w; sr_w++) hme_level_0(c, li, ri, sa_w, sa_h, sr_w, sr_h);
            if (p->enable_me_sr_adjustment && p->distance_based_hme_resizing)
                c->hme_l0_sa = base;
            if (p->prehme_enable) {
                uint8_t wq = 0, hq = 0;
                worst_quadrant(c, li, ri, &wq, &hq);
                int sr = c->phme[li][ri][0].sad <= c->phme[li][ri][1].sad ? 0 : 1;
                if (c->phme[li][ri][sr].sad < c->l0sad[li][ri][wq][hq]) {
                    c->l0sad[li][ri][wq][hq] = c->phme[li][ri][sr].sad;
                    c->l0x[li][ri][wq][hq]   = c->phme[li][ri][sr].col;
                    c->l0y[li][ri][wq][hq]   = c->phme[li][ri][sr].row;
                }
            }
        }
}

/* ME:923-1022 hme_level_1 / ME:1025-1113 hme_level_2: refinement around the previous level's centre.
 * level 1: 1/4 plane, block b64/2, result x2. level 2: full plane, pad 63, result x1. */
static void hme_refine(B64Ctx *c, int level, int li, int ri, int16_t sa_w, int16_t sa_h, int16_t cx, int16_t cy,
                       uint64_t *sad, int16_t *outx, int16_t *outy) {
    const OraclePlane *rp   = level == 1 ? &c->ref[li][ri]->quarter : &c->ref[li][ri]->full;
    const OraclePlane *sp   = level == 1 ? &c->src->quarter : &c->src->full;
    int                sh   = level == 1 ? 1 : 0;
    int                ox   = (int16_t)c->org_x >> sh, oy = (int16_t)c->org_y >> sh;
    int                padx = level == 1 ? rp->org_x - 1 : 63, pady = level == 1 ? rp->org_y - 1 : 63;
    sa_w                    = (int16_t)((sa_w + 7) & ~0x07);
    int16_t x0 = (int16_t)(-(sa_w >> 1) + cx), y0 = (int16_t)(-(sa_h >> 1) + cy);
    clamp_axis(ox, &x0, &sa_w, padx, rp->width);
    sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
    clamp_axis(oy, &y0, &sa_h, pady, rp->height);
    search_window(c, plane_at(sp, ox, oy), sp->stride, rp, rp->org_x + ox + x0, rp->org_y + oy + y0, c->b64_w >> sh,
                  c->b64_h >> sh, c->p->hme_search_method == ME_GPU_SUB_SAD_SEARCH, 0, sa_w, sa_h, sad, outx, outy);
    *outx = (int16_t)((*outx + x0) * (level == 1 ? 2 : 1));
    *outy = (int16_t)((*outy + y0) * (level == 1 ? 2 : 1));
}

/* ME:2041-2122 hme_level1_b64 */
static void stage_hme_l1(B64Ctx *c) {
    const MeGpuParams *p = c->p;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (!(p->temporal_layer_index > 0 || li == 0))
                continue;
            if (p->me_early_exit_th && c->zz_sad[li][ri] < (p->me_early_exit_th >> 2)) {
                fill_quadrants(c->l1x[li][ri], c->l1y[li][ri], c->l1sad[li][ri], 0, 0, 0);
                continue;
            }
            if (!c->do_ref[li][ri]) {
                fill_quadrants(c->l1x[li][ri], c->l1y[li][ri], c->l1sad[li][ri], 0, 0, OR_MAX_U32);
                continue;
            }
            for (int sr_h = 0; sr_h < p->num_hme_sa_h; sr_h++)
                for (int sr_w = 0; sr_w < p->num_hme_sa_w; sr_w++) {
                    if (p->prev_me_stage_based_exit_th &&
                        c->l0sad[li][ri][sr_w][sr_h] < (p->prev_me_stage_based_exit_th >> 5)) {
                        c->l1x[li][ri][sr_w][sr_h]   = c->l0x[li][ri][sr_w][sr_h];
                        c->l1y[li][ri][sr_w][sr_h]   = c->l0y[li][ri][sr_w][sr_h];
                        c->l1sad[li][ri][sr_w][sr_h] = c->l0sad[li][ri][sr_w][sr_h];
                        continue;
                    }
                    hme_refine(c, 1, li, ri, (int16_t)p->hme_l1_sa.width, (int16_t)p->hme_l1_sa.height,
                               c->l0x[li][ri][sr_w][sr_h] >> 1, c->l0y[li][ri][sr_w][sr_h] >> 1,
                               &c->l1sad[li][ri][sr_w][sr_h], &c->l1x[li][ri][sr_w][sr_h], &c->l1y[li][ri][sr_w][sr_h]);
                }
        }
}

/* ME:2127-2177 hme_level2_b64 (no do_ref / zz early-outs at this level, as in the reference) */
static void stage_hme_l2(B64Ctx *c) {
    const MeGpuParams *p = c->p;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (!(p->temporal_layer_index > 0 || li == 0))
                continue;
            for (int sr_h = 0; sr_h < p->num_hme_sa_h; sr_h++)
                for (int sr_w = 0; sr_w < p->num_hme_sa_w; sr_w++) {
                    if (p->prev_me_stage_based_exit_th &&
                        c->l1sad[li][ri][sr_w][sr_h] < (p->prev_me_stage_based_exit_th >> 2)) {
                        c->l2x[li][ri][sr_w][sr_h]   = c->l1x[li][ri][sr_w][sr_h];
                        c->l2y[li][ri][sr_w][sr_h]   = c->l1y[li][ri][sr_w][sr_h];
                        c->l2sad[li][ri][sr_w][sr_h] = c->l1sad[li][ri][sr_w][sr_h];
                        continue;
                    }
                    hme_refine(c, 2, li, ri, (int16_t)p->hme_l2_sa.width, (int16_t)p->hme_l2_sa.height,
                               c->l1x[li][ri][sr_w][sr_h], c->l1y[li][ri][sr_w][sr_h], &c->l2sad[li][ri][sr_w][sr_h],
                               &c->l2x[li][ri][sr_w][sr_h], &c->l2y[li][ri][sr_w][sr_h]);
                }
        }
}

/* ME:2182-2380 set_final_seach_centre_sb: best quadrant of the deepest enabled level, scanned
 * [1][0],[0][1],[1][1] after [0][0] with strict '<'. The running (centre, sad) deliberately carries over
 * from the previous reference when a branch does not assign it (ME:2361-2371). */
static void stage_final_centre(B64Ctx *c) {
    const MeGpuParams *p = c->p;
    int16_t  hx = 0, hy = 0, cx = 0, cy = 0;
    uint64_t hsad = 0;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (p->temporal_layer_index > 0 || li == 0) {
                if (p->enable_hme_flag) {
                    const int16_t (*X)[2]  = NULL, (*Y)[2] = NULL;
                    const uint64_t (*S)[2] = NULL;
                    /* the three blocks of the reference are evaluated in order; a later one overrides */
                    if (p->enable_hme_level0_flag && !p->enable_hme_level1_flag && !p->enable_hme_level2_flag) {
                        X = c->l0x[li][ri]; Y = c->l0y[li][ri]; S = c->l0sad[li][ri];
                    }
                    if (p->enable_hme_level1_flag && !p->enable_hme_level2_flag) {
                        X = c->l1x[li][ri]; Y = c->l1y[li][ri]; S = c->l1sad[li][ri];
                    }
                    if (p->enable_hme_level2_flag) {
                        X = c->l2x[li][ri]; Y = c->l2y[li][ri]; S = c->l2sad[li][ri];
                    }
                    if (S) {
                        hx = X[0][0]; hy = Y[0][0]; hsad = S[0][0];
                        uint32_t w = 1, h = 0;
                        while (h < p->num_hme_sa_h) {
                            while (w < p->num_hme_sa_w) {
                                if (S[w][h] < hsad) { hx = X[w][h]; hy = Y[w][h]; hsad = S[w][h]; }
                                w++;
                            }
                            w = 0;
                            h++;
                        }
                    }
                    cx = hx;
                    cy = hy;
                }
            } else {
                cx = 0;
                cy = 0;
            }
            c->sc_x[li][ri]    = cx;
            c->sc_y[li][ri]    = cy;
            c->hme_sad[li][ri] = hsad;
        }
}

/* ME:2477-2518 hme_prune_ref_and_adjust_sr: loops run over all 2x4 slots, used or not */
static void stage_hme_prune(B64Ctx *c) {
    const MeGpuParams *p  = c->p;
    uint16_t           th = p->prune_ref_if_hme_sad_dev_bigger_than_th;
    if (p->enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        uint64_t best = ~(uint64_t)0;
        for (int i = 0; i < NL; i++)
            for (int j = 0; j < NR; j++) best = OR_MIN(best, c->hme_sad[i][j]);
        for (int li = 0; li < NL; li++)
            for (int ri = 1; ri < NR; ri++)
                if ((c->hme_sad[li][ri] - best) * 100 > (uint64_t)th * best)
                    c->do_ref[li][ri] = 0;
    }
    if (p->enable_me_sr_adjustment)
        for (int li = 0; li < NL; li++)
            for (int ri = 0; ri < NR; ri++) {
                if (OR_ABS(c->sc_x[li][ri]) <= p->reduce_me_sr_based_on_mv_length_th &&
                    OR_ABS(c->sc_y[li][ri]) <= p->reduce_me_sr_based_on_mv_length_th &&
                    c->hme_sad[li][ri] < p->stationary_hme_sad_abs_th)
                    c->sr_div[li][ri] = p->stationary_me_sr_divisor;
                else if (c->hme_sad[li][ri] < p->reduce_me_sr_based_on_hme_sad_abs_th)
                    c->sr_div[li][ri] = p->me_sr_divisor_for_low_hme_sad;
            }
}

/* z-order index of the 8x8 (level 3), 16x16 (2), 32x32 (1) block at raster (bx,by) */
static int z_index(int bx, int by, int levels) {
    int z = 0;
    for (int b = 0; b < levels; b++) z |= (((bx >> b) & 1) << (2 * b)) | (((by >> b) & 1) << (2 * b + 1));
    return z;
}

/* One search position: the 85 sub-sampled (or full) SADs and the strict '<' best update.
 * ME:98-205 (single position) and ME:210-425 (eight positions) compute the same values with the same
 * update rule; only the loop nest differs, which cannot change which position wins because every PU is
 * visited in scan order. PU order in best arrays: 0 = 64x64, 1..4 32x32, 5..20 16x16 (z), 21..84 8x8 (z). */
static void eval_position(B64Ctx *c, int li, int ri, const uint8_t *src, int src_stride, const uint8_t *ref,
                          int ref_stride, int16_t mvx, int16_t mvy, int sub) {
    uint32_t s8[64], s16[16], s32[4], s64 = 0;
    memset(s16, 0, sizeof s16);
    memset(s32, 0, sizeof s32);
    for (int by = 0; by < 8; by++)
        for (int bx = 0; bx < 8; bx++) {
            const uint8_t *s = src + (ptrdiff_t)(8 * by) * src_stride + 8 * bx;
            const uint8_t *r = ref + (ptrdiff_t)(8 * by) * ref_stride + 8 * bx;
            uint32_t       v = sub ? oracle_nxm_sad(s, 2 * src_stride, r, 2 * ref_stride, 4, 8) << 1
                                   : oracle_nxm_sad(s, src_stride, r, ref_stride, 8, 8);
            s8[z_index(bx, by, 3)] = v;
            s16[z_index(bx >> 1, by >> 1, 2)] += v;
            s32[z_index(bx >> 2, by >> 2, 1)] += v;
            s64 += v;
        }
    uint32_t  mv = ((uint32_t)(uint16_t)mvy << 16) | (uint16_t)mvx;
    uint32_t *bs = c->best_sad[li][ri], *bm = c->best_mv[li][ri];
    if (s64 < bs[0]) { bs[0] = s64; bm[0] = mv; }
    for (int i = 0; i < 4; i++)
        if (s32[i] < bs[1 + i]) { bs[1 + i] = s32[i]; bm[1 + i] = mv; }
    for (int i = 0; i < 16; i++)
        if (s16[i] < bs[5 + i]) { bs[5 + i] = s16[i]; bm[5 + i] = mv; }
    for (int i = 0; i < 64; i++)
        if (s8[i] < bs[21 + i]) { bs[21 + i] = s8[i]; bm[21 + i] = mv; }
    c->ops += sub ? 64 * 32 : 64 * 64;
}

/* ME:781-817 open_loop_me_fullpel_search_sblock: y-major, x-minor scan */
static void fullpel_search(B64Ctx *c, int li, int ri, int16_t x0, int16_t y0, int w, int h) {
    const OraclePlane *rp  = &c->ref[li][ri]->full;
    const uint8_t     *src = plane_at(&c->src->full, c->org_x, c->org_y);
    int                sub = c->p->me_search_method == ME_GPU_SUB_SAD_SEARCH;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++)
            eval_position(c, li, ri, src, c->src->full.stride, plane_at(rp, c->org_x + x0 + x, c->org_y + y0 + y),
                          rp->stride, (int16_t)(x0 + x), (int16_t)(y0 + y), sub);
}

/* ME:1139-1206 check_00_center */
static uint32_t check_00_center(B64Ctx *c, int li, int ri, int16_t *cx, int16_t *cy) {
    const OraclePlane *rp  = &c->ref[li][ri]->full;
    const uint8_t     *src = plane_at(&c->src->full, c->org_x, c->org_y);
    int                ss  = c->src->full.stride;
    int16_t            ox = (int16_t)c->org_x, oy = (int16_t)c->org_y;
    uint32_t           zero_sad;
    if (c->p->me_early_exit_th)
        zero_sad = c->zz_sad[li][ri];
    else {
        zero_sad = oracle_nxm_sad(src, ss << 1, plane_at(rp, ox, oy), rp->stride << 1, c->b64_h >> 1, c->b64_w);
        c->ops += (uint64_t)(c->b64_h >> 1) * c->b64_w;
    }
    zero_sad <<= 1;
    if (ox + *cx < -63) *cx = (int16_t)(-63 - ox);
    if (ox + *cx > rp->width - 1) *cx = (int16_t)(*cx - ((ox + *cx) - (rp->width - 1)));
    if (oy + *cy < -63) *cy = (int16_t)(-63 - oy);
    if (oy + *cy > rp->height - 1) *cy = (int16_t)(*cy - ((oy + *cy) - (rp->height - 1)));
    uint32_t hme_sad = oracle_nxm_sad(src, ss << 1, plane_at(rp, ox + *cx, oy + *cy), rp->stride << 1, c->b64_h >> 1,
                                      c->b64_w)
        << 1;
    c->ops += (uint64_t)(c->b64_h >> 1) * c->b64_w;
    /* COST_PRECISION shift applies to both sides, so comparing the SADs is equivalent (ME:1188-1204) */
    if ((uint64_t)zero_sad <= (uint64_t)hme_sad) {
        *cx = 0;
        *cy = 0;
    }
    return hme_sad;
}

/* ME:1249-1516 integer_search_b64 */
static void stage_integer_search(B64Ctx *c) {
    const MeGpuParams *p = c->p;
    int16_t            ox = (int16_t)c->org_x, oy = (int16_t)c->org_y;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (!c->do_ref[li][ri])
                continue;
            int16_t  cx = c->sc_x[li][ri], cy = c->sc_y[li][ri];
            uint16_t dist = ref_distance(c, li, ri);
            if (p->me_type != ME_GPU_ME_MCTF) /* ME:1300-1302 */
                dist = scaled_distance(dist);
            int16_t  w = (int16_t)OR_MIN(p->me_sa.sa_min.width * dist, p->me_sa.sa_max.width);
            int16_t  h = (int16_t)OR_MIN(p->me_sa.sa_min.height * dist, p->me_sa.sa_max.height);
            if (p->mv_sa_adj_enabled && (!p->mv_sa_adj_nearest_ref_only || ri == 0)) {
                if (OR_ABS(cx) > p->mv_sa_adj_mv_size_th) w = (int16_t)(w * p->mv_sa_adj_sa_multiplier);
                if (OR_ABS(cy) > p->mv_sa_adj_mv_size_th) h = (int16_t)(h * p->mv_sa_adj_sa_multiplier);
            }
            /* w, h are int16 and the divisor uint32, so the division is unsigned (ME:1314-1318) */
            w = (int16_t)((OR_MAX(1, ((uint32_t)w / c->sr_div[li][ri])) + 7) & ~0x07);
            h = (int16_t)OR_MAX(3, ((uint32_t)h / c->sr_div[li][ri]));
            const int16_t h_before = h, w_before = w;
            uint64_t      best_hme_sad = ~(uint64_t)0;
            if (p->me_early_exit_th) {
                if (c->zz_sad[li][ri] < p->me_early_exit_th / 6)
                    w = h = 1;
            } else {
                int hme_accurate = 1;
                if ((cx != 0 || cy != 0) && p->is_ref) {
                    best_hme_sad = check_00_center(c, li, ri, &cx, &cy);
                    if (cx == 0 && cy == 0)
                        hme_accurate = 0;
                }
                if (p->enable_me_sr_adjustment == 2) {
                    if ((hme_accurate && best_hme_sad < 24 * 24) || (p->is_ref && c->hme_sad[li][ri] < 24 * 24))
                        h = (int16_t)(h / 2);
                    if ((li || ri) && c->best_sad[0][0][0] < 5000 && h == h_before && w == w_before) {
                        h = (int16_t)(h >> 1);
                        w = (int16_t)(w >> 1);
                    }
                }
            }
            /* svt_initialize_buffer_32bits(..., 21, 1, MAX_SAD_VALUE) = 85 words (ME:1366) */
            for (int k = 0; k < NPU; k++) c->best_sad[li][ri][k] = OR_MAX_SAD_VALUE;
            /* 8x8-variance probe of the centre position (ME:1391-1439); its result stays in the best arrays */
            if (p->me_8x8_var_enabled && (w * h > 24)) {
                fullpel_search(c, li, ri, cx, cy, 1, 1);
                const uint32_t *b8   = &c->best_sad[li][ri][21];
                const uint32_t  mean = c->best_sad[li][ri][0] / 64;
                uint32_t        ssq  = 0;
                for (int i = 0; i < 64; i++) {
                    int32_t d = (int32_t)b8[i] - (int32_t)mean;
                    ssq += (uint32_t)(d * d);
                }
                uint32_t var = ssq / 64;
                if (var > p->me_sr_mult2_th) {
                    w = (int16_t)((OR_MAX(1, w * 3 / 2) + 7) & ~0x7);
                    h = (int16_t)OR_MAX(1, h * 3 / 2);
                }
                if (var < p->me_sr_div4_th) {
                    w = (int16_t)((OR_MAX(1, w >> 2) + 7) & ~0x7);
                    h = (int16_t)OR_MAX(1, h >> 2);
                    h = (int16_t)OR_MAX(3, h);
                } else if (var < p->me_sr_div2_th) {
                    w = (int16_t)((OR_MIN(w, w >> 1) + 7) & ~0x7);
                    h = (int16_t)OR_MIN(h, h >> 1);
                    h = (int16_t)OR_MAX(3, h);
                }
            }
            int16_t x0 = (int16_t)(cx - (w >> 1)), y0 = (int16_t)(cy - (h >> 1));
            clamp_axis(ox, &x0, &w, 63, c->pic_w);
            w = (int16_t)((w < 8) ? w : w & ~0x07);
            clamp_axis(oy, &y0, &h, 63, c->pic_h);
            c->sa_final[li][ri][0] = x0; c->sa_final[li][ri][1] = y0;
            c->sa_final[li][ri][2] = w;  c->sa_final[li][ri][3] = h;
            fullpel_search(c, li, ri, x0, y0, w, h);
        }
}

/* ME:1522-1565 me_prune_ref */
static void stage_me_prune(B64Ctx *c) {
    const MeGpuParams *p = c->p;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (!c->do_ref[li][ri]) {
                c->hme_sad[li][ri] = (uint64_t)(OR_MAX_SAD_VALUE * 64);
                continue;
            }
            uint64_t s = 0;
            for (int k = 0; k < 64; k++) s += c->best_sad[li][ri][21 + k];
            c->hme_sad[li][ri] = s;
        }
    uint16_t th = p->prune_ref_if_me_sad_dev_bigger_than_th;
    if (p->enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        uint64_t best = ~(uint64_t)0;
        for (int i = 0; i < NL; i++)
            for (int j = 0; j < NR; j++) best = OR_MIN(best, c->hme_sad[i][j]);
        for (int li = 0; li < NL; li++)
            for (int ri = 1; ri < NR; ri++)
                if ((c->hme_sad[li][ri] - best) * 100 > (uint64_t)th * best)
                    c->do_ref[li][ri] = 0;
    }
}

/* z-order PU index (MeContext order) -> raster PU index (output order); equals the reference's
 * z_to_raster[85] table (ME:2520-2531), derived instead of tabulated. */
static uint8_t z_to_raster(int n) {
    if (n < 5)
        return (uint8_t)n;
    int base = n < 21 ? 5 : 21, bits = n < 21 ? 2 : 3, z = n - base, bx = 0, by = 0;
    for (int b = 0; b < bits; b++) {
        bx |= ((z >> (2 * b)) & 1) << b;
        by |= ((z >> (2 * b + 1)) & 1) << b;
    }
    return (uint8_t)(base + by * (1 << bits) + bx);
}

/* MeCandidate bitfield byte image (Codec/me_sb_results.h:28-34): direction:2, ref_idx_l0:2, ref_idx_l1:2,
 * ref0_list:1, ref1_list:1, allocated from the least significant bit by gcc on x86-64. */
static uint8_t cand_byte(int direction, int ref_idx_l0, int ref_idx_l1, int ref0_list, int ref1_list) {
    return (uint8_t)((direction & 3) | ((ref_idx_l0 & 3) << 2) | ((ref_idx_l1 & 3) << 4) | ((ref0_list & 1) << 6) |
                     ((ref1_list & 1) << 7));
}

static int pu_exported(const MeGpuParams *p, int n_idx) { /* use_me_pu, ME:2704-2706 */
    return p->enable_me_16x16 ? (p->enable_me_8x8 || n_idx < 21) : n_idx < 5;
}
static int exported_pu_count(const MeGpuParams *p) { /* ME:2552-2554 */
    return p->enable_me_16x16 ? (p->enable_me_8x8 ? p->max_number_of_pus_per_sb : 21) : 5;
}

typedef struct B64Out { /* views into the caller's slabs for this b64 */
    uint8_t  *total;
    uint8_t  *cand;
    uint32_t *mv;
} B64Out;

/* ME:2646-2697 construct_me_candidate_array_single_ref */
static void cand_single_ref(B64Ctx *c, B64Out *o) {
    const MeGpuParams *p = c->p;
    memset(o->total, 1, exported_pu_count(p));
    for (int n = 0; n < p->max_number_of_pus_per_sb; n++) {
        int pu                = z_to_raster(n);
        c->me_distortion[pu]  = c->best_sad[0][0][n];
        if (!c->do_ref[0][0] || !pu_exported(p, n))
            continue;
        o->cand[pu * p->max_cand] = cand_byte(0, 0, 0, 0, 0);
        o->mv[pu * p->max_refs]   = c->best_mv[0][0][n];
    }
}

/* ME:2532-2645 construct_me_candidate_array_mrp_off (one reference per list) */
static void cand_mrp_off(B64Ctx *c, B64Out *o) {
    const MeGpuParams *p     = c->p;
    uint32_t           lists = p->num_of_list_to_search;
    uint8_t            org[2];
    org[0] = c->do_ref[0][0];
    org[1] = (lists == 1) ? 0 : c->do_ref[1][0];
    if (lists < 2 || !c->do_ref[1][0])
        lists = 1;
    const uint32_t prune_th = (org[0] && org[1]) ? (uint32_t)p->prune_me_candidates_th : 0;
    memset(o->total, 1, exported_pu_count(p));
    for (int n = 0; n < p->max_number_of_pus_per_sb; n++) {
        int      pu = z_to_raster(n), use = pu_exported(p, n), off = 0;
        uint8_t  blk[2] = {org[0], org[1]};
        uint32_t s0 = c->best_sad[0][0][n], s1 = c->best_sad[1][0][n];
        uint32_t best = (org[0] && org[1]) ? OR_MIN(s0, s1) : org[0] ? s0 : s1;
        c->me_distortion[pu] = best;
        int min_list = -1;
        if (p->use_best_unipred_cand_only && blk[0] && blk[1])
            min_list = s0 < s1 ? 0 : 1;
        for (uint32_t li = 0; li < lists && (use || off == 0); li++) {
            if (!blk[li])
                continue;
            if (prune_th > 0) {
                uint32_t d = (c->best_sad[li][0][n] - best) * 100;
                if (d > best * prune_th) {
                    blk[li] = 0;
                    continue;
                }
            }
            if (min_list != -1 && min_list != (int)li) {
                if (use)
                    o->mv[pu * p->max_refs + (li ? p->max_l0 : 0)] = c->best_mv[li][0][n];
                continue;
            }
            if (use) {
                o->cand[pu * p->max_cand + off]                = cand_byte(li, 0, 0, li == 0 ? li : 24, li == 1 ? li : 24);
                o->mv[pu * p->max_refs + (li ? p->max_l0 : 0)] = c->best_mv[li][0][n];
            }
            off++;
        }
        if (blk[0] && blk[1] && use) {
            o->cand[pu * p->max_cand + off] = cand_byte(2 /*BI_PRED*/, 0, 0, 0, 1);
            o->total[pu]                    = (uint8_t)(off + 1);
        }
    }
}

/* ME:2698-2835 construct_me_candidate_array (general MRP case) */
static void cand_general(B64Ctx *c, B64Out *o) {
    const MeGpuParams *p     = c->p;
    const uint32_t     lists = p->num_of_list_to_search;
    const uint8_t     *nref  = p->num_of_ref_pic_to_search;
    for (int n = 0; n < p->max_number_of_pus_per_sb; n++) {
        int            pu  = (n > 4) ? z_to_raster(n) : n;
        int            use = pu_exported(p, n), off = 0;
        uint8_t        blk[NL][NR];
        const uint32_t prune_th = (uint32_t)p->prune_me_candidates_th;
        uint32_t       best     = OR_MAX_U32;
        memset(blk, 0, sizeof blk);
        for (uint32_t li = 0; li < lists; li++)
            for (uint32_t ri = 0; ri < nref[li]; ri++) {
                blk[li][ri] = c->do_ref[li][ri];
                if (blk[li][ri] && c->best_sad[li][ri][n] < best)
                    best = c->best_sad[li][ri][n];
            }
        c->me_distortion[pu] = best;
        uint8_t *cand = use ? &o->cand[pu * p->max_cand] : NULL;
        for (uint32_t li = 0; li < lists && (use || off == 0); li++)
            for (uint32_t ri = 0; ri < nref[li] && (use || off == 0); ri++) {
                if (!blk[li][ri])
                    continue;
                if (prune_th > 0) {
                    uint32_t d = (c->best_sad[li][ri][n] - best) * 100;
                    if (d > best * prune_th) {
                        blk[li][ri] = 0;
                        continue;
                    }
                }
                if (use) {
                    cand[off] = cand_byte(li, ri, ri, li == 0 ? li : 24, li == 1 ? li : 24);
                    o->mv[pu * p->max_refs + (li ? p->max_l0 : 0) + ri] = c->best_mv[li][ri][n];
                }
                off++;
            }
        if (lists == 2 && use) {
            for (uint32_t a = 0; a < nref[0]; a++)
                for (uint32_t b = 0; b < nref[1]; b++) {
                    if (p->only_l_bwd && (a > 0 || b > 0))
                        continue;
                    if (blk[0][a] && blk[1][b])
                        cand[off++] = cand_byte(2, a, b, 0, 1);
                }
            if (!p->only_l_bwd) {
                for (uint32_t a = 1; a < nref[0]; a++)
                    if (blk[0][0] && blk[0][a])
                        cand[off++] = cand_byte(2, 0, a, 0, 0);
                if (nref[1] == 3 && blk[1][0] && blk[1][2])
                    cand[off++] = cand_byte(2, 0, 2, 1, 1);
            }
        }
        if (use)
            o->total[pu] = (uint8_t)off;
    }
}

/* ME:2964-3007 compute_distortion */
static void stage_distortion(B64Ctx *c, MeGpuOutputs *out, uint32_t b) {
    const uint32_t *d = c->me_distortion;
    uint32_t d64 = d[0], d32 = 0, d16 = 0, d8 = 0;
    for (int i = 0; i < 4; i++) d32 += d[1 + i];
    for (int i = 0; i < 16; i++) d16 += d[5 + i];
    for (int i = 0; i < 64; i++) d8 += d[21 + i];
    uint64_t mean = d8 / 64, ssq = 0;
    for (int i = 0; i < 64; i++) {
        int64_t diff = (int64_t)d[21 + i] - (int64_t)mean;
        ssq += (uint64_t)(diff * diff);
    }
    out->me_8x8_cost_variance[b] = (uint32_t)(ssq / 64);
    out->rc_me_distortion[b]     = (c->p->input_resolution <= 2 /*INPUT_SIZE_480p_RANGE*/) ? d8 : d16;
    const uint32_t pix           = c->b64_w * c->b64_h; /* b64_geom width*height */
    out->me_64x64_distortion[b]  = (d64 * 4096u) / pix;
    out->me_32x32_distortion[b]  = (d32 * 4096u) / pix;
    out->me_16x16_distortion[b]  = (d16 * 4096u) / pix;
    out->me_8x8_distortion[b]    = (d8 * 4096u) / pix;
}

/* ME:2838-2961 perform_gm_detection. As in the reference, the candidate is fetched with the z-order
 * n_idx used as if it were a raster PU index, and the MV is read from the z-order best-MV array. */
static void stage_gm_detection(B64Ctx *c, const B64Out *o, MeGpuOutputs *out, uint32_t b) {
    const MeGpuParams *p = c->p;
    uint64_t stationary = 0, total = 0, cnt[NL][NR][2][2];
    memset(cnt, 0, sizeof cnt);
    const int small = p->input_resolution <= 2; /* INPUT_SIZE_480p_RANGE */
    const int count = small ? 64 : 16;
    for (int i = 0; i < count; i++) {
        int n = small ? 21 + i : 5 + i;
        if (small) {
            /* me_idx_85_8x8_to_16x16_conversion / me_idx_16x16_to_parent_32x32_conversion
             * (Codec/definitions.h:2613-2633) are raster tables: entry (row, col) -> parent at (row/2, col/2). */
            if (!p->enable_me_8x8) {
                int r = n - 21, row = r / 8, col = r % 8;
                n     = 5 + (row / 2) * 4 + (col / 2);
                if (!p->enable_me_16x16)
                    n = 1 + ((n - 5) / 4 / 2) * 2 + (((n - 5) % 4) / 2);
            }
        } else if (!p->enable_me_16x16) {
            int r = n - 5, row = r / 4, col = r % 4;
            n     = 1 + (row / 2) * 2 + (col / 2);
        }
        uint8_t  cb  = o->cand[n * p->max_cand];
        int      dir = cb & 3, l0 = (cb >> 2) & 3, l1 = (cb >> 4) & 3, r0 = (cb >> 6) & 1, r1 = (cb >> 7) & 1;
        uint32_t li  = (dir == 0 || dir == 2) ? r0 : r1;
        uint32_t ri  = (dir == 0 || dir == 2) ? l0 : l1;
        uint64_t ref_num = c->ref[li][ri] ? c->ref[li][ri]->picture_number : 0;
        int      active_th;
        if (small) {
            uint64_t a = p->picture_number, bnum = ref_num;
            int16_t  d16  = (int16_t)(OR_MAX(a, bnum) - OR_MIN(a, bnum));
            uint16_t dist = (uint16_t)OR_ABS(d16);
            active_th     = p->gm_use_distance_based_active_th ? OR_MAX(dist >> 1, 4) : 4;
        } else {
            int16_t  d16  = (int16_t)(p->picture_number - ref_num);
            uint16_t dist = (uint16_t)OR_ABS(d16);
            active_th     = p->gm_use_distance_based_active_th ? OR_MAX(dist * 16, 32) : 32;
        }
        uint32_t mv = c->best_mv[li][ri][n];
        int      mx = (int16_t)(mv & 0xFFFF) << 2, my = (int16_t)(mv >> 16) << 2;
        if (mx < -active_th) cnt[li][ri][0][0]++;
        else if (mx > active_th) cnt[li][ri][0][1]++;
        if (my < -active_th) cnt[li][ri][1][0]++;
        else if (my > active_th) cnt[li][ri][1][1]++;
        int st = small ? 0 : 4;
        if (abs(mx) <= st && abs(my) <= st)
            stationary++;
        total++;
    }
    if (stationary > (total * 5) / 100)
        out->stationary_block_present_sb[b] = 1;
    for (int l = 0; l < NL; l++)
        for (int r = 0; r < NR; r++)
            for (int k = 0; k < 2; k++)
                for (int s = 0; s < 2; s++)
                    if (cnt[l][r][k][s] > total / 2)
                        out->rc_me_allow_gm[b] = 1;
}

/* what the temporal filter reads from the MeContext after an ME_MCTF call (temporal_filtering.c:1867-1870, 2640-2690) */
static void tf_outputs(const B64Ctx *c, MeGpuOutputs *out, uint32_t b, int early_exit) {
    if (out->tf_best_sad) memcpy(out->tf_best_sad + (size_t)b * NPU, c->best_sad[0][0], sizeof c->best_sad[0][0]);
    if (out->tf_best_mv) memcpy(out->tf_best_mv + (size_t)b * NPU, c->best_mv[0][0], sizeof c->best_mv[0][0]);
    if (out->tf_hme_centre) {
        out->tf_hme_centre[2 * b]     = c->sc_x[0][0];
        out->tf_hme_centre[2 * b + 1] = c->sc_y[0][0];
    }
    if (out->tf_hme_sad) out->tf_hme_sad[b] = (uint32_t)c->hme_sad[0][0];
    if (out->tf_early_exit) out->tf_early_exit[b] = (uint8_t)early_exit;
}

static void parity_taps(const B64Ctx *c, MeGpuOutputs *out, uint32_t b);

/* ME:3076-3153 svt_aom_motion_estimation_b64 for ME_OPEN_LOOP and ME_MCTF */
static void me_b64(B64Ctx *c, MeGpuOutputs *out, uint32_t b, uint32_t pu_count) {
    const MeGpuParams *p = c->p;
    init_b64(c);
    /* ME:2441-2475 hme_b64 */
    if (p->me_early_exit_th || p->me_safe_limit_zz_th)
        stage_zz(c);
    if (p->prehme_enable)
        stage_prehme(c);
    if (p->enable_hme_flag) {
        if (p->enable_hme_level0_flag) stage_hme_l0(c);
        if (p->enable_hme_level1_flag) stage_hme_l1(c);
        if (p->enable_hme_level2_flag) stage_hme_l2(c);
    }
    stage_final_centre(c);
    if (p->me_type == ME_GPU_ME_MCTF) {
        /* ME:2469-2474 (end of hme_b64), :3109-3113 early exit, :3126 nothing after the integer search */
        if (OR_ABS(c->sc_x[0][0]) > OR_ABS(c->sc_y[0][0]))
            out->tf_tot_horz_blks++;
        else
            out->tf_tot_vert_blks++;
        const int early_exit = c->hme_sad[0][0] < p->tf_me_exit_th;
        if (!early_exit)
            stage_integer_search(c);
        tf_outputs(c, out, b, early_exit);
        parity_taps(c, out, b);
        return;
    }
    const int prune_ref = p->enable_hme_flag; /* && me_type != ME_MCTF, ME:3103 */
    if (prune_ref)
        stage_hme_prune(c);
    stage_integer_search(c);
    if (prune_ref && p->enable_me_hme_ref_pruning)
        stage_me_prune(c);

    B64Out o;
    o.total = out->total_me_candidate_index + (size_t)b * pu_count;
    o.cand  = out->me_candidate_array + (size_t)b * pu_count * p->max_cand;
    o.mv    = out->me_mv_array + (size_t)b * pu_count * p->max_refs;
    if (p->num_of_ref_pic_to_search[0] == 1 && p->num_of_ref_pic_to_search[1] == 0)
        cand_single_ref(c, &o);
    else if (p->num_of_ref_pic_to_search[0] == 1 && p->num_of_ref_pic_to_search[1] == 1)
        cand_mrp_off(c, &o);
    else
        cand_general(c, &o);
    stage_distortion(c, out, b);
    out->stationary_block_present_sb[b] = 0;
    out->rc_me_allow_gm[b]              = 0;
    if (p->gm_enabled)
        stage_gm_detection(c, &o, out, b);

    parity_taps(c, out, b);
}

static void parity_taps(const B64Ctx *c, MeGpuOutputs *out, uint32_t b) {
    if (out->p_sb_best_sad) memcpy(out->p_sb_best_sad + (size_t)b * NL * NR * NPU, c->best_sad, sizeof c->best_sad);
    if (out->p_sb_best_mv) memcpy(out->p_sb_best_mv + (size_t)b * NL * NR * NPU, c->best_mv, sizeof c->best_mv);
    for (int li = 0; li < NL; li++)
        for (int ri = 0; ri < NR; ri++) {
            size_t k = ((size_t)b * NL + li) * NR + ri;
            if (out->hme_search_centre) {
                out->hme_search_centre[2 * k]     = c->sc_x[li][ri];
                out->hme_search_centre[2 * k + 1] = c->sc_y[li][ri];
            }
            if (out->hme_sad) out->hme_sad[k] = c->hme_sad[li][ri];
            if (out->do_ref) out->do_ref[k] = c->do_ref[li][ri];
            if (out->search_area) memcpy(out->search_area + 4 * k, c->sa_final[li][ri], 4 * sizeof(int16_t));
        }
}

int oracle_me_b64_range(const OraclePicture *src, const OraclePicture *const refs[NL][NR], const MeGpuParams *params,
                        MeGpuOutputs *out, uint32_t first, uint32_t count, uint64_t *byte_sad_ops) {
    if (!src || !params || !out)
        return ME_GPU_ERR_BAD_ARG;
    if (params->num_hme_sa_w != 2 || params->num_hme_sa_h != 2) /* get_worst_quadrant, ME:1875 */
        return ME_GPU_ERR_UNSUPPORTED;
    const int pic_w = (src->full.width + 7) & ~7, pic_h = (src->full.height + 7) & ~7;
    const uint32_t bw = (pic_w + 63) / 64, bh = (pic_h + 63) / 64, total = bw * bh;
    for (int li = 0; li < params->num_of_list_to_search; li++)
        for (int ri = 0; ri < params->num_of_ref_pic_to_search[li]; ri++)
            if (!refs[li][ri])
                return ME_GPU_ERR_BAD_ARG;
    if (out->b64_count < total)
        return ME_GPU_ERR_BAD_ARG;
    B64Ctx *c = (B64Ctx *)calloc(1, sizeof *c);
    if (!c)
        return ME_GPU_ERR_NO_MEMORY;
    c->p = params;
    c->src = src;
    memcpy(c->ref, refs, sizeof c->ref);
    c->pic_w = pic_w;
    c->pic_h = pic_h;
    const uint32_t pu_count = (uint32_t)exported_pu_count(params);
    if (params->me_type == ME_GPU_ME_MCTF) {
        if (params->num_of_list_to_search != 1 || params->num_of_ref_pic_to_search[0] != 1)
            return ME_GPU_ERR_BAD_ARG;
        if (first == 0)
            out->tf_tot_horz_blks = out->tf_tot_vert_blks = 0;
    } else if (params->me_type != ME_GPU_ME_OPEN_LOOP)
        return ME_GPU_ERR_BAD_ARG;
    for (uint32_t b = first; b < first + count && b < total; b++) {
        c->org_x = (int32_t)(b % bw) * 64;
        c->org_y = (int32_t)(b / bw) * 64;
        c->b64_w = (uint32_t)OR_MIN(64, pic_w - c->org_x); /* ME:3093-3100 */
        c->b64_h = (uint32_t)OR_MIN(64, pic_h - c->org_y);
        me_b64(c, out, b, pu_count);
    }
    if (byte_sad_ops)
        *byte_sad_ops += c->ops;
    out->b64_count = total;
    out->pu_count  = pu_count;
    free(c);
    return ME_GPU_OK;
}

int oracle_me_picture(const OraclePicture *src, const OraclePicture *const refs[NL][NR], const MeGpuParams *params,
                      MeGpuOutputs *out) {
    return oracle_me_b64_range(src, refs, params, out, 0, 0xFFFFFFFFu / 2, NULL);
}
