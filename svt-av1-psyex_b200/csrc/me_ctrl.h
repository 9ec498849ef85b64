/*
 * me_ctrl.h - the integer control logic of open-loop ME between the search kernels.
 *
 * One thread per 64x64 block runs these stages on the device (me_kernels.cu, k_ctrl); each stage consumes
 * the results of the search tasks the previous stage emitted and emits the next ones. The functions are
 * plain integer code (__host__ __device__), so tests/emul can run the identical logic on the CPU with a
 * scalar task executor - that is a test harness, not a product path.
 *
 * Reference (ME = /root/reference/Source/Lib/Codec/motion_estimation.c):
 *   stage_init           init_me_hme_data                     ME:3010-3070
 *   zz                   init_zz_sad / get_zz_sad             ME:2382-2437, 1667-1689
 *   pre-HME              prehme_b64 / prehme_core             ME:1722-1796, 1568-1666, 1693-1719
 *   HME L0               hme_level0_b64 / hme_level_0         ME:1906-2036, 820-920, 1800-1901
 *   HME L1 / L2          hme_level1_b64 / hme_level2_b64      ME:2041-2177, 923-1113
 *   final centre         set_final_seach_centre_sb            ME:2182-2380
 *   HME prune, SR adjust hme_prune_ref_and_adjust_sr          ME:2477-2518
 *   integer search prep  integer_search_b64, check_00_center  ME:1249-1516, 1139-1206
 *   epilogue             me_prune_ref, construct_me_candidate_array*, compute_distortion,
 *                        perform_gm_detection                 ME:1522-1565, 2532-2835, 2964-3007, 2838-2961
 */
#ifndef ME_CTRL_H
#define ME_CTRL_H

#include "me_types.h"

#define MC_MIN(a, b) ((a) < (b) ? (a) : (b))
#define MC_MAX(a, b) ((a) > (b) ? (a) : (b))
#define MC_ABS(a) ((a) < 0 ? -(a) : (a))

enum MeStage {
    ST_INIT = 0, /* init state, emit zero-MV SAD tasks */
    ST_PREHME0, /* emit pre-HME tasks of list 0 (and list 1 unless l1_early_exit) */
    ST_PREHME1, /* emit pre-HME tasks of list 1 (l1_early_exit only) */
    ST_HME_L0,
    ST_HME_L0B, /* HME level 0 of the references other than (0, 0), where their search area depends on (0, 0)'s result (ME:1810-1825) */
    ST_HME_L1,
    ST_HME_L2,
    ST_CENTRE, /* final centre, HME prune, search-area derivation, emit check_00_center SADs */
    ST_PROBE, /* centre decision, emit 8x8-variance probes (or the main search where there is no probe) */
    ST_MAIN, /* probe variance -> final area, emit the main search */
    ST_COUNT
};

enum MePending { PEND_NONE = 0, PEND_ZZ, PEND_PHME, PEND_L0, PEND_L1, PEND_L2, PEND_C00, PEND_PROBE };

typedef struct CtrlCtx {
    const MeJob        *j;
    const MeGpuParams  *p;
    B64State           *s;
    SearchTask         *t; /* [ME_SLOTS] of this b64 */
    const SearchResult *r; /* [ME_SLOTS] */
    FsTask             *f; /* [ME_NREF] */
    unsigned long long *keys; /* [ME_NREF][ME_NPU] */
    int32_t             org_x, org_y;
    uint32_t            b64_w, b64_h;
    int32_t             pic_w, pic_h;
    /* A stage can be run by several cooperating threads of one warp (k_hme): the loops over (list, reference) are
     * independent per reference, so thread `lane` takes reference slot `only` (= li * ME_NR + ri; -1: all of them) and
     * the steps that look across references are run by lane 0 alone, with a warp barrier between steps. One thread
     * (host emulation, legacy kernel): only = -1, lane = 0, nlanes = 1. */
    int32_t             only, lane, nlanes;
    /* integer-search stages (ST_PROBE / ST_MAIN): 0 = every reference; 1 = reference (0, 0) only, 2 = all the others. With
     * enable_me_sr_adjustment == 2 the search area of the other references depends on the FINISHED search of (0, 0)
     * (ME:1355-1364), so the picture takes the two stages twice, with the tiled searches of (0, 0) in between. */
    int32_t             fs_which;
} CtrlCtx;

#define MC_FS_MINE(c, li, ri) ((c)->fs_which == 0 || ((c)->fs_which == 1) == (!(li) && !(ri)))
/* the references other than (0, 0) size their HME level 0 area by (0, 0)'s level 0 centre (low delay, ME:1810-1825) */
ME_HD int mc_l0_two_phase(const MeGpuParams *p) {
    return p->enable_me_sr_adjustment && p->distance_based_hme_resizing && p->reduce_hme_l0_sr_th_min && p->reduce_hme_l0_sr_th_max;
}
ME_HD int mc_fs_two_pass(const MeGpuParams *p) { return p->enable_me_sr_adjustment == 2 && !p->me_early_exit_th; }

#define MC_MINE(c, li, ri) ((c)->only < 0 || (c)->only == (li) * ME_NR + (ri))
/* loop bounds over lists / references / task slots: everything for only < 0, else exactly this thread's reference, so that
 * cooperating lanes run the loop body together (SIMT) instead of one after the other */
#define MC_L0(c) ((c)->only < 0 ? 0 : (c)->only / ME_NR)
#define MC_L1(c, n) ((c)->only < 0 ? (int)(n) : MC_MIN((int)(n), (c)->only / ME_NR + 1))
#define MC_R0(c) ((c)->only < 0 ? 0 : (c)->only % ME_NR)
#define MC_R1(c, n) ((c)->only < 0 ? (int)(n) : MC_MIN((int)(n), (c)->only % ME_NR + 1))
#define MC_K0(c, per) ((c)->only < 0 ? 0 : (c)->only * (per))
#define MC_K1(c, per) ((c)->only < 0 ? ME_NREF * (per) : MC_MIN(ME_NREF * (per), ((c)->only + 1) * (per)))
#if defined(__CUDA_ARCH__)
#define MC_SYNC(c)                \
    do {                          \
        if ((c)->nlanes > 1)      \
            __syncwarp();         \
    } while (0)
#define MC_OPS_ADD(p, v) atomicAdd(reinterpret_cast<unsigned long long *>(p), (unsigned long long)(v))
#else
#define MC_SYNC(c) ((void)0)
#define MC_OPS_ADD(p, v) (*(p) += (v))
#endif

/* ME:1239-1243 svt_aom_get_scaled_picture_distance */
ME_HD uint16_t mc_scaled_distance(uint16_t dist) { return (uint16_t)(((dist * 5) / 8) + ((dist % 8) ? 1 : 0)); }

/* ME:1232-1234: ABS((int16_t)(pcs->picture_number - ref picture_number)) */
ME_HD uint16_t mc_ref_distance(const CtrlCtx *c, int li, int ri) {
    /* truncated to 16 bits BEFORE the absolute value, as the reference writes it: only the low 16 bits of the numbers count,
       so a caller may keep an identity tag in the high bits of the picture numbers it hands to the engine */
    const int16_t d = (int16_t)(c->p->picture_number - c->j->ref_picture_number[li][ri]);
    return (uint16_t)MC_ABS(d);
}

/* window clamp shared by ME:848-880, 949-982, 1050-1078, 1585-1629, 1445-1491. The left/top clamp moves the
 * origin only: the reference recomputes the overshoot with the already-moved origin, i.e. subtracts 0. */
ME_HD void mc_clamp_axis(int org, int16_t *origin, int16_t *size, int pad, int dim) {
    if (org + *origin < -pad)
        *origin = (int16_t)(-pad - org);
    if (org + *origin > dim - 1)
        *origin = (int16_t)(*origin - ((org + *origin) - (dim - 1)));
    if (org + *origin + *size > dim)
        *size = (int16_t)MC_MAX(1, *size - ((org + *origin + *size) - dim));
}

ME_HD void mc_clear_tasks(CtrlCtx *c) {
    for (int i = c->lane; i < ME_SLOTS; i += c->nlanes) c->t[i].valid = 0;
}

/* describe one svt_sad_loop_kernel call; sub halves the rows and doubles the strides (ME:891-909) */
ME_HD void mc_emit(CtrlCtx *c, int slot, int level, int li, int ri, int ox, int oy, int16_t x0, int16_t y0, uint32_t bw,
                   uint32_t bh, int sub, int skip, int16_t sa_w, int16_t sa_h) {
    SearchTask *t = &c->t[slot];
    t->valid      = 1;
    t->level      = (uint8_t)level;
    t->bw         = (uint8_t)bw;
    t->rows       = (uint8_t)(sub ? bh >> 1 : bh);
    t->row_step   = (uint8_t)(sub ? 2 : 1);
    t->skip_line  = (uint8_t)skip;
    t->ref_slot   = c->j->ref_slot[li][ri];
    t->sx         = (int16_t)ox;
    t->sy         = (int16_t)oy;
    t->rx         = (int16_t)(ox + x0);
    t->ry         = (int16_t)(oy + y0);
    t->sa_w       = sa_w;
    t->sa_h       = sa_h;
    t->x0         = x0;
    t->y0         = y0;
    uint64_t lines = (uint64_t)sa_h;
    if (bw == 16 && t->rows <= 16 && skip)
        lines = (uint64_t)(sa_h / 2);
    MC_OPS_ADD(&c->s->ops, lines * (uint64_t)sa_w * bw * t->rows);
}

/* result of a task as the callers of svt_sad_loop_kernel see it: sad doubled for sub-sampled SAD; x / y
 * only overwrite the previous value when some position beat the initial 0xffffff */
ME_HD uint32_t mc_result(const CtrlCtx *c, int slot, int16_t *x, int16_t *y) {
    const SearchResult *r = &c->r[slot];
    if (r->found) {
        *x = r->x;
        *y = r->y;
    }
    return c->t[slot].row_step == 2 ? r->sad * 2 : r->sad;
}

ME_HD int mc_ref_active(const CtrlCtx *c, int li) { return c->p->temporal_layer_index > 0 || li == 0; }

/* p_sb_best_sad = MAX_SAD_VALUE for every PU (svt_initialize_buffer_32bits, ME:1366), as keys. Separate from
 * mc_stage_init so that the device can spread the 680 stores over a whole CTA. */
ME_HD void mc_init_keys(unsigned long long *keys, int first, int stride) {
    for (int k = first; k < ME_NREF * ME_NPU; k += stride) keys[k] = ME_KEY_INIT;
}

#if defined(__CUDA_ARCH__)
#define MC_KEY_LOAD(p) __ldcg(p) /* keys are written with atomics at L2: read them there */
#else
#define MC_KEY_LOAD(p) (*(p))
#endif

/* ------------------------------------------------------------------------------------------------ */
ME_HD void mc_stage_init(CtrlCtx *c) {
    B64State *s = c->s;
    for (int li = MC_L0(c); li < MC_L1(c, ME_NL); li++)
        for (int ri = MC_R0(c); ri < MC_R1(c, ME_NR); ri++) {
            for (int a = 0; a < 2; a++)
                for (int b = 0; b < 2; b++) {
                    s->l0x[li][ri][a][b] = s->l0y[li][ri][a][b] = 0;
                    s->l1x[li][ri][a][b] = s->l1y[li][ri][a][b] = 0;
                    s->l2x[li][ri][a][b] = s->l2y[li][ri][a][b] = 0;
                    s->l0sad[li][ri][a][b] = s->l1sad[li][ri][a][b] = s->l2sad[li][ri][a][b] = 0;
                }
            s->sc_x[li][ri] = s->sc_y[li][ri] = 0;
            s->hme_sad[li][ri] = ME_MAX_U32;
            s->do_ref[li][ri]  = 1;
            s->sr_div[li][ri]  = 1;
            s->zz_sad[li][ri]  = ME_MAX_U32;
            for (int sr = 0; sr < 2; sr++) {
                MePreHme *d = &s->phme[li][ri][sr];
                d->sa_w = d->sa_h = 0;
                d->col = d->row = 0;
                d->sad          = 0;
                d->valid        = 0;
                s->performed_phme[li][ri][sr] = 0;
            }
            s->cx[li][ri] = s->cy[li][ri] = 0;
            s->sw[li][ri] = s->sh[li][ri] = 0;
            for (int k = 0; k < 4; k++) s->sa_final[li][ri][k] = 0;
            s->probed[li][ri] = s->searched[li][ri] = s->pend00[li][ri] = 0;
            c->f[li * ME_NR + ri].valid = 0;
        }
    if (c->lane == 0) {
        s->ops        = 0;
        s->ops_fs     = 0;
        s->ops_fs_inline = 0;
        s->tf_exit    = 0;
        c->s->pending = PEND_NONE;
    }
}

/* ME:2382-2437 / 1667-1689: zero-MV SAD of the block, every other row */
ME_HD void mc_zz_emit(CtrlCtx *c) {
    const MeGpuParams *p = c->p;
    if (!(p->me_early_exit_th || p->me_safe_limit_zz_th))
        return;
    for (int li = MC_L0(c); li < MC_L1(c, p->num_of_list_to_search); li++)
        for (int ri = MC_R0(c); ri < MC_R1(c, p->num_of_ref_pic_to_search[li]); ri++) {
            if (!mc_ref_active(c, li))
                continue;
            mc_emit(c, li * ME_NR + ri, 0, li, ri, c->org_x, c->org_y, 0, 0, c->b64_w, c->b64_h, 1, 0, 1, 1);
        }
    c->s->pending = PEND_ZZ;
}

ME_HD void mc_zz_consume(CtrlCtx *c) {
    const MeGpuParams *p = c->p;
    B64State          *s = c->s;
    uint32_t           best = ME_MAX_U32;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (!mc_ref_active(c, li))
                continue;
            int16_t  dx = 0, dy = 0;
            uint32_t zz = mc_result(c, li * ME_NR + ri, &dx, &dy); /* nxm_sad << 1 */
            zz          = (zz * 64 * 64) / (c->b64_w * c->b64_h);
            s->zz_sad[li][ri] = zz;
            best              = MC_MIN(best, zz);
        }
    if (p->temporal_layer_index > 0 && best < p->zz_sad_th)
        for (int li = 0; li < p->num_of_list_to_search; li++)
            for (int ri = 1; ri < p->num_of_ref_pic_to_search[li]; ri++)
                if ((uint32_t)((s->zz_sad[li][ri] - best) * 100) > (uint32_t)(p->zz_sad_pct * best))
                    s->do_ref[li][ri] = 0;
    if (p->me_safe_limit_zz_th) {
        int limit = p->hierarchical_levels > 0 && p->num_of_list_to_search == 2 &&
            p->pcs_temporal_layer_index >= p->hierarchical_levels && p->similar_brightness_refs &&
            s->zz_sad[0][0] < p->me_safe_limit_zz_th && s->zz_sad[1][0] < p->me_safe_limit_zz_th;
        if (limit)
            for (int li = 0; li < p->num_of_list_to_search; li++)
                for (int ri = 1; ri < p->num_of_ref_pic_to_search[li]; ri++) s->do_ref[li][ri] = 0;
    }
}

/* ME:1693-1719 check_prehme_early_exit */
ME_HD int mc_prehme_early_exit(CtrlCtx *c, int li, int ri, int sr) {
    const MeGpuParams *p = c->p;
    MePreHme          *d = &c->s->phme[li][ri][sr];
    if (p->me_early_exit_th && c->s->zz_sad[li][ri] < p->me_early_exit_th) {
        d->col = d->row = 0;
        d->sad   = 0;
        d->valid = 1;
        return 1;
    }
    if (p->prehme_l1_early_exit) {
        const MePreHme *l0 = &c->s->phme[0][ri][sr];
        if (li == 1 && l0->valid && (l0->sad < 32 * 32 || (MC_ABS(l0->col) < 16 && MC_ABS(l0->row) < 16))) {
            d->col   = (int16_t)-l0->col;
            d->row   = (int16_t)-l0->row;
            d->sad   = l0->sad;
            d->valid = 1;
            return 1;
        }
    }
    return 0;
}

/* ME:1722-1796 prehme_b64 (search part of list li) + ME:1568-1666 prehme_core */
ME_HD void mc_prehme_emit_list(CtrlCtx *c, int li) {
    const MeGpuParams *p = c->p;
    B64State          *s = c->s;
    if (li >= p->num_of_list_to_search || !mc_ref_active(c, li))
        return;
    const MePlaneGeom *pl = &c->j->g.plane[2];
    const int          ox = (int16_t)c->org_x >> 2, oy = (int16_t)c->org_y >> 2;
    if (c->only >= 0 && c->only / ME_NR != li)
        return;
    for (int ri = MC_R0(c); ri < MC_R1(c, p->num_of_ref_pic_to_search[li]); ri++) {
        uint32_t f = mc_scaled_distance(mc_ref_distance(c, li, ri));
        for (int sr = 0; sr < 2; sr++) {
            if (mc_prehme_early_exit(c, li, ri, sr))
                continue;
            MePreHme *d = &s->phme[li][ri][sr];
            if (!s->do_ref[li][ri]) {
                d->col = d->row = 0;
                d->sad = ME_MAX_U32;
                continue;
            }
            const MeGpuSearchAreaMinMax *cfg = &p->prehme_sa_cfg[sr];
            d->sa_w     = (uint16_t)MC_MIN(cfg->sa_min.width * f, (uint32_t)cfg->sa_max.width);
            d->sa_h     = (uint16_t)MC_MIN(cfg->sa_min.height * f, (uint32_t)cfg->sa_max.height);
            int16_t w   = (int16_t)d->sa_w, h = (int16_t)d->sa_h;
            int16_t x0  = (int16_t) - (w >> 1), y0 = (int16_t) - (h >> 1);
            mc_clamp_axis(ox, &x0, &w, pl->pad - 1, pl->width);
            mc_clamp_axis(oy, &y0, &h, pl->pad - 1, pl->height);
            mc_emit(c, (li * ME_NR + ri) * 2 + sr, 2, li, ri, ox, oy, x0, y0, c->b64_w >> 2, c->b64_h >> 2,
                    p->hme_search_method == ME_GPU_SUB_SAD_SEARCH, p->prehme_skip_search_line, w, h);
            s->performed_phme[li][ri][sr] = 1;
        }
    }
}

ME_HD void mc_prehme_consume_tasks(CtrlCtx *c) {
    for (int k = MC_K0(c, 2); k < MC_K1(c, 2); k++) {
        if (!c->t[k].valid)
            continue;
        MePreHme *d = &c->s->phme[(k >> 1) / ME_NR][(k >> 1) % ME_NR][k & 1];
        d->sad      = mc_result(c, k, &d->col, &d->row);
        d->col      = (int16_t)((d->col + c->t[k].x0) * 4);
        d->row      = (int16_t)((d->row + c->t[k].y0) * 4);
        d->valid    = 1;
    }
}

/* the part of prehme_b64 that follows the searches: list-1 mirror at temporal layer 0 and the prune */
ME_HD void mc_prehme_finish(CtrlCtx *c) {
    const MeGpuParams *p = c->p;
    B64State          *s = c->s;
    uint32_t           best = ME_MAX_U32;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (mc_ref_active(c, li)) {
                uint32_t m = MC_MIN(s->phme[li][ri][0].sad, s->phme[li][ri][1].sad);
                best       = MC_MIN(best, m);
            } else {
                for (int sr = 0; sr < 2; sr++) {
                    s->phme[1][ri][sr].col = (int16_t)-s->phme[0][ri][sr].col;
                    s->phme[1][ri][sr].row = (int16_t)-s->phme[0][ri][sr].row;
                    s->phme[1][ri][sr].sad = s->phme[0][ri][sr].sad;
                }
            }
        }
    if (p->temporal_layer_index > 0 && best < p->phme_sad_th)
        for (int li = 0; li < p->num_of_list_to_search; li++)
            for (int ri = 1; ri < p->num_of_ref_pic_to_search[li]; ri++) {
                if (!s->do_ref[li][ri])
                    continue;
                uint32_t v = MC_MIN(s->phme[li][ri][0].sad, s->phme[li][ri][1].sad);
                if ((uint32_t)((v - best) * 100) > (uint32_t)(p->phme_sad_pct * best))
                    s->do_ref[li][ri] = 0;
            }
}

ME_HD void mc_fill_quadrants(int16_t x[2][2], int16_t y[2][2], uint32_t s[2][2], int16_t vx, int16_t vy, uint32_t vs) {
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++) {
            x[a][b] = vx;
            y[a][b] = vy;
            s[a][b] = vs;
        }
}

/* ME:1800-1867 get_hme_l0_search_area. The reference edits me_ctx->hme_l0_sa in place and the caller
 * restores it (ME:1998-2003); a local copy is the same thing. */
ME_HD void mc_hme_l0_area(const CtrlCtx *c, int li, int ri, uint16_t dist, int16_t *sa_w, int16_t *sa_h) {
    const MeGpuParams    *p  = c->p;
    MeGpuSearchAreaMinMax sa = p->hme_l0_sa;
    if (p->enable_me_sr_adjustment && p->distance_based_hme_resizing) {
        uint8_t is_hor = 1, is_ver = 1, is_still = 0;
        if (p->reduce_hme_l0_sr_th_min && p->reduce_hme_l0_sr_th_max && (li || ri)) {
            int mvx = c->s->l0x[0][0][0][0], mvy = c->s->l0y[0][0][0][0];
            is_ver   = MC_ABS(mvx) < p->reduce_hme_l0_sr_th_min && MC_ABS(mvy) > p->reduce_hme_l0_sr_th_max;
            is_hor   = MC_ABS(mvx) > p->reduce_hme_l0_sr_th_max && MC_ABS(mvy) < p->reduce_hme_l0_sr_th_min;
            is_still = MC_ABS(mvx) < p->reduce_hme_l0_sr_th_min * 3 && MC_ABS(mvy) < p->reduce_hme_l0_sr_th_min * 3;
        }
        uint8_t x_off = is_hor ? 1 : 2, y_off = is_ver ? 1 : 2;
        if (p->enable_me_sr_adjustment == 2 && is_still)
            x_off = y_off = 4;
        sa.sa_min.width  = (uint16_t)(sa.sa_min.width / (x_off + ri));
        sa.sa_min.height = (uint16_t)(sa.sa_min.height / (y_off + ri));
        sa.sa_max.width  = (uint16_t)(sa.sa_max.width / (x_off + ri));
        sa.sa_max.height = (uint16_t)(sa.sa_max.height / (y_off + ri));
    }
    int32_t f = mc_scaled_distance(dist);
    int16_t w = (int16_t)(sa.sa_min.width / p->num_hme_sa_w);
    w         = (int16_t)MC_MIN(((w * f) + 15) & ~0x0F, ((sa.sa_max.width / p->num_hme_sa_w) + 15) & ~0x0F);
    int16_t h = (int16_t)(sa.sa_min.height / p->num_hme_sa_h);
    h         = (int16_t)MC_MIN(h * f, sa.sa_max.height / p->num_hme_sa_h);
    *sa_w     = w;
    *sa_h     = h;
}

/* ME:1906-2036 hme_level0_b64 (decisions + task emission); ME:820-920 hme_level_0 */
ME_HD void mc_l0_emit(CtrlCtx *c, int which /* 0 all, 1 reference (0, 0) only, 2 the others */) {
    const MeGpuParams *p  = c->p;
    B64State          *s  = c->s;
    const MePlaneGeom *pl = &c->j->g.plane[2];
    const int          ox = (int16_t)c->org_x >> 2, oy = (int16_t)c->org_y >> 2;
    for (int li = MC_L0(c); li < MC_L1(c, p->num_of_list_to_search); li++)
        for (int ri = MC_R0(c); ri < MC_R1(c, p->num_of_ref_pic_to_search[li]); ri++) {
            if (which && (which == 1) != (!li && !ri))
                continue;
            if (p->me_early_exit_th && s->zz_sad[li][ri] < (p->me_early_exit_th >> 2)) {
                mc_fill_quadrants(s->l0x[li][ri], s->l0y[li][ri], s->l0sad[li][ri], 0, 0, 0);
                continue;
            }
            if (p->prev_me_stage_based_exit_th) {
                int sr = s->phme[li][ri][0].sad <= s->phme[li][ri][1].sad ? 0 : 1;
                if (s->performed_phme[li][ri][sr] && s->phme[li][ri][sr].sad < (p->prev_me_stage_based_exit_th >> 4)) {
                    mc_fill_quadrants(s->l0x[li][ri], s->l0y[li][ri], s->l0sad[li][ri], s->phme[li][ri][sr].col,
                                      s->phme[li][ri][sr].row, s->phme[li][ri][sr].sad);
                    continue;
                }
            }
            if (!s->do_ref[li][ri]) {
                mc_fill_quadrants(s->l0x[li][ri], s->l0y[li][ri], s->l0sad[li][ri], 0, 0, ME_MAX_U32);
                continue;
            }
            if (!mc_ref_active(c, li))
                continue;
            int16_t sa_w = 0, sa_h = 0;
            mc_hme_l0_area(c, li, ri, mc_ref_distance(c, li, ri), &sa_w, &sa_h);
            for (int sr_h = 0; sr_h < p->num_hme_sa_h; sr_h++)
                for (int sr_w = 0; sr_w < p->num_hme_sa_w; sr_w++) {
                    int16_t w  = (int16_t)((sa_w + 7) & ~0x07), h = sa_h;
                    int16_t x0 = (int16_t)(-(int16_t)((w * p->num_hme_sa_w) >> 1) + (int16_t)(w * sr_w));
                    int16_t y0 = (int16_t)(-(int16_t)((h * p->num_hme_sa_h) >> 1) + (int16_t)(h * sr_h));
                    mc_clamp_axis(ox, &x0, &w, pl->pad - 1, pl->width);
                    w = (int16_t)((w < 8) ? w : w & ~0x07);
                    mc_clamp_axis(oy, &y0, &h, pl->pad - 1, pl->height);
                    mc_emit(c, (li * ME_NR + ri) * 4 + sr_w * 2 + sr_h, 2, li, ri, ox, oy, x0, y0, c->b64_w >> 2,
                            c->b64_h >> 2, p->hme_search_method == ME_GPU_SUB_SAD_SEARCH, 0, w, h);
                }
        }
}

/* ME:1872-1901 get_worst_quadrant: the last comparison does not update max_sad (kept as written) */
ME_HD void mc_worst_quadrant(const B64State *s, int li, int ri, uint8_t *w, uint8_t *h) {
    uint32_t m = 0;
    if (s->l0sad[li][ri][0][0] > m) { m = s->l0sad[li][ri][0][0]; *w = 0; *h = 0; }
    if (s->l0sad[li][ri][1][0] > m) { m = s->l0sad[li][ri][1][0]; *w = 1; *h = 0; }
    if (s->l0sad[li][ri][0][1] > m) { m = s->l0sad[li][ri][0][1]; *w = 0; *h = 1; }
    if (s->l0sad[li][ri][1][1] > m) { *w = 1; *h = 1; }
}

ME_HD void mc_l0_consume(CtrlCtx *c) {
    const MeGpuParams *p = c->p;
    B64State          *s = c->s;
    for (int li = MC_L0(c); li < MC_L1(c, ME_NL); li++)
        for (int ri = MC_R0(c); ri < MC_R1(c, ME_NR); ri++) {
            const int base = (li * ME_NR + ri) * 4;
            if (!c->t[base].valid)
                continue;
            for (int sr_w = 0; sr_w < 2; sr_w++)
                for (int sr_h = 0; sr_h < 2; sr_h++) {
                    const int k = base + sr_w * 2 + sr_h;
                    int16_t  *x = &s->l0x[li][ri][sr_w][sr_h], *y = &s->l0y[li][ri][sr_w][sr_h];
                    s->l0sad[li][ri][sr_w][sr_h] = mc_result(c, k, x, y);
                    *x = (int16_t)((*x + c->t[k].x0) * 4);
                    *y = (int16_t)((*y + c->t[k].y0) * 4);
                }
            if (p->prehme_enable) {
                uint8_t wq = 0, hq = 0;
                mc_worst_quadrant(s, li, ri, &wq, &hq);
                int sr = s->phme[li][ri][0].sad <= s->phme[li][ri][1].sad ? 0 : 1;
                if (s->phme[li][ri][sr].sad < s->l0sad[li][ri][wq][hq]) {
                    s->l0sad[li][ri][wq][hq] = s->phme[li][ri][sr].sad;
                    s->l0x[li][ri][wq][hq]   = s->phme[li][ri][sr].col;
                    s->l0y[li][ri][wq][hq]   = s->phme[li][ri][sr].row;
                }
            }
        }
}

/* ME:923-1022 hme_level_1 / ME:1025-1113 hme_level_2: refinement window around the previous centre */
ME_HD void mc_refine_emit(CtrlCtx *c, int level, int slot, int li, int ri, int16_t sa_w, int16_t sa_h, int16_t cx,
                          int16_t cy) {
    const MePlaneGeom *pl   = &c->j->g.plane[level == 1 ? 1 : 0];
    const int          sh   = level == 1 ? 1 : 0;
    const int          ox   = (int16_t)c->org_x >> sh, oy = (int16_t)c->org_y >> sh;
    const int          padx = level == 1 ? pl->pad - 1 : 63, pady = padx;
    sa_w                    = (int16_t)((sa_w + 7) & ~0x07);
    int16_t x0 = (int16_t)(-(sa_w >> 1) + cx), y0 = (int16_t)(-(sa_h >> 1) + cy);
    mc_clamp_axis(ox, &x0, &sa_w, padx, pl->width);
    sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
    mc_clamp_axis(oy, &y0, &sa_h, pady, pl->height);
    mc_emit(c, slot, level == 1 ? 1 : 0, li, ri, ox, oy, x0, y0, c->b64_w >> sh, c->b64_h >> sh,
            c->p->hme_search_method == ME_GPU_SUB_SAD_SEARCH, 0, sa_w, sa_h);
}

/* ME:2041-2122 hme_level1_b64 */
ME_HD void mc_l1_emit(CtrlCtx *c) {
    const MeGpuParams *p = c->p;
    B64State          *s = c->s;
    for (int li = MC_L0(c); li < MC_L1(c, p->num_of_list_to_search); li++)
        for (int ri = MC_R0(c); ri < MC_R1(c, p->num_of_ref_pic_to_search[li]); ri++) {
            if (!mc_ref_active(c, li))
                continue;
            if (p->me_early_exit_th && s->zz_sad[li][ri] < (p->me_early_exit_th >> 2)) {
                mc_fill_quadrants(s->l1x[li][ri], s->l1y[li][ri], s->l1sad[li][ri], 0, 0, 0);
                continue;
            }
            if (!s->do_ref[li][ri]) {
                mc_fill_quadrants(s->l1x[li][ri], s->l1y[li][ri], s->l1sad[li][ri], 0, 0, ME_MAX_U32);
                continue;
            }
            for (int sr_h = 0; sr_h < p->num_hme_sa_h; sr_h++)
                for (int sr_w = 0; sr_w < p->num_hme_sa_w; sr_w++) {
                    if (p->prev_me_stage_based_exit_th &&
                        s->l0sad[li][ri][sr_w][sr_h] < (p->prev_me_stage_based_exit_th >> 5)) {
                        s->l1x[li][ri][sr_w][sr_h]   = s->l0x[li][ri][sr_w][sr_h];
                        s->l1y[li][ri][sr_w][sr_h]   = s->l0y[li][ri][sr_w][sr_h];
                        s->l1sad[li][ri][sr_w][sr_h] = s->l0sad[li][ri][sr_w][sr_h];
                        continue;
                    }
                    mc_refine_emit(c, 1, (li * ME_NR + ri) * 4 + sr_w * 2 + sr_h, li, ri, (int16_t)p->hme_l1_sa.width,
                                   (int16_t)p->hme_l1_sa.height, s->l0x[li][ri][sr_w][sr_h] >> 1,
                                   s->l0y[li][ri][sr_w][sr_h] >> 1);
                }
        }
}

/* ME:2127-2177 hme_level2_b64 (no do_ref / zz early-outs at this level, as in the reference) */
ME_HD void mc_l2_emit(CtrlCtx *c) {
    const MeGpuParams *p = c->p;
    B64State          *s = c->s;
    for (int li = MC_L0(c); li < MC_L1(c, p->num_of_list_to_search); li++)
        for (int ri = MC_R0(c); ri < MC_R1(c, p->num_of_ref_pic_to_search[li]); ri++) {
            if (!mc_ref_active(c, li))
                continue;
            for (int sr_h = 0; sr_h < p->num_hme_sa_h; sr_h++)
                for (int sr_w = 0; sr_w < p->num_hme_sa_w; sr_w++) {
                    if (p->prev_me_stage_based_exit_th &&
                        s->l1sad[li][ri][sr_w][sr_h] < (p->prev_me_stage_based_exit_th >> 2)) {
                        s->l2x[li][ri][sr_w][sr_h]   = s->l1x[li][ri][sr_w][sr_h];
                        s->l2y[li][ri][sr_w][sr_h]   = s->l1y[li][ri][sr_w][sr_h];
                        s->l2sad[li][ri][sr_w][sr_h] = s->l1sad[li][ri][sr_w][sr_h];
                        continue;
                    }
                    mc_refine_emit(c, 2, (li * ME_NR + ri) * 4 + sr_w * 2 + sr_h, li, ri, (int16_t)p->hme_l2_sa.width,
                                   (int16_t)p->hme_l2_sa.height, s->l1x[li][ri][sr_w][sr_h], s->l1y[li][ri][sr_w][sr_h]);
                }
        }
}

ME_HD void mc_refine_consume(CtrlCtx *c, int level) {
    B64State *s = c->s;
    for (int k = MC_K0(c, 4); k < MC_K1(c, 4); k++) {
        if (!c->t[k].valid)
            continue;
        const int li = (k >> 2) / ME_NR, ri = (k >> 2) % ME_NR, sr_w = (k >> 1) & 1, sr_h = k & 1;
        int16_t  *x   = level == 1 ? &s->l1x[li][ri][sr_w][sr_h] : &s->l2x[li][ri][sr_w][sr_h];
        int16_t  *y   = level == 1 ? &s->l1y[li][ri][sr_w][sr_h] : &s->l2y[li][ri][sr_w][sr_h];
        uint32_t *sad = level == 1 ? &s->l1sad[li][ri][sr_w][sr_h] : &s->l2sad[li][ri][sr_w][sr_h];
        *sad          = mc_result(c, k, x, y);
        *x            = (int16_t)((*x + c->t[k].x0) * (level == 1 ? 2 : 1));
        *y            = (int16_t)((*y + c->t[k].y0) * (level == 1 ? 2 : 1));
    }
}

/* ME:2182-2380 set_final_seach_centre_sb: best quadrant of the deepest enabled level, scanned
 * [1][0],[0][1],[1][1] after [0][0] with strict '<'. The running (centre, sad) carries over from the
 * previous reference when a branch does not assign it (ME:2361-2371). */
ME_HD void mc_final_centre(CtrlCtx *c) {
    const MeGpuParams *p = c->p;
    B64State          *s = c->s;
    int16_t  hx = 0, hy = 0, cx = 0, cy = 0;
    uint32_t hsad = 0;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (mc_ref_active(c, li)) {
                if (p->enable_hme_flag) {
                    const int16_t(*X)[2]  = 0;
                    const int16_t(*Y)[2]  = 0;
                    const uint32_t(*S)[2] = 0;
                    if (p->enable_hme_level0_flag && !p->enable_hme_level1_flag && !p->enable_hme_level2_flag) {
                        X = s->l0x[li][ri]; Y = s->l0y[li][ri]; S = s->l0sad[li][ri];
                    }
                    if (p->enable_hme_level1_flag && !p->enable_hme_level2_flag) {
                        X = s->l1x[li][ri]; Y = s->l1y[li][ri]; S = s->l1sad[li][ri];
                    }
                    if (p->enable_hme_level2_flag) {
                        X = s->l2x[li][ri]; Y = s->l2y[li][ri]; S = s->l2sad[li][ri];
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
            s->sc_x[li][ri]    = cx;
            s->sc_y[li][ri]    = cy;
            s->hme_sad[li][ri] = hsad;
        }
}

/* ME:2477-2518 hme_prune_ref_and_adjust_sr: loops run over all 2x4 slots, used or not */
ME_HD void mc_hme_prune(CtrlCtx *c) {
    const MeGpuParams *p  = c->p;
    B64State          *s  = c->s;
    uint16_t           th = p->prune_ref_if_hme_sad_dev_bigger_than_th;
    if (p->enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        uint64_t best = ~(uint64_t)0;
        for (int i = 0; i < ME_NL; i++)
            for (int k = 0; k < ME_NR; k++) best = MC_MIN(best, (uint64_t)s->hme_sad[i][k]);
        for (int li = 0; li < ME_NL; li++)
            for (int ri = 1; ri < ME_NR; ri++)
                if (((uint64_t)s->hme_sad[li][ri] - best) * 100 > (uint64_t)th * best)
                    s->do_ref[li][ri] = 0;
    }
    if (p->enable_me_sr_adjustment)
        for (int li = 0; li < ME_NL; li++)
            for (int ri = 0; ri < ME_NR; ri++) {
                if (MC_ABS(s->sc_x[li][ri]) <= p->reduce_me_sr_based_on_mv_length_th &&
                    MC_ABS(s->sc_y[li][ri]) <= p->reduce_me_sr_based_on_mv_length_th &&
                    s->hme_sad[li][ri] < p->stationary_hme_sad_abs_th)
                    s->sr_div[li][ri] = p->stationary_me_sr_divisor;
                else if (s->hme_sad[li][ri] < p->reduce_me_sr_based_on_hme_sad_abs_th)
                    s->sr_div[li][ri] = p->me_sr_divisor_for_low_hme_sad;
            }
}

/* ME:1249-1330 integer_search_b64, first part: search-area size per reference, then either the zz early
 * exit or the two SADs of check_00_center (ME:1139-1206) as tasks */
ME_HD void mc_centre_emit(CtrlCtx *c) {
    const MeGpuParams *p  = c->p;
    B64State          *s  = c->s;
    const MePlaneGeom *pl = &c->j->g.plane[0];
    const int16_t      ox = (int16_t)c->org_x, oy = (int16_t)c->org_y;
    for (int li = MC_L0(c); li < MC_L1(c, p->num_of_list_to_search); li++)
        for (int ri = MC_R0(c); ri < MC_R1(c, p->num_of_ref_pic_to_search[li]); ri++) {
            if (!s->do_ref[li][ri])
                continue;
            int16_t  cx = s->sc_x[li][ri], cy = s->sc_y[li][ri];
            uint16_t dist = mc_ref_distance(c, li, ri);
            if (p->me_type != ME_GPU_ME_MCTF) /* ME:1300-1302 */
                dist = mc_scaled_distance(dist);
            int16_t  w = (int16_t)MC_MIN(p->me_sa.sa_min.width * dist, p->me_sa.sa_max.width);
            int16_t  h = (int16_t)MC_MIN(p->me_sa.sa_min.height * dist, p->me_sa.sa_max.height);
            if (p->mv_sa_adj_enabled && (!p->mv_sa_adj_nearest_ref_only || ri == 0)) {
                if (MC_ABS(cx) > p->mv_sa_adj_mv_size_th) w = (int16_t)(w * p->mv_sa_adj_sa_multiplier);
                if (MC_ABS(cy) > p->mv_sa_adj_mv_size_th) h = (int16_t)(h * p->mv_sa_adj_sa_multiplier);
            }
            /* int16 operands against a uint32 divisor: the division is unsigned (ME:1314-1318) */
            w = (int16_t)((MC_MAX(1u, ((uint32_t)w / s->sr_div[li][ri])) + 7) & ~0x07u);
            h = (int16_t)MC_MAX(3u, ((uint32_t)h / s->sr_div[li][ri]));
            if (p->me_early_exit_th) {
                if (s->zz_sad[li][ri] < p->me_early_exit_th / 6)
                    w = h = 1;
            } else if ((cx != 0 || cy != 0) && p->is_ref) {
                /* check_00_center: clamp the centre, then SAD at (0,0) and at the centre */
                if (ox + cx < -63) cx = (int16_t)(-63 - ox);
                if (ox + cx > pl->width - 1) cx = (int16_t)(cx - ((ox + cx) - (pl->width - 1)));
                if (oy + cy < -63) cy = (int16_t)(-63 - oy);
                if (oy + cy > pl->height - 1) cy = (int16_t)(cy - ((oy + cy) - (pl->height - 1)));
                const int k = (li * ME_NR + ri) * 2;
                mc_emit(c, k, 0, li, ri, ox, oy, 0, 0, c->b64_w, c->b64_h, 1, 0, 1, 1);
                mc_emit(c, k + 1, 0, li, ri, ox, oy, cx, cy, c->b64_w, c->b64_h, 1, 0, 1, 1);
                s->pend00[li][ri] = 1;
            }
            s->cx[li][ri] = cx;
            s->cy[li][ri] = cy;
            s->sw[li][ri] = w;
            s->sh[li][ri] = h;
        }
}

ME_HD void mc_fs_emit(CtrlCtx *c, int li, int ri, int16_t x0, int16_t y0, int16_t w, int16_t h, int probe) {
    FsTask *f   = &c->f[li * ME_NR + ri];
    f->valid    = 1;
    f->probe    = (uint8_t)probe;
    f->ref_slot = c->j->ref_slot[li][ri];
    f->x0       = x0;
    f->y0       = y0;
    f->w        = w;
    f->h        = h;
    const uint64_t n = (uint64_t)w * h * (c->p->me_search_method == ME_GPU_SUB_SAD_SEARCH ? 64 * 32 : 64 * 64);
    MC_OPS_ADD(&c->s->ops, n);
    MC_OPS_ADD(&c->s->ops_fs, n);
}

/* ME:1441-1510: centre the area, clamp to the picture +-63, emit the main search */
ME_HD void mc_main_emit(CtrlCtx *c, int li, int ri, int16_t w, int16_t h) {
    B64State *s  = c->s;
    int16_t   x0 = (int16_t)(s->cx[li][ri] - (w >> 1)), y0 = (int16_t)(s->cy[li][ri] - (h >> 1));
    mc_clamp_axis((int16_t)c->org_x, &x0, &w, 63, c->pic_w);
    w = (int16_t)((w < 8) ? w : w & ~0x07);
    mc_clamp_axis((int16_t)c->org_y, &y0, &h, 63, c->pic_h);
    s->sa_final[li][ri][0] = x0;
    s->sa_final[li][ri][1] = y0;
    s->sa_final[li][ri][2] = w;
    s->sa_final[li][ri][3] = h;
    s->searched[li][ri]    = 1;
    mc_fs_emit(c, li, ri, x0, y0, w, h, 0);
}

/* ME:1330-1440: check_00_center decision, then the 8x8-variance probe where it applies */
ME_HD void mc_probe_emit(CtrlCtx *c) {
    const MeGpuParams *p = c->p;
    B64State          *s = c->s;
    for (int k = MC_K0(c, 1); k < MC_K1(c, 1); k++) c->f[k].valid = 0;
    for (int li = MC_L0(c); li < MC_L1(c, p->num_of_list_to_search); li++)
        for (int ri = MC_R0(c); ri < MC_R1(c, p->num_of_ref_pic_to_search[li]); ri++) {
            if (!s->do_ref[li][ri] || !MC_FS_MINE(c, li, ri))
                continue;
            int16_t  w = s->sw[li][ri], h = s->sh[li][ri];
            int      hme_accurate = 1;
            uint64_t best_hme_sad = ~(uint64_t)0;
            if (s->pend00[li][ri]) {
                const int k  = (li * ME_NR + ri) * 2;
                int16_t   dx = 0, dy = 0;
                /* zero_sad = nxm_sad then <<= 1; hme_sad = nxm_sad << 1: both are the doubled task results.
                 * COST_PRECISION scales both sides alike (ME:1188-1204). */
                uint32_t zero_sad = mc_result(c, k, &dx, &dy);
                uint32_t hme_sad  = mc_result(c, k + 1, &dx, &dy);
                if ((uint64_t)zero_sad <= (uint64_t)hme_sad)
                    s->cx[li][ri] = s->cy[li][ri] = 0;
                s->pend00[li][ri] = 0;
                best_hme_sad      = hme_sad; /* check_00_center returns the SAD at the HME centre (ME:1205) */
                if (s->cx[li][ri] == 0 && s->cy[li][ri] == 0)
                    hme_accurate = 0;
            }
            if (p->enable_me_sr_adjustment == 2 && !p->me_early_exit_th) { /* ME:1349-1364 */
                const int16_t w_before = w, h_before = h;
                if ((hme_accurate && best_hme_sad < 24 * 24) || (p->is_ref && s->hme_sad[li][ri] < 24 * 24))
                    h = (int16_t)(h / 2);
                /* p_sb_best_sad[0][0][0]: the 64x64 result of reference (0, 0), whose whole search precedes this one */
                if ((li || ri) && (uint32_t)(MC_KEY_LOAD(&c->keys[0]) >> 32) < 5000 && h == h_before && w == w_before) {
                    h = (int16_t)(h >> 1);
                    w = (int16_t)(w >> 1);
                }
                s->sw[li][ri] = w;
                s->sh[li][ri] = h;
            }
            if (p->me_8x8_var_enabled && (w * h > 24)) {
                s->probed[li][ri]   = 1;
                s->searched[li][ri] = 1;
                mc_fs_emit(c, li, ri, s->cx[li][ri], s->cy[li][ri], 1, 1, 1);
            } else
                mc_main_emit(c, li, ri, w, h);
        }
}

/* ME:1391-1439: variance of the 64 8x8 SADs of the centre position scales the area */
ME_HD void mc_main_emit_after_probe(CtrlCtx *c) {
    const MeGpuParams *p = c->p;
    B64State          *s = c->s;
    for (int k = MC_K0(c, 1); k < MC_K1(c, 1); k++) c->f[k].valid = 0;
    for (int li = MC_L0(c); li < MC_L1(c, p->num_of_list_to_search); li++)
        for (int ri = MC_R0(c); ri < MC_R1(c, p->num_of_ref_pic_to_search[li]); ri++) {
            if (!s->do_ref[li][ri] || !s->probed[li][ri] || !MC_FS_MINE(c, li, ri))
                continue;
            const unsigned long long *key = &c->keys[(li * ME_NR + ri) * ME_NPU];
            const uint32_t            mean = (uint32_t)(MC_KEY_LOAD(&key[0]) >> 32) / 64;
            uint32_t                  ssq  = 0;
            for (int i = 0; i < 64; i++) {
                int32_t d = (int32_t)(uint32_t)(MC_KEY_LOAD(&key[21 + i]) >> 32) - (int32_t)mean;
                ssq += (uint32_t)(d * d);
            }
            const uint32_t var = ssq / 64;
            int16_t        w = s->sw[li][ri], h = s->sh[li][ri];
            if (var > p->me_sr_mult2_th) {
                w = (int16_t)((MC_MAX(1, w * 3 / 2) + 7) & ~0x7);
                h = (int16_t)MC_MAX(1, h * 3 / 2);
            }
            if (var < p->me_sr_div4_th) {
                w = (int16_t)((MC_MAX(1, w >> 2) + 7) & ~0x7);
                h = (int16_t)MC_MAX(1, h >> 2);
                h = (int16_t)MC_MAX(3, h);
            } else if (var < p->me_sr_div2_th) {
                w = (int16_t)((MC_MIN(w, w >> 1) + 7) & ~0x7);
                h = (int16_t)MC_MIN(h, h >> 1);
                h = (int16_t)MC_MAX(3, h);
            }
            mc_main_emit(c, li, ri, w, h);
        }
}

/* consume the results of the previous stage's tasks. The zero-MV decision looks across references: lane 0 alone. */
ME_HD void mc_consume_pending(CtrlCtx *c) {
    const int pending = c->s->pending;
    MC_SYNC(c); /* every lane has read `pending` before lane 0 resets it */
    switch (pending) {
    case PEND_ZZ:
        if (c->lane == 0) {
            const int only = c->only;
            c->only        = -1;
            mc_zz_consume(c);
            c->only = only;
        }
        break;
    case PEND_PHME: mc_prehme_consume_tasks(c); break;
    case PEND_L0: mc_l0_consume(c); break;
    case PEND_L1: mc_refine_consume(c, 1); break;
    case PEND_L2: mc_refine_consume(c, 2); break;
    default: break;
    }
    if (c->lane == 0)
        c->s->pending = PEND_NONE;
    MC_SYNC(c);
}

/* a step that looks across references: lane 0 alone, over all references */
#define MC_SERIAL(c, call)         \
    do {                           \
        if ((c)->lane == 0) {      \
            const int only_ = (c)->only; \
            (c)->only = -1;        \
            call;                  \
            (c)->only = only_;     \
        }                          \
        MC_SYNC(c);                \
    } while (0)

/* one control stage of one b64 (see CtrlCtx: run by one thread, or by the lanes of a warp with a barrier per step) */
ME_HD void mc_run_stage(CtrlCtx *c, int stage) {
    const MeGpuParams *p = c->p;
    switch (stage) {
    case ST_INIT:
        mc_stage_init(c);
        mc_clear_tasks(c);
        MC_SYNC(c);
        mc_zz_emit(c);
        break;
    case ST_PREHME0:
        mc_consume_pending(c);
        mc_clear_tasks(c);
        MC_SYNC(c);
        if (p->prehme_enable) {
            mc_prehme_emit_list(c, 0);
            if (!p->prehme_l1_early_exit)
                mc_prehme_emit_list(c, 1);
            if (c->lane == 0)
                c->s->pending = PEND_PHME;
        }
        break;
    case ST_PREHME1:
        mc_consume_pending(c);
        mc_clear_tasks(c);
        MC_SYNC(c);
        if (p->prehme_enable && p->prehme_l1_early_exit) {
            mc_prehme_emit_list(c, 1);
            if (c->lane == 0)
                c->s->pending = PEND_PHME;
        }
        break;
    case ST_HME_L0:
        mc_consume_pending(c);
        if (p->prehme_enable)
            MC_SERIAL(c, mc_prehme_finish(c));
        mc_clear_tasks(c);
        MC_SYNC(c);
        if (p->enable_hme_flag && p->enable_hme_level0_flag) {
            mc_l0_emit(c, mc_l0_two_phase(p) ? 1 : 0);
            if (c->lane == 0)
                c->s->pending = PEND_L0;
        }
        break;
    case ST_HME_L0B:
        if (p->enable_hme_flag && p->enable_hme_level0_flag && mc_l0_two_phase(p)) {
            mc_consume_pending(c); /* (0, 0)'s level 0 centres, pre-HME replacement included */
            mc_clear_tasks(c);
            MC_SYNC(c);
            mc_l0_emit(c, 2);
            if (c->lane == 0)
                c->s->pending = PEND_L0;
        }
        break;
    case ST_HME_L1:
        mc_consume_pending(c);
        mc_clear_tasks(c);
        MC_SYNC(c);
        if (p->enable_hme_flag && p->enable_hme_level1_flag) {
            mc_l1_emit(c);
            if (c->lane == 0)
                c->s->pending = PEND_L1;
        }
        break;
    case ST_HME_L2:
        mc_consume_pending(c);
        mc_clear_tasks(c);
        MC_SYNC(c);
        if (p->enable_hme_flag && p->enable_hme_level2_flag) {
            mc_l2_emit(c);
            if (c->lane == 0)
                c->s->pending = PEND_L2;
        }
        break;
    case ST_CENTRE:
        mc_consume_pending(c);
        mc_clear_tasks(c);
        MC_SERIAL(c, mc_final_centre(c));
        if (p->me_type == ME_GPU_ME_MCTF) { /* ME:3109-3113: a low HME SAD ends the block before the integer search */
            if (c->lane == 0)
                c->s->tf_exit = c->s->hme_sad[0][0] < p->tf_me_exit_th;
            MC_SYNC(c);
        } else if (p->enable_hme_flag) /* prune_ref = enable_hme_flag && me_type != ME_MCTF, ME:3103 */
            MC_SERIAL(c, mc_hme_prune(c));
        if (!c->s->tf_exit)
            mc_centre_emit(c);
        break;
    case ST_PROBE:
        if (!c->s->tf_exit)
            mc_probe_emit(c);
        break;
    case ST_MAIN:
        if (!c->s->tf_exit)
            mc_main_emit_after_probe(c);
        break;
    default: break;
    }
    MC_SYNC(c);
}

/* ------------------------------------------------------------------------------------------------
 * epilogue: p_sb_best_* reconstruction, me_prune_ref, candidates, distortion, GM detection
 * ---------------------------------------------------------------------------------------------- */

typedef struct EpiCtx {
    const MeJob       *j;
    const MeGpuParams *p;
    B64State          *s;
    const unsigned long long *keys;
    uint32_t           b; /* b64 index */
    uint32_t           b64_w, b64_h;
    uint32_t (*best_sad)[ME_NPU]; /* [ME_NREF][ME_NPU] scratch (shared memory on the device) */
    uint32_t (*best_mv)[ME_NPU];
    uint32_t *me_distortion; /* [ME_NPU] */
} EpiCtx;

/* key -> (sad, mv) of PU n of (list, ref) slot k. Scan order 0 is the probe position (the search centre),
 * order 1 + y * w + x a position of the main search area (ME:448-450 for the MV packing). */
ME_HD void me_epi_decode(EpiCtx *e, int k, int n) {
    const B64State *s = e->s;
    const int       li = k / ME_NR, ri = k % ME_NR;
    uint32_t        sad = ME_MAX_SAD_VALUE, mv = 0;
    if (s->searched[li][ri]) {
        const unsigned long long key = e->keys[k * ME_NPU + n];
        const uint32_t           ord = (uint32_t)key;
        sad                          = (uint32_t)(key >> 32);
        int16_t x, y;
        if (ord == 0) {
            x = s->cx[li][ri];
            y = s->cy[li][ri];
        } else {
            const uint32_t idx = ord - 1, w = (uint32_t)s->sa_final[li][ri][2];
            x = (int16_t)(s->sa_final[li][ri][0] + (int)(idx % w));
            y = (int16_t)(s->sa_final[li][ri][1] + (int)(idx / w));
        }
        mv = ((uint32_t)(uint16_t)y << 16) | (uint16_t)x;
    }
    e->best_sad[k][n] = sad;
    e->best_mv[k][n]  = mv;
}

/* ME:1522-1565 me_prune_ref, first half: hme_sad = sum of the 64 best 8x8 SADs (the device sums them with a warp per
 * reference, me_kernels.cu) */
ME_HD void me_epi_prune_sums(EpiCtx *e) {
    const MeGpuParams *p = e->p;
    B64State          *s = e->s;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            if (!s->do_ref[li][ri]) {
                s->hme_sad[li][ri] = ME_MAX_SAD_VALUE * 64u;
                continue;
            }
            uint32_t v = 0;
            for (int k = 0; k < 64; k++) v += e->best_sad[li * ME_NR + ri][21 + k];
            s->hme_sad[li][ri] = v;
        }
}

/* second half: the deviation test over all 2x4 slots */
ME_HD void me_epi_prune_decide(EpiCtx *e) {
    const MeGpuParams *p = e->p;
    B64State          *s = e->s;
    uint16_t th = p->prune_ref_if_me_sad_dev_bigger_than_th;
    if (p->enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        uint64_t best = ~(uint64_t)0;
        for (int i = 0; i < ME_NL; i++)
            for (int k = 0; k < ME_NR; k++) best = MC_MIN(best, (uint64_t)s->hme_sad[i][k]);
        for (int li = 0; li < ME_NL; li++)
            for (int ri = 1; ri < ME_NR; ri++)
                if (((uint64_t)s->hme_sad[li][ri] - best) * 100 > (uint64_t)th * best)
                    s->do_ref[li][ri] = 0;
    }
}

ME_HD void me_epi_prune(EpiCtx *e) {
    me_epi_prune_sums(e);
    me_epi_prune_decide(e);
}

/* z-order PU index (MeContext order) -> raster PU index (output order), ME:2520-2531 */
ME_HD uint8_t me_z_to_raster(int n) {
    if (n < 5)
        return (uint8_t)n;
    int base = n < 21 ? 5 : 21, bits = n < 21 ? 2 : 3, z = n - base, bx = 0, by = 0;
    for (int b = 0; b < bits; b++) {
        bx |= ((z >> (2 * b)) & 1) << b;
        by |= ((z >> (2 * b + 1)) & 1) << b;
    }
    return (uint8_t)(base + by * (1 << bits) + bx);
}

/* MeCandidate bitfield byte (Codec/me_sb_results.h:28-34): direction:2, ref_idx_l0:2, ref_idx_l1:2,
 * ref0_list:1, ref1_list:1 from the least significant bit (gcc, x86-64). The reference stores 24 in the
 * 1-bit list field of the unused side, which truncates to 0. */
ME_HD uint8_t me_cand_byte(int direction, int ref_idx_l0, int ref_idx_l1, int ref0_list, int ref1_list) {
    return (uint8_t)((direction & 3) | ((ref_idx_l0 & 3) << 2) | ((ref_idx_l1 & 3) << 4) | ((ref0_list & 1) << 6) |
                     ((ref1_list & 1) << 7));
}

ME_HD int me_pu_exported(const MeGpuParams *p, int n_idx) { /* use_me_pu, ME:2704-2706 */
    return p->enable_me_16x16 ? (p->enable_me_8x8 || n_idx < 21) : n_idx < 5;
}
ME_HD int me_exported_pu_count(const MeGpuParams *p) { /* ME:2552-2554 */
    return p->enable_me_16x16 ? (p->enable_me_8x8 ? p->max_number_of_pus_per_sb : 21) : 5;
}

/* candidate construction of PU n_idx (every PU is independent of the others). `mode`: 0 single reference
 * (ME:2646-2697), 1 one reference per list (ME:2532-2645), 2 general (ME:2698-2835). */
ME_HD void me_epi_candidates(EpiCtx *e, int n) {
    const MeGpuParams *p  = e->p;
    const B64State    *s  = e->s;
    const uint32_t     P  = e->j->pu_count;
    uint8_t  *total = e->j->total_me_candidate_index + (size_t)e->b * P;
    uint8_t  *cands = e->j->me_candidate_array + (size_t)e->b * P * p->max_cand;
    uint32_t *mvs   = e->j->me_mv_array + (size_t)e->b * P * p->max_refs;
    const int use   = me_pu_exported(p, n);
    const uint8_t *nref = p->num_of_ref_pic_to_search;
    if (nref[0] == 1 && nref[1] == 0) {
        const int pu         = me_z_to_raster(n);
        e->me_distortion[pu] = e->best_sad[0][n];
        if (use)
            total[pu] = 1;
        if (!s->do_ref[0][0] || !use)
            return;
        cands[pu * p->max_cand] = me_cand_byte(0, 0, 0, 0, 0);
        mvs[pu * p->max_refs]   = e->best_mv[0][n];
    } else if (nref[0] == 1 && nref[1] == 1) {
        uint32_t lists = p->num_of_list_to_search;
        uint8_t  org[2];
        org[0] = s->do_ref[0][0];
        org[1] = (lists == 1) ? 0 : s->do_ref[1][0];
        if (lists < 2 || !s->do_ref[1][0])
            lists = 1;
        const uint32_t prune_th = (org[0] && org[1]) ? (uint32_t)p->prune_me_candidates_th : 0;
        const int      pu = me_z_to_raster(n);
        int            off = 0;
        uint8_t        blk[2] = {org[0], org[1]};
        const uint32_t s0 = e->best_sad[0][n], s1 = e->best_sad[ME_NR][n];
        const uint32_t best = (org[0] && org[1]) ? MC_MIN(s0, s1) : org[0] ? s0 : s1;
        e->me_distortion[pu] = best;
        if (use)
            total[pu] = 1;
        int min_list = -1;
        if (p->use_best_unipred_cand_only && blk[0] && blk[1])
            min_list = s0 < s1 ? 0 : 1;
        for (uint32_t li = 0; li < lists && (use || off == 0); li++) {
            if (!blk[li])
                continue;
            if (prune_th > 0) {
                uint32_t d = (e->best_sad[li * ME_NR][n] - best) * 100;
                if (d > best * prune_th) {
                    blk[li] = 0;
                    continue;
                }
            }
            if (min_list != -1 && min_list != (int)li) {
                if (use)
                    mvs[pu * p->max_refs + (li ? p->max_l0 : 0)] = e->best_mv[li * ME_NR][n];
                continue;
            }
            if (use) {
                cands[pu * p->max_cand + off] = me_cand_byte((int)li, 0, 0, li == 0 ? (int)li : 24, li == 1 ? (int)li : 24);
                mvs[pu * p->max_refs + (li ? p->max_l0 : 0)] = e->best_mv[li * ME_NR][n];
            }
            off++;
        }
        if (blk[0] && blk[1] && use) {
            cands[pu * p->max_cand + off] = me_cand_byte(2 /*BI_PRED*/, 0, 0, 0, 1);
            total[pu]                     = (uint8_t)(off + 1);
        }
    } else {
        const uint32_t lists = p->num_of_list_to_search;
        const int      pu    = (n > 4) ? me_z_to_raster(n) : n;
        int            off   = 0;
        uint8_t        blk[ME_NL][ME_NR];
        const uint32_t prune_th = (uint32_t)p->prune_me_candidates_th;
        uint32_t       best     = ME_MAX_U32;
        for (int a = 0; a < ME_NL; a++)
            for (int b = 0; b < ME_NR; b++) blk[a][b] = 0;
        for (uint32_t li = 0; li < lists; li++)
            for (uint32_t ri = 0; ri < nref[li]; ri++) {
                blk[li][ri] = s->do_ref[li][ri];
                if (blk[li][ri] && e->best_sad[li * ME_NR + ri][n] < best)
                    best = e->best_sad[li * ME_NR + ri][n];
            }
        e->me_distortion[pu] = best;
        uint8_t *cand        = use ? &cands[pu * p->max_cand] : 0;
        for (uint32_t li = 0; li < lists && (use || off == 0); li++)
            for (uint32_t ri = 0; ri < nref[li] && (use || off == 0); ri++) {
                if (!blk[li][ri])
                    continue;
                if (prune_th > 0) {
                    uint32_t d = (e->best_sad[li * ME_NR + ri][n] - best) * 100;
                    if (d > best * prune_th) {
                        blk[li][ri] = 0;
                        continue;
                    }
                }
                if (use) {
                    cand[off] = me_cand_byte((int)li, (int)ri, (int)ri, li == 0 ? (int)li : 24, li == 1 ? (int)li : 24);
                    mvs[pu * p->max_refs + (li ? p->max_l0 : 0) + ri] = e->best_mv[li * ME_NR + ri][n];
                }
                off++;
            }
        if (lists == 2 && use) {
            for (uint32_t a = 0; a < nref[0]; a++)
                for (uint32_t b = 0; b < nref[1]; b++) {
                    if (p->only_l_bwd && (a > 0 || b > 0))
                        continue;
                    if (blk[0][a] && blk[1][b])
                        cand[off++] = me_cand_byte(2, (int)a, (int)b, 0, 1);
                }
            if (!p->only_l_bwd) {
                for (uint32_t a = 1; a < nref[0]; a++)
                    if (blk[0][0] && blk[0][a])
                        cand[off++] = me_cand_byte(2, 0, (int)a, 0, 0);
                if (nref[1] == 3 && blk[1][0] && blk[1][2])
                    cand[off++] = me_cand_byte(2, 0, 2, 1, 1);
            }
        }
        if (use)
            total[pu] = (uint8_t)off;
    }
}

/* ME:2964-3007 compute_distortion */
ME_HD void me_epi_distortion(EpiCtx *e) {
    const uint32_t *d  = e->me_distortion;
    const uint32_t  nb = (uint32_t)e->j->g.b64_count, b = e->b;
    uint32_t d64 = d[0], d32 = 0, d16 = 0, d8 = 0;
    for (int i = 0; i < 4; i++) d32 += d[1 + i];
    for (int i = 0; i < 16; i++) d16 += d[5 + i];
    for (int i = 0; i < 64; i++) d8 += d[21 + i];
    uint64_t mean = d8 / 64, ssq = 0;
    for (int i = 0; i < 64; i++) {
        int64_t diff = (int64_t)d[21 + i] - (int64_t)mean;
        ssq += (uint64_t)(diff * diff);
    }
    const uint32_t pix        = e->b64_w * e->b64_h;
    e->j->dist[0 * nb + b]    = (d64 * 4096u) / pix;
    e->j->dist[1 * nb + b]    = (d32 * 4096u) / pix;
    e->j->dist[2 * nb + b]    = (d16 * 4096u) / pix;
    e->j->dist[3 * nb + b]    = (d8 * 4096u) / pix;
    e->j->dist[4 * nb + b]    = (uint32_t)(ssq / 64);
    e->j->dist[5 * nb + b]    = (e->p->input_resolution <= 2 /*INPUT_SIZE_480p_RANGE*/) ? d8 : d16;
}

/* ME:2838-2961 perform_gm_detection. As in the reference the candidate is fetched with the z-order n_idx
 * used as a raster PU index, and the MV comes from the z-order best-MV array. */
ME_HD void me_epi_gm(EpiCtx *e) {
    const MeGpuParams *p  = e->p;
    const uint32_t     nb = (uint32_t)e->j->g.b64_count, b = e->b;
    const uint32_t     P  = e->j->pu_count;
    const uint8_t     *cands = e->j->me_candidate_array + (size_t)b * P * p->max_cand;
    e->j->flags[0 * nb + b] = 0;
    e->j->flags[1 * nb + b] = 0;
    if (!p->gm_enabled)
        return;
    uint32_t stationary = 0, total = 0;
    uint32_t cnt[ME_NREF][2][2];
    for (int k = 0; k < ME_NREF; k++) cnt[k][0][0] = cnt[k][0][1] = cnt[k][1][0] = cnt[k][1][1] = 0;
    const int small = p->input_resolution <= 2;
    const int count = small ? 64 : 16;
    for (int i = 0; i < count; i++) {
        int n = small ? 21 + i : 5 + i;
        if (small) {
            /* me_idx_85_8x8_to_16x16_conversion / me_idx_16x16_to_parent_32x32_conversion
             * (Codec/definitions.h:2613-2633): raster tables, entry (row, col) -> parent at (row/2, col/2) */
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
        const uint8_t  cb  = cands[n * p->max_cand];
        const int      dir = cb & 3, l0 = (cb >> 2) & 3, l1 = (cb >> 4) & 3, r0 = (cb >> 6) & 1, r1 = (cb >> 7) & 1;
        const uint32_t li  = (dir == 0 || dir == 2) ? (uint32_t)r0 : (uint32_t)r1;
        const uint32_t ri  = (dir == 0 || dir == 2) ? (uint32_t)l0 : (uint32_t)l1;
        const uint64_t ref_num = e->j->ref_picture_number[li][ri];
        int            active_th;
        if (small) {
            uint64_t a = p->picture_number, bn = ref_num;
            int16_t  d16  = (int16_t)(MC_MAX(a, bn) - MC_MIN(a, bn));
            uint16_t dist = (uint16_t)MC_ABS(d16);
            active_th     = p->gm_use_distance_based_active_th ? MC_MAX(dist >> 1, 4) : 4;
        } else {
            int16_t  d16  = (int16_t)(p->picture_number - ref_num);
            uint16_t dist = (uint16_t)MC_ABS(d16);
            active_th     = p->gm_use_distance_based_active_th ? MC_MAX(dist * 16, 32) : 32;
        }
        const uint32_t mv = e->best_mv[li * ME_NR + ri][n];
        const int      mx = (int16_t)(mv & 0xFFFF) << 2, my = (int16_t)(mv >> 16) << 2;
        if (mx < -active_th) cnt[li * ME_NR + ri][0][0]++;
        else if (mx > active_th) cnt[li * ME_NR + ri][0][1]++;
        if (my < -active_th) cnt[li * ME_NR + ri][1][0]++;
        else if (my > active_th) cnt[li * ME_NR + ri][1][1]++;
        const int st = small ? 0 : 4;
        if (MC_ABS(mx) <= st && MC_ABS(my) <= st)
            stationary++;
        total++;
    }
    if (stationary > (total * 5) / 100)
        e->j->flags[0 * nb + b] = 1;
    for (int k = 0; k < ME_NREF; k++)
        for (int a = 0; a < 2; a++)
            for (int d = 0; d < 2; d++)
                if (cnt[k][a][d] > total / 2)
                    e->j->flags[1 * nb + b] = 1;
}

/* ME_MCTF: what the temporal filter reads after the call. Per PU n (any thread), then per block (one thread). */
ME_HD void me_epi_tf_pu(EpiCtx *e, int n) {
    const MeJob *j = e->j;
    j->tf_best_sad[(size_t)e->b * ME_NPU + n] = e->best_sad[0][n];
    j->tf_best_mv[(size_t)e->b * ME_NPU + n]  = e->best_mv[0][n];
}
ME_HD void me_epi_tf_block(EpiCtx *e) {
    const MeJob    *j = e->j;
    const B64State *s = e->s;
    j->tf_hme_centre[2 * e->b]     = s->sc_x[0][0];
    j->tf_hme_centre[2 * e->b + 1] = s->sc_y[0][0];
    j->tf_hme_sad[e->b]            = s->hme_sad[0][0];
    j->tf_early_exit[e->b]         = s->tf_exit;
    /* ME:2469-2474: the block counts as horizontal or vertical motion (exited blocks too) */
    MC_OPS_ADD(&j->ops_total[MC_ABS(s->sc_x[0][0]) > MC_ABS(s->sc_y[0][0]) ? 2 : 3], 1);
}

/* parity taps (MeGpuOutputs debug arrays) */
ME_HD void me_epi_taps(EpiCtx *e) {
    const B64State *s = e->s;
    const MeJob    *j = e->j;
    for (int k = 0; k < ME_NREF; k++) {
        const size_t q = (size_t)e->b * ME_NREF + k;
        const int    li = k / ME_NR, ri = k % ME_NR;
        if (j->dbg_centre) {
            j->dbg_centre[2 * q]     = s->sc_x[li][ri];
            j->dbg_centre[2 * q + 1] = s->sc_y[li][ri];
        }
        if (j->dbg_hme_sad) j->dbg_hme_sad[q] = s->hme_sad[li][ri];
        if (j->dbg_do_ref) j->dbg_do_ref[q] = s->do_ref[li][ri];
        if (j->dbg_area)
            for (int i = 0; i < 4; i++) j->dbg_area[4 * q + i] = s->sa_final[li][ri][i];
    }
}

#endif /* ME_CTRL_H */
