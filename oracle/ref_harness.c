/*
 * ref_harness.c - thin driver around the UNMODIFIED reference sources (compiled in place from
 * /root/reference by oracle/Makefile into oracle/_ref/libme_ref.so). TEST INFRASTRUCTURE ONLY.
 *
 * It builds the minimum PictureParentControlSet / SequenceControlSet / MeContext state that
 * svt_aom_motion_estimation_b64 (Codec/motion_estimation.c:3076) reads, calls the reference function
 * for every 64x64 block, and returns results in the MeGpuOutputs layout so that the restatement
 * (me_oracle*.c) and the CUDA engine can be compared with it byte for byte.
 *
 * No reference source is copied: this file only includes the reference headers and calls its symbols.
 */
#include <stdlib.h>
#include <string.h>

#define RTCD_C
#include "common_dsp_rtcd.h"
#include "aom_dsp_rtcd.h"
#include "pcs.h"
#include "sequence_control_set.h"
#include "motion_estimation.h"
#include "me_context.h"
#include "me_sad_calculation.h"
#include "compute_sad_c.h"
#include "pic_operators.h"
#include "enc_mode_config.h"

#include "me_oracle.h"

#define REF_API __attribute__((visibility("default")))

/* symbols the few reference translation units we link expect from the rest of the library */
void svt_aom_downsample_filtering_input_picture(PictureParentControlSet *pcs, EbPictureBufferDesc *input_padded_pic,
                                                EbPictureBufferDesc *quarter_picture_ptr,
                                                EbPictureBufferDesc *sixteenth_picture_ptr);


static int g_simd_level = -1;

/* svt_memcpy is an RTCD pointer (Codec/common_dsp_rtcd.h:1071); plain memcpy has its semantics */
static void harness_memcpy(void *dst, void const *src, size_t n) { memcpy(dst, src, n); }

/* RTCD pointers on the ME path (Codec/aom_dsp_rtcd.c:501-515). level 0 = the _c specification. */
static void set_rtcd(int level) {
    if (g_simd_level == level)
        return;
    svt_sad_loop_kernel                       = svt_sad_loop_kernel_c;
    svt_ext_all_sad_calculation_8x8_16x16     = svt_ext_all_sad_calculation_8x8_16x16_c;
    svt_ext_eight_sad_calculation_32x32_64x64 = svt_ext_eight_sad_calculation_32x32_64x64_c;
    svt_ext_sad_calculation_8x8_16x16         = svt_ext_sad_calculation_8x8_16x16_c;
    svt_ext_sad_calculation_32x32_64x64       = svt_ext_sad_calculation_32x32_64x64_c;
    svt_nxm_sad_kernel                        = svt_nxm_sad_kernel_helper_c;
    svt_initialize_buffer_32bits              = svt_initialize_buffer_32bits_c;
    downsample_2d                             = svt_aom_downsample_2d_c;
    svt_memcpy                                = harness_memcpy;
#ifdef REF_HAVE_SIMD
    if (level >= 1) {
        svt_sad_loop_kernel                       = svt_sad_loop_kernel_avx2_intrin;
        svt_ext_all_sad_calculation_8x8_16x16     = svt_ext_all_sad_calculation_8x8_16x16_avx2;
        svt_ext_eight_sad_calculation_32x32_64x64 = svt_ext_eight_sad_calculation_32x32_64x64_avx2;
        svt_ext_sad_calculation_8x8_16x16         = svt_ext_sad_calculation_8x8_16x16_avx2_intrin;
        svt_ext_sad_calculation_32x32_64x64       = svt_ext_sad_calculation_32x32_64x64_sse4_intrin;
        svt_nxm_sad_kernel                        = svt_nxm_sad_kernel_helper_avx2;
        svt_initialize_buffer_32bits              = svt_initialize_buffer_32bits_sse2_intrin;
    }
    if (level >= 2) {
        svt_sad_loop_kernel = svt_sad_loop_kernel_avx512_intrin;
        svt_nxm_sad_kernel  = svt_nxm_sad_kernel_helper_avx512;
    }
#endif
    g_simd_level = level;
}

REF_API int ref_simd_available(void) {
#ifdef REF_HAVE_SIMD
    return 1;
#else
    return 0;
#endif
}

static void desc_from_plane(EbPictureBufferDesc *d, const OraclePlane *p) {
    memset(d, 0, sizeof *d);
    d->buffer_y   = p->buf;
    d->stride_y   = (uint16_t)p->stride;
    d->org_x      = (uint16_t)p->org_x;
    d->org_y      = (uint16_t)p->org_y;
    d->width      = (uint16_t)p->width;
    d->height     = (uint16_t)p->height;
    d->max_width  = (uint16_t)p->width;
    d->max_height = (uint16_t)p->height;
    d->bit_depth  = EB_EIGHT_BIT;
}

/* leaf kernels of the reference, for the kernel-level tests --------------------------------- */
REF_API void ref_downsample_2d(uint8_t *in, uint32_t in_stride, uint32_t w, uint32_t h, uint8_t *out, uint32_t out_stride,
                       uint32_t step) {
    svt_aom_downsample_2d_c(in, in_stride, w, h, out, out_stride, step);
}
REF_API void ref_sad_loop(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride, uint32_t bh, uint32_t bw,
                  uint64_t *best, int16_t *x, int16_t *y, uint32_t raw, uint8_t skip, int16_t sw, int16_t sh) {
    svt_sad_loop_kernel_c(src, src_stride, ref, ref_stride, bh, bw, best, x, y, raw, skip, sw, sh);
}

/* svt_aom_downsample_filtering_input_picture + paddings on caller-provided planes */
REF_API int ref_build_pyramid(OraclePicture *pic) {
    set_rtcd(0);
    PictureParentControlSet *pcs = calloc(1, sizeof *pcs);
    if (!pcs)
        return -1;
    pcs->enable_hme_flag = pcs->enable_hme_level0_flag = pcs->enable_hme_level1_flag = 1;
    EbPictureBufferDesc f, q, s;
    desc_from_plane(&f, &pic->full);
    desc_from_plane(&q, &pic->quarter);
    desc_from_plane(&s, &pic->sixteenth);
    svt_aom_downsample_filtering_input_picture(pcs, &f, &q, &s);
    free(pcs);
    return 0;
}
/* svt_aom_generate_padding on the full plane (what svt_aom_pad_input_pictures does for luma) */
REF_API void ref_pad_full(OraclePicture *pic) {
    svt_aom_generate_padding(pic->full.buf, pic->full.stride, pic->full.width, pic->full.height, pic->full.org_x,
                             pic->full.org_y);
}

static void ctx_from_params(MeContext *m, const MeGpuParams *p) {
    m->me_type                = (EbMeType)p->me_type; /* ME_GPU_ME_* carry the EbMeType values */
    m->tf_me_exit_th          = p->tf_me_exit_th;
    m->num_of_list_to_search  = p->num_of_list_to_search;
    m->num_of_ref_pic_to_search[0] = p->num_of_ref_pic_to_search[0];
    m->num_of_ref_pic_to_search[1] = p->num_of_ref_pic_to_search[1];
    m->temporal_layer_index   = p->temporal_layer_index;
    m->is_ref                 = p->is_ref;
    m->enable_hme_flag        = p->enable_hme_flag;
    m->enable_hme_level0_flag = p->enable_hme_level0_flag;
    m->enable_hme_level1_flag = p->enable_hme_level1_flag;
    m->enable_hme_level2_flag = p->enable_hme_level2_flag;
    m->hme_search_method      = p->hme_search_method;
    m->me_search_method       = p->me_search_method;
    m->use_best_unipred_cand_only = p->use_best_unipred_cand_only;
    m->reduce_hme_l0_sr_th_min = p->reduce_hme_l0_sr_th_min;
    m->reduce_hme_l0_sr_th_max = p->reduce_hme_l0_sr_th_max;
    m->num_hme_sa_w = p->num_hme_sa_w;
    m->num_hme_sa_h = p->num_hme_sa_h;
    memcpy(&m->me_sa, &p->me_sa, sizeof m->me_sa);
    memcpy(&m->hme_l0_sa, &p->hme_l0_sa, sizeof m->hme_l0_sa);
    memcpy(&m->hme_l1_sa, &p->hme_l1_sa, sizeof m->hme_l1_sa);
    memcpy(&m->hme_l2_sa, &p->hme_l2_sa, sizeof m->hme_l2_sa);
    m->me_hme_prune_ctrls.enable_me_hme_ref_pruning               = p->enable_me_hme_ref_pruning;
    m->me_hme_prune_ctrls.prune_ref_if_hme_sad_dev_bigger_than_th = p->prune_ref_if_hme_sad_dev_bigger_than_th;
    m->me_hme_prune_ctrls.prune_ref_if_me_sad_dev_bigger_than_th  = p->prune_ref_if_me_sad_dev_bigger_than_th;
    m->me_hme_prune_ctrls.zz_sad_th    = p->zz_sad_th;
    m->me_hme_prune_ctrls.zz_sad_pct   = p->zz_sad_pct;
    m->me_hme_prune_ctrls.phme_sad_th  = p->phme_sad_th;
    m->me_hme_prune_ctrls.phme_sad_pct = p->phme_sad_pct;
    m->me_sr_adjustment_ctrls.enable_me_sr_adjustment              = p->enable_me_sr_adjustment;
    m->me_sr_adjustment_ctrls.reduce_me_sr_based_on_mv_length_th   = p->reduce_me_sr_based_on_mv_length_th;
    m->me_sr_adjustment_ctrls.stationary_hme_sad_abs_th            = p->stationary_hme_sad_abs_th;
    m->me_sr_adjustment_ctrls.stationary_me_sr_divisor             = p->stationary_me_sr_divisor;
    m->me_sr_adjustment_ctrls.reduce_me_sr_based_on_hme_sad_abs_th = p->reduce_me_sr_based_on_hme_sad_abs_th;
    m->me_sr_adjustment_ctrls.me_sr_divisor_for_low_hme_sad        = p->me_sr_divisor_for_low_hme_sad;
    m->me_sr_adjustment_ctrls.distance_based_hme_resizing          = p->distance_based_hme_resizing;
    m->me_8x8_var_ctrls.enabled        = p->me_8x8_var_enabled;
    m->me_8x8_var_ctrls.me_sr_div4_th  = p->me_sr_div4_th;
    m->me_8x8_var_ctrls.me_sr_div2_th  = p->me_sr_div2_th;
    m->me_8x8_var_ctrls.me_sr_mult2_th = p->me_sr_mult2_th;
    m->mv_based_sa_adj.enabled          = p->mv_sa_adj_enabled;
    m->mv_based_sa_adj.nearest_ref_only = p->mv_sa_adj_nearest_ref_only;
    m->mv_based_sa_adj.mv_size_th       = p->mv_sa_adj_mv_size_th;
    m->mv_based_sa_adj.sa_multiplier    = p->mv_sa_adj_sa_multiplier;
    m->prehme_ctrl.enable           = p->prehme_enable;
    m->prehme_ctrl.skip_search_line = p->prehme_skip_search_line;
    m->prehme_ctrl.l1_early_exit    = p->prehme_l1_early_exit;
    memcpy(m->prehme_ctrl.prehme_sa_cfg, p->prehme_sa_cfg, sizeof m->prehme_ctrl.prehme_sa_cfg);
    m->prune_me_candidates_th      = p->prune_me_candidates_th;
    m->me_early_exit_th            = p->me_early_exit_th;
    m->me_safe_limit_zz_th         = p->me_safe_limit_zz_th;
    m->prev_me_stage_based_exit_th = p->prev_me_stage_based_exit_th;
}

/* MeContext -> MeGpuParams is the reference-side binding's own mapping (integration/svt_me_gpu_hook.c): the derive-params
 * tests below therefore pin the code a maintainer ships, not a private copy of it */
#include "../integration/svt_me_gpu_hook.h"
static void params_from_ctx(MeGpuParams *p, const MeContext *m) { svt_me_gpu_params_from_me_ctx(p, m); }

/* Extra inputs of svt_aom_sig_deriv_me that come from enc_handle.c's static set_mrp_ctrl() and cannot be
 * called from outside the reference: passed in by the caller (the host-side restatement computes them,
 * and tests/test_derive_params.py pins them against enc_handle.c:3397-3670 by table). */
typedef struct RefMrpInputs {
    uint8_t  safe_limit_nref;
    uint32_t safe_limit_zz_th;
} RefMrpInputs;

/* Runs the reference's svt_aom_sig_deriv_me (Codec/enc_mode_config.c:681) plus the exported helper
 * functions that set the pcs-level inputs, and flattens the result into MeGpuParams. Fields that only
 * enc_handle.c / pd_process.c know (reference counts, max_cand...) are left untouched in *out. */
REF_API int ref_derive_params(const MeGpuEncoderConfig *cfg, const RefMrpInputs *mrp, MeGpuParams *out) {
    PictureParentControlSet *pcs = calloc(1, sizeof *pcs);
    SequenceControlSet      *scs = calloc(1, sizeof *scs);
    MeContext               *m   = calloc(1, sizeof *m);
    if (!pcs || !scs || !m)
        return -1;
    pcs->scs                         = scs;
    pcs->enc_mode                    = (EncMode)cfg->enc_mode;
    pcs->sc_class1                   = cfg->sc_class1;
    pcs->temporal_layer_index        = cfg->temporal_layer_index;
    pcs->hierarchical_levels         = cfg->hierarchical_levels;
    pcs->slice_type                  = (SliceType)cfg->slice_type;
    scs->static_config.pred_structure = cfg->rtc_tune ? SVT_AV1_PRED_LOW_DELAY_B : SVT_AV1_PRED_RANDOM_ACCESS;
    scs->static_config.qp            = cfg->qp;
    scs->static_config.enc_mode      = cfg->enc_mode;
    scs->frame_rate                  = cfg->low_frame_rate ? (1u << 16) : 30;
    scs->mrp_ctrls.safe_limit_nref   = mrp->safe_limit_nref;
    scs->mrp_ctrls.safe_limit_zz_th  = mrp->safe_limit_zz_th;
    svt_aom_derive_input_resolution(&scs->input_resolution, (uint32_t)cfg->width * cfg->height);
    pcs->input_resolution = scs->input_resolution;
    /* HME level flags: Codec/enc_mode_config.c:1633-1645 (inside svt_aom_sig_deriv_multi_processes, which
     * needs a fully built encoder; the three assignments are restated here) */
    pcs->enable_hme_flag        = 1;
    pcs->enable_hme_level0_flag = 1;
    pcs->enable_hme_level1_flag = 1;
    pcs->enable_hme_level2_flag = (cfg->sc_class1 || cfg->enc_mode <= ENC_M6) ? 1 : 0;
    /* Codec/enc_mode_config.c:1845-1848 */
    pcs->use_best_me_unipred_cand_only = cfg->enc_mode <= ENC_M3 ? 0 : 1;
    svt_aom_set_gm_controls(pcs, svt_aom_derive_gm_level(pcs, true));
    svt_aom_sig_deriv_me(scs, pcs, m);

    params_from_ctx(out, m);
    out->picture_number                  = cfg->picture_number;
    out->input_resolution                = (uint8_t)scs->input_resolution;
    out->enable_me_8x8                   = svt_aom_get_enable_me_8x8(pcs->enc_mode, cfg->rtc_tune, scs->input_resolution);
    out->enable_me_16x16                 = svt_aom_get_enable_me_16x16(pcs->enc_mode);
    out->max_number_of_pus_per_sb        = SQUARE_PU_COUNT; /* resource_coordination_process.c:425 */
    out->gm_enabled                      = pcs->gm_ctrls.enabled;
    out->gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
    out->hierarchical_levels             = cfg->hierarchical_levels;
    out->pcs_temporal_layer_index        = cfg->temporal_layer_index;
    out->temporal_layer_index            = cfg->temporal_layer_index;
    out->similar_brightness_refs         = cfg->similar_brightness_refs;
    out->is_ref                          = cfg->is_ref;
    out->me_type                         = ME_GPU_ME_OPEN_LOOP; /* me_process.c:215 */
    free(m);
    free(scs);
    free(pcs);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * One picture's ME state as the encoder holds it: pcs / scs / pa_me_data / MeSbResults / b64 geometry / picture
 * descriptors, built ONCE per picture; the blocks are then run against it by any number of threads, each with its own
 * MeContext (me_process.c:75-88: one MeContext per ME thread).
 * ---------------------------------------------------------------------------------------------- */
typedef struct RefPicCtx {
    PictureParentControlSet *pcs;
    SequenceControlSet      *scs;
    MotionEstimationData    *md;
    MeSbResults             *res;
    MeSbResults            **rp;
    B64Geom                 *geo;
    EbPictureBufferDesc      sf, sq, ss, rf[2][4], rq[2][4], rs[2][4];
    const OraclePicture     *refs[2][4];
    const MeGpuParams       *p;
    MeGpuOutputs            *out;
    uint32_t                 total, P;
} RefPicCtx;

static void ref_pic_ctx_free(RefPicCtx *c) {
    if (!c)
        return;
    free(c->geo); free(c->rp); free(c->res); free(c->md); free(c->scs); free(c->pcs); free(c);
}

static RefPicCtx *ref_pic_ctx_new(const OraclePicture *src, const OraclePicture *const refs[2][4], const MeGpuParams *p, MeGpuOutputs *out) {
    const int pic_w = (src->full.width + 7) & ~7, pic_h = (src->full.height + 7) & ~7;
    const uint32_t bw = (pic_w + 63) / 64, bh = (pic_h + 63) / 64, total = bw * bh;
    if (out->b64_count < total)
        return NULL;
    RefPicCtx *c = calloc(1, sizeof *c);
    if (!c)
        return NULL;
    PictureParentControlSet *pcs = c->pcs = calloc(1, sizeof *pcs);
    SequenceControlSet      *scs = c->scs = calloc(1, sizeof *scs);
    MotionEstimationData    *md  = c->md  = calloc(1, sizeof *md);
    MeSbResults             *res = c->res = calloc(total, sizeof *res);
    MeSbResults            **rp  = c->rp  = calloc(total, sizeof *rp);
    B64Geom                 *geo = c->geo = calloc(total, sizeof *geo);
    if (!pcs || !scs || !md || !res || !rp || !geo) {
        ref_pic_ctx_free(c);
        return NULL;
    }
    const uint32_t P = p->enable_me_16x16 ? (p->enable_me_8x8 ? p->max_number_of_pus_per_sb : 21) : 5;
    c->total = total; c->P = P; c->p = p; c->out = out;
    pcs->scs                      = scs;
    pcs->picture_number           = p->picture_number;
    pcs->aligned_width            = (uint16_t)pic_w;
    pcs->aligned_height           = (uint16_t)pic_h;
    pcs->enable_me_8x8            = p->enable_me_8x8;
    pcs->enable_me_16x16          = p->enable_me_16x16;
    pcs->max_number_of_pus_per_sb = p->max_number_of_pus_per_sb;
    pcs->gm_ctrls.enabled         = p->gm_enabled;
    pcs->gm_ctrls.use_distance_based_active_th = p->gm_use_distance_based_active_th;
    pcs->hierarchical_levels      = p->hierarchical_levels;
    pcs->temporal_layer_index     = p->pcs_temporal_layer_index;
    pcs->similar_brightness_refs  = p->similar_brightness_refs;
    pcs->b64_geom                 = geo;
    pcs->pa_me_data               = md;
    pcs->me_64x64_distortion      = out->me_64x64_distortion;
    pcs->me_32x32_distortion      = out->me_32x32_distortion;
    pcs->me_16x16_distortion      = out->me_16x16_distortion;
    pcs->me_8x8_distortion        = out->me_8x8_distortion;
    pcs->me_8x8_cost_variance     = out->me_8x8_cost_variance;
    pcs->rc_me_distortion         = out->rc_me_distortion;
    pcs->stationary_block_present_sb = out->stationary_block_present_sb;
    pcs->rc_me_allow_gm           = out->rc_me_allow_gm;
    scs->mrp_ctrls.only_l_bwd     = p->only_l_bwd;
    scs->input_resolution         = (EbInputResolution)p->input_resolution;
    scs->b64_size                 = 64;
    md->me_results                = rp;
    md->max_cand                  = p->max_cand;
    md->max_refs                  = p->max_refs;
    md->max_l0                    = p->max_l0;
    for (uint32_t b = 0; b < total; b++) {
        rp[b]                        = &res[b];
        res[b].total_me_candidate_index = out->total_me_candidate_index + (size_t)b * P;
        res[b].me_candidate_array    = (MeCandidate *)(out->me_candidate_array + (size_t)b * P * p->max_cand);
        res[b].me_mv_array           = (MvCandidate *)(out->me_mv_array + (size_t)b * P * p->max_refs);
        geo[b].org_x                 = (uint16_t)((b % bw) * 64);
        geo[b].org_y                 = (uint16_t)((b / bw) * 64);
        geo[b].width                 = (uint8_t)((pic_w - geo[b].org_x) < 64 ? pic_w - geo[b].org_x : 64);
        geo[b].height                = (uint8_t)((pic_h - geo[b].org_y) < 64 ? pic_h - geo[b].org_y : 64);
    }
    desc_from_plane(&c->sf, &src->full);
    desc_from_plane(&c->sq, &src->quarter);
    desc_from_plane(&c->ss, &src->sixteenth);
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            c->refs[li][ri] = refs[li][ri];
            desc_from_plane(&c->rf[li][ri], &refs[li][ri]->full);
            desc_from_plane(&c->rq[li][ri], &refs[li][ri]->quarter);
            desc_from_plane(&c->rs[li][ri], &refs[li][ri]->sixteenth);
        }
    return c;
}

/* what svt_aom_motion_estimation_kernel does when it takes an item of this picture (me_process.c:121, 219-262) */
static void ref_pic_ctx_bind(RefPicCtx *c, MeContext *m) {
    const MeGpuParams *p = c->p;
    ctx_from_params(m, p);
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++) {
            m->me_ds_ref_array[li][ri].picture_ptr           = &c->rf[li][ri];
            m->me_ds_ref_array[li][ri].quarter_picture_ptr   = &c->rq[li][ri];
            m->me_ds_ref_array[li][ri].sixteenth_picture_ptr = &c->rs[li][ri];
            m->me_ds_ref_array[li][ri].picture_number        = c->refs[li][ri]->picture_number;
        }
}

/* blocks [first, first + count) of the picture on MeContext m (bound to it). canonical: reset the state the reference
 * leaves stale between blocks (never read for an output) so that the parity taps are reproducible; the timed arm
 * leaves it as the encoder does. */
static void ref_pic_ctx_run(RefPicCtx *c, MeContext *m, uint32_t first, uint32_t count, int canonical) {
    const MeGpuParams *p   = c->p;
    MeGpuOutputs      *out = c->out;
    EbPictureBufferDesc *sf = &c->sf, *sq = &c->sq, *ss = &c->ss;
    for (uint32_t b = first; b < first + count && b < c->total; b++) {
        const uint32_t ox = c->geo[b].org_x, oy = c->geo[b].org_y;
        /* me_process.c:183-214 */
        m->b64_src_ptr                 = &sf->buffer_y[(sf->org_y + oy) * sf->stride_y + sf->org_x + ox];
        m->b64_src_stride              = sf->stride_y;
        m->quarter_b64_buffer          = &sq->buffer_y[(sq->org_y + (oy >> 1)) * sq->stride_y + sq->org_x + (ox >> 1)];
        m->quarter_b64_buffer_stride   = sq->stride_y;
        m->sixteenth_b64_buffer        = &ss->buffer_y[(ss->org_y + (oy >> 2)) * ss->stride_y + ss->org_x + (ox >> 2)];
        m->sixteenth_b64_buffer_stride = ss->stride_y;
        m->me_type                     = (EbMeType)p->me_type;
        m->tf_use_pred_64x64_only_th   = 1; /* temporal_filtering.c:3174 sets tf_ctrls.use_pred_64x64_only_th; 0xFF marks the exit */
        if (canonical) {
            for (int li = 0; li < 2; li++)
                for (int ri = 0; ri < 4; ri++)
                    for (int k = 0; k < SQUARE_PU_COUNT; k++) m->p_sb_best_sad[li][ri][k] = MAX_SAD_VALUE;
            memset(m->prehme_data, 0, sizeof m->prehme_data);
            for (int li = 0; li < 2; li++)
                for (int ri = 0; ri < 4; ri++) m->search_results[li][ri].hme_sc_x = m->search_results[li][ri].hme_sc_y = 0;
        }

        svt_aom_motion_estimation_b64(c->pcs, b, ox, oy, m, sf);

        if (p->me_type == ME_GPU_ME_MCTF) {
            if (out->tf_best_sad) memcpy(out->tf_best_sad + (size_t)b * 85, m->p_sb_best_sad[0][0], 85 * sizeof(uint32_t));
            if (out->tf_best_mv) memcpy(out->tf_best_mv + (size_t)b * 85, m->p_sb_best_mv[0][0], 85 * sizeof(uint32_t));
            if (out->tf_hme_centre) {
                out->tf_hme_centre[2 * b]     = m->search_results[0][0].hme_sc_x;
                out->tf_hme_centre[2 * b + 1] = m->search_results[0][0].hme_sc_y;
            }
            if (out->tf_hme_sad) out->tf_hme_sad[b] = (uint32_t)m->search_results[0][0].hme_sad;
            if (out->tf_early_exit) out->tf_early_exit[b] = m->tf_use_pred_64x64_only_th == (uint8_t)~0;
        }
        if (out->p_sb_best_sad) memcpy(out->p_sb_best_sad + (size_t)b * 2 * 4 * 85, m->p_sb_best_sad, sizeof m->p_sb_best_sad);
        if (out->p_sb_best_mv) memcpy(out->p_sb_best_mv + (size_t)b * 2 * 4 * 85, m->p_sb_best_mv, sizeof m->p_sb_best_mv);
        if (out->hme_search_centre || out->hme_sad || out->do_ref)
            for (int li = 0; li < 2; li++)
                for (int ri = 0; ri < 4; ri++) {
                    size_t k = ((size_t)b * 2 + li) * 4 + ri;
                    if (out->hme_search_centre) {
                        out->hme_search_centre[2 * k]     = m->search_results[li][ri].hme_sc_x;
                        out->hme_search_centre[2 * k + 1] = m->search_results[li][ri].hme_sc_y;
                    }
                    if (out->hme_sad) out->hme_sad[k] = m->search_results[li][ri].hme_sad;
                    if (out->do_ref) out->do_ref[k] = m->search_results[li][ri].do_ref;
                }
    }
}

/* ME of b64s [first, first+count) with the reference's own svt_aom_motion_estimation_b64.
 * simd: 0 = C kernels (the specification), 1 = AVX2, 2 = AVX-512 sad_loop on top of AVX2. */
REF_API int ref_me_b64_range(const OraclePicture *src, const OraclePicture *const refs[2][4], const MeGpuParams *p,
                     MeGpuOutputs *out, uint32_t first, uint32_t count, int simd) {
    set_rtcd(simd);
    RefPicCtx *c = ref_pic_ctx_new(src, refs, p, out);
    MeContext *m = calloc(1, sizeof *m);
    if (!c || !m) {
        ref_pic_ctx_free(c);
        free(m);
        return c ? -1 : -2;
    }
    ref_pic_ctx_bind(c, m);
    ref_pic_ctx_run(c, m, first, count, 1);
    out->b64_count = c->total;
    out->pu_count  = c->P;
    if (p->me_type == ME_GPU_ME_MCTF) { /* counted by hme_b64 (ME:2469-2474) over the blocks of this call */
        out->tf_tot_horz_blks = m->tf_tot_horz_blks;
        out->tf_tot_vert_blks = m->tf_tot_vert_blks;
    }
    ref_pic_ctx_free(c);
    free(m);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * The timed CPU arm (bench.py --impl reference, cpu_baseline): open-loop ME of a whole step of pictures on `nthreads`
 * threads, the way the encoder's ME stage runs it - every picture cut into its ME segments (enc_handle.c:396-410:
 * 6 x 8 = 48 at 1080p and above), the items of all pictures in one queue (pd_process.c:3545-3557), each thread with a
 * private MeContext taking the next item (enc_handle.c:731-773) - minus the FIFO and object-pool overhead. The picture
 * state is built once per picture per step, outside the b64 loop. every > 1: only every `every`-th segment of a picture
 * is run (a bounded sample for the slow configurations); *b64_done = blocks actually searched.
 * ---------------------------------------------------------------------------------------------- */
#include <dlfcn.h>
#include <pthread.h>
typedef struct RefStepJob {
    const OraclePicture *src;
    const OraclePicture *const (*refs)[4];
    const MeGpuParams   *p;
    MeGpuOutputs        *out;
} RefStepJob;
typedef struct RefStepShared {
    RefPicCtx **ctx;
    int         njobs, seg_cols, seg_rows, every;
    volatile int next; /* next (picture, segment) item */
    volatile long b64_done;
} RefStepShared;

static void *ref_step_thread(void *arg) {
    RefStepShared *sh = (RefStepShared *)arg;
    MeContext     *m  = calloc(1, sizeof *m);
    int            bound = -1;
    const int      nseg = sh->seg_cols * sh->seg_rows;
    for (;;) {
        const int item = __sync_fetch_and_add(&sh->next, 1);
        if (item >= sh->njobs * nseg)
            break;
        const int pic = item / nseg, seg = item % nseg;
        if (sh->every > 1 && (seg + 7 * pic) % sh->every)
            continue;
        RefPicCtx *c = sh->ctx[pic];
        if (bound != pic) {
            ref_pic_ctx_bind(c, m);
            bound = pic;
        }
        const uint32_t bw = (c->pcs->aligned_width + 63) / 64, bh = (c->pcs->aligned_height + 63) / 64;
        const uint32_t sx = seg % sh->seg_cols, sy = seg / sh->seg_cols;
        /* SEGMENT_START_IDX / SEGMENT_END_IDX, Codec/definitions.h */
        const uint32_t x0 = (sx * bw) / sh->seg_cols, x1 = ((sx + 1) * bw) / sh->seg_cols;
        const uint32_t y0 = (sy * bh) / sh->seg_rows, y1 = ((sy + 1) * bh) / sh->seg_rows;
        for (uint32_t y = y0; y < y1; y++) ref_pic_ctx_run(c, m, y * bw + x0, x1 - x0, 0);
        __sync_fetch_and_add(&sh->b64_done, (long)((x1 - x0) * (y1 - y0)));
    }
    free(m);
    return NULL;
}

REF_API int ref_me_step_mt(const RefStepJob *jobs, int njobs, int nthreads, int simd, int every, long *b64_done) {
    set_rtcd(simd);
    if (njobs <= 0 || nthreads <= 0 || nthreads > 256)
        return -3;
    RefStepShared sh;
    memset(&sh, 0, sizeof sh);
    sh.ctx   = calloc((size_t)njobs, sizeof *sh.ctx);
    sh.njobs = njobs;
    sh.every = every < 1 ? 1 : every;
    int rc   = 0;
    for (int i = 0; i < njobs && !rc; i++) {
        sh.ctx[i] = ref_pic_ctx_new(jobs[i].src, jobs[i].refs, jobs[i].p, jobs[i].out);
        if (!sh.ctx[i])
            rc = -1;
    }
    if (!rc) {
        /* enc_handle.c:394-397: me_seg_h = ((h + 32) / 64 < 6) ? 1 : 8, me_seg_w = ((w + 32) / 64 < 10) ? 1 : 6 */
        const uint32_t w = sh.ctx[0]->pcs->aligned_width, h = sh.ctx[0]->pcs->aligned_height;
        sh.seg_cols = ((w + 32) / 64) < 10 ? 1 : 6;
        sh.seg_rows = ((h + 32) / 64) < 6 ? 1 : 8;
        pthread_t th[256];
        int       started = 0;
        for (; started < nthreads; started++)
            if (pthread_create(&th[started], NULL, ref_step_thread, &sh))
                break;
        for (int i = 0; i < started; i++) pthread_join(th[i], NULL);
        if (!started)
            rc = -4;
        for (int i = 0; i < njobs; i++) {
            jobs[i].out->b64_count = sh.ctx[i]->total;
            jobs[i].out->pu_count  = sh.ctx[i]->P;
        }
    }
    if (b64_done)
        *b64_done = sh.b64_done;
    for (int i = 0; i < njobs; i++) ref_pic_ctx_free(sh.ctx[i]);
    free(sh.ctx);
    return rc;
}

/* ------------------------------------------------------------------------------------------------
 * The reference-side binding (integration/svt_me_gpu_hook.c) driven the way the encoder would drive it: a pcs with
 * individually allocated MeSbResults (pcs.c:1425-1444, me_sb_results_ctor), pictures as EbPictureBufferDesc, controls
 * in a MeContext. svt_me_gpu_picture() stands where the b64 loop of svt_aom_motion_estimation_kernel stands; what it
 * leaves in the pcs is copied into `out` (same layout as ref_me_b64_range) for the byte-for-byte comparison with the
 * CPU loop. Needs a GPU: called from the `-m gpu` tests only.
 * ---------------------------------------------------------------------------------------------- */
REF_API int ref_hook_me_picture(const OraclePicture *src, const OraclePicture *const refs[2][4], const MeGpuParams *p,
                                MeGpuOutputs *out, const char *gpu_library_path) {
    const int pic_w = (src->full.width + 7) & ~7, pic_h = (src->full.height + 7) & ~7;
    const uint32_t bw = (pic_w + 63) / 64, bh = (pic_h + 63) / 64, total = bw * bh;
    if (out->b64_count < total)
        return -2;
    PictureParentControlSet *pcs = calloc(1, sizeof *pcs);
    SequenceControlSet      *scs = calloc(1, sizeof *scs);
    MeContext               *m   = calloc(1, sizeof *m);
    MotionEstimationData    *md  = calloc(1, sizeof *md);
    MeSbResults            **rp  = calloc(total, sizeof *rp);
    if (!pcs || !scs || !m || !md || !rp)
        return -1;
    const uint32_t P = p->enable_me_16x16 ? (p->enable_me_8x8 ? p->max_number_of_pus_per_sb : 21) : 5;
    ctx_from_params(m, p);
    pcs->scs                      = scs;
    pcs->picture_number           = p->picture_number;
    pcs->aligned_width            = (uint16_t)pic_w;
    pcs->aligned_height           = (uint16_t)pic_h;
    pcs->enable_me_8x8            = p->enable_me_8x8;
    pcs->enable_me_16x16          = p->enable_me_16x16;
    pcs->max_number_of_pus_per_sb = p->max_number_of_pus_per_sb;
    pcs->gm_ctrls.enabled         = p->gm_enabled;
    pcs->gm_ctrls.use_distance_based_active_th = p->gm_use_distance_based_active_th;
    pcs->hierarchical_levels      = p->hierarchical_levels;
    pcs->temporal_layer_index     = p->pcs_temporal_layer_index;
    pcs->similar_brightness_refs  = p->similar_brightness_refs;
    pcs->pa_me_data               = md;
    pcs->me_64x64_distortion      = calloc(total, sizeof(uint32_t)); /* pcs.c allocates these per picture, b64_total_count long */
    pcs->me_32x32_distortion      = calloc(total, sizeof(uint32_t));
    pcs->me_16x16_distortion      = calloc(total, sizeof(uint32_t));
    pcs->me_8x8_distortion        = calloc(total, sizeof(uint32_t));
    pcs->me_8x8_cost_variance     = calloc(total, sizeof(uint32_t));
    pcs->rc_me_distortion         = calloc(total, sizeof(uint32_t));
    pcs->stationary_block_present_sb = calloc(total, 1);
    pcs->rc_me_allow_gm           = calloc(total, 1);
    scs->mrp_ctrls.only_l_bwd     = p->only_l_bwd;
    scs->input_resolution         = (EbInputResolution)p->input_resolution;
    scs->b64_size                 = 64;
    md->me_results                = rp;
    md->max_cand                  = p->max_cand;
    md->max_refs                  = p->max_refs;
    md->max_l0                    = p->max_l0;
    for (uint32_t b = 0; b < total; b++) { /* me_sb_results_ctor: three arrays per b64, each its own allocation */
        rp[b]                           = calloc(1, sizeof(MeSbResults));
        rp[b]->me_mv_array              = calloc((size_t)P * p->max_refs, sizeof(MvCandidate));
        rp[b]->me_candidate_array       = calloc((size_t)P * p->max_cand, sizeof(MeCandidate));
        rp[b]->total_me_candidate_index = calloc(P, 1);
    }
    int rc = 0;
    if (svt_me_gpu_hook_init(gpu_library_path, pic_w, pic_h, 4, 3, 10) != EB_ErrorNone)
        rc = -10;
    /* picture analysis side: every picture is registered once, from its padded y8b descriptor */
    EbPictureBufferDesc d;
    PictureParentControlSet *rpcs = calloc(1, sizeof *rpcs);
    uint64_t registered[9];
    int      nreg = 0;
    for (int k = -1; !rc && k < 8; k++) {
        const OraclePicture *pic = k < 0 ? src : refs[k / 4][k % 4];
        if (k >= 0 && (k / 4 >= p->num_of_list_to_search || k % 4 >= p->num_of_ref_pic_to_search[k / 4]))
            continue;
        int seen = 0;
        for (int i = 0; i < nreg; i++) seen |= registered[i] == pic->picture_number;
        if (seen)
            continue;
        desc_from_plane(&d, &pic->full);
        rpcs->picture_number = pic->picture_number;
        rpcs->aligned_width  = (uint16_t)pic_w;
        rpcs->aligned_height = (uint16_t)pic_h;
        if (svt_me_gpu_register_picture(rpcs, &d) != EB_ErrorNone)
            rc = -11;
        else
            registered[nreg++] = pic->picture_number;
    }
    /* ME kernel side: me_process.c:215-262 */
    m->me_type = (EbMeType)p->me_type;
    for (int li = 0; li < p->num_of_list_to_search; li++)
        for (int ri = 0; ri < p->num_of_ref_pic_to_search[li]; ri++)
            m->me_ds_ref_array[li][ri].picture_number = refs[li][ri]->picture_number;
    if (!rc && svt_me_gpu_picture(pcs, m) != EB_ErrorNone)
        rc = -12;
    for (int i = 0; i < nreg; i++) svt_me_gpu_release_picture(registered[i]);
    svt_me_gpu_hook_shutdown();

    for (uint32_t b = 0; b < total; b++) {
        if (!rc) {
            memcpy(out->total_me_candidate_index + (size_t)b * P, rp[b]->total_me_candidate_index, P);
            memcpy(out->me_candidate_array + (size_t)b * P * p->max_cand, rp[b]->me_candidate_array, (size_t)P * p->max_cand);
            memcpy(out->me_mv_array + (size_t)b * P * p->max_refs, rp[b]->me_mv_array, (size_t)P * p->max_refs * sizeof(uint32_t));
        }
        free(rp[b]->total_me_candidate_index); free(rp[b]->me_candidate_array); free(rp[b]->me_mv_array); free(rp[b]);
    }
    if (!rc) {
        memcpy(out->me_64x64_distortion, pcs->me_64x64_distortion, total * sizeof(uint32_t));
        memcpy(out->me_32x32_distortion, pcs->me_32x32_distortion, total * sizeof(uint32_t));
        memcpy(out->me_16x16_distortion, pcs->me_16x16_distortion, total * sizeof(uint32_t));
        memcpy(out->me_8x8_distortion, pcs->me_8x8_distortion, total * sizeof(uint32_t));
        memcpy(out->me_8x8_cost_variance, pcs->me_8x8_cost_variance, total * sizeof(uint32_t));
        memcpy(out->rc_me_distortion, pcs->rc_me_distortion, total * sizeof(uint32_t));
        memcpy(out->stationary_block_present_sb, pcs->stationary_block_present_sb, total);
        memcpy(out->rc_me_allow_gm, pcs->rc_me_allow_gm, total);
        out->b64_count = total;
        out->pu_count  = P;
    }
    free(pcs->me_64x64_distortion); free(pcs->me_32x32_distortion); free(pcs->me_16x16_distortion); free(pcs->me_8x8_distortion);
    free(pcs->me_8x8_cost_variance); free(pcs->rc_me_distortion); free(pcs->stationary_block_present_sb); free(pcs->rc_me_allow_gm);
    free(rpcs); free(rp); free(md); free(m); free(scs); free(pcs);
    return rc;
}

/* register one OraclePicture through the binding (needs only picture number and aligned size of a pcs) */
static int hook_register(const OraclePicture *pic, int pic_w, int pic_h) {
    EbPictureBufferDesc      d;
    PictureParentControlSet *rpcs = calloc(1, sizeof *rpcs);
    if (!rpcs)
        return -1;
    desc_from_plane(&d, &pic->full);
    rpcs->picture_number = pic->picture_number;
    rpcs->aligned_width  = (uint16_t)pic_w;
    rpcs->aligned_height = (uint16_t)pic_h;
    const int rc = svt_me_gpu_register_picture(rpcs, &d) == EB_ErrorNone ? 0 : -11;
    free(rpcs);
    return rc;
}

/* Temporal-filter ME of one (centre, neighbour) pair through the binding: svt_me_gpu_tf_pair before the block loop,
 * svt_me_gpu_tf_load_block inside it; what the MeContext then holds per block goes to out->tf_* (same layout as
 * ref_me_b64_range with me_type = ME_MCTF). GPU tests only. */
REF_API int ref_hook_tf_pair(const OraclePicture *src, const OraclePicture *ref, const MeGpuParams *p, MeGpuOutputs *out,
                             const char *gpu_library_path) {
    const int pic_w = (src->full.width + 7) & ~7, pic_h = (src->full.height + 7) & ~7;
    const uint32_t total = (uint32_t)((pic_w + 63) / 64) * ((pic_h + 63) / 64);
    if (out->b64_count < total)
        return -2;
    PictureParentControlSet *pcs = calloc(1, sizeof *pcs);
    SequenceControlSet      *scs = calloc(1, sizeof *scs);
    MeContext               *m   = calloc(1, sizeof *m);
    MotionEstimationData    *md  = calloc(1, sizeof *md);
    if (!pcs || !scs || !m || !md)
        return -1;
    ctx_from_params(m, p);
    pcs->scs                      = scs;
    pcs->picture_number           = p->picture_number;
    pcs->aligned_width            = (uint16_t)pic_w;
    pcs->aligned_height           = (uint16_t)pic_h;
    pcs->enable_me_8x8            = p->enable_me_8x8;
    pcs->enable_me_16x16          = p->enable_me_16x16;
    pcs->max_number_of_pus_per_sb = p->max_number_of_pus_per_sb;
    pcs->hierarchical_levels      = p->hierarchical_levels;
    pcs->temporal_layer_index     = p->pcs_temporal_layer_index;
    pcs->pa_me_data               = md;
    scs->input_resolution         = (EbInputResolution)p->input_resolution;
    md->max_cand                  = p->max_cand;
    md->max_refs                  = p->max_refs;
    md->max_l0                    = p->max_l0;
    int rc = svt_me_gpu_hook_init(gpu_library_path, pic_w, pic_h, 4, 3, 4) == EB_ErrorNone ? 0 : -10;
    if (!rc) rc = hook_register(src, pic_w, pic_h);
    if (!rc && ref->picture_number != src->picture_number) rc = hook_register(ref, pic_w, pic_h);
    SvtMeGpuTfPair pair = {total, out->tf_best_sad, out->tf_best_mv, out->tf_hme_centre, out->tf_hme_sad, out->tf_early_exit};
    /* the pair's arrays are filled by the engine; the loop below reads them back through the MeContext, block by block */
    if (!rc && svt_me_gpu_tf_pair(pcs, m, ref->picture_number, &pair) != EB_ErrorNone)
        rc = -12;
    if (!rc) {
        uint32_t *sad = malloc((size_t)total * 85 * 4), *mv = malloc((size_t)total * 85 * 4);
        for (uint32_t b = 0; b < total; b++) {
            m->tf_use_pred_64x64_only_th = 1;
            svt_me_gpu_tf_load_block(m, &pair, b);
            memcpy(sad + (size_t)b * 85, m->p_sb_best_sad[0][0], 85 * 4);
            memcpy(mv + (size_t)b * 85, m->p_sb_best_mv[0][0], 85 * 4);
            out->tf_hme_centre[2 * b]     = m->search_results[0][0].hme_sc_x;
            out->tf_hme_centre[2 * b + 1] = m->search_results[0][0].hme_sc_y;
            out->tf_hme_sad[b]            = (uint32_t)m->search_results[0][0].hme_sad;
            out->tf_early_exit[b]         = m->tf_use_pred_64x64_only_th == (uint8_t)~0;
        }
        memcpy(out->tf_best_sad, sad, (size_t)total * 85 * 4);
        memcpy(out->tf_best_mv, mv, (size_t)total * 85 * 4);
        free(sad); free(mv);
        out->tf_tot_horz_blks = m->tf_tot_horz_blks;
        out->tf_tot_vert_blks = m->tf_tot_vert_blks;
        out->b64_count        = total;
    }
    svt_me_gpu_release_picture(src->picture_number);
    if (ref->picture_number != src->picture_number) svt_me_gpu_release_picture(ref->picture_number);
    svt_me_gpu_hook_shutdown();
    free(md); free(m); free(scs); free(pcs);
    return rc;
}

/* early_hme's detector round trip through the binding: metrics land in src_pcs->dg_detector->metrics. GPU tests only. */
REF_API int ref_hook_dg_detector(const OraclePicture *src, const OraclePicture *ref, uint8_t input_resolution, MeGpuDgMetrics *out,
                                 const char *gpu_library_path) {
    const int pic_w = (src->full.width + 7) & ~7, pic_h = (src->full.height + 7) & ~7;
    PictureParentControlSet *pcs = calloc(1, sizeof *pcs), *rpcs = calloc(1, sizeof *rpcs);
    DGDetectorSeg           *dg  = calloc(1, sizeof *dg);
    if (!pcs || !rpcs || !dg)
        return -1;
    pcs->picture_number   = src->picture_number;
    rpcs->picture_number  = ref->picture_number;
    pcs->input_resolution = (EbInputResolution)input_resolution;
    pcs->dg_detector      = dg;
    dg->ref_pic           = rpcs;
    int rc = svt_me_gpu_hook_init(gpu_library_path, pic_w, pic_h, 4, 3, 4) == EB_ErrorNone ? 0 : -10;
    if (!rc) rc = hook_register(src, pic_w, pic_h);
    if (!rc) rc = hook_register(ref, pic_w, pic_h);
    if (!rc && svt_me_gpu_dg_detector(pcs) != EB_ErrorNone)
        rc = -12;
    out->tot_dist       = dg->metrics.tot_dist;
    out->tot_cplx       = dg->metrics.tot_cplx;
    out->tot_active     = dg->metrics.tot_active;
    out->sum_in_vectors = dg->metrics.sum_in_vectors;
    svt_me_gpu_release_picture(src->picture_number);
    svt_me_gpu_release_picture(ref->picture_number);
    svt_me_gpu_hook_shutdown();
    free(dg); free(rpcs); free(pcs);
    return rc;
}

/* picture analysis takes the device-built pyramid through the binding. `pic`: full plane filled, quarter and
 * sixteenth planes to be written (padding included). hist_cpu / hist_gpu[256]: the reference's own calculate_histogram
 * (pic_analysis_process.c:166, the statistic picture analysis computes on the 1/16 plane at :1495) over the interior of
 * the sixteenth plane, once on the reference's CPU-built plane (`cpu`) and once on the fetched one. GPU tests only. */
void calculate_histogram(uint8_t *input_samples, uint32_t input_area_width, uint32_t input_area_height, uint32_t stride,
                         uint8_t decim_step, uint32_t *histogram, uint64_t *sum);
REF_API int ref_hook_pyramid_for_analysis(const OraclePicture *pic, const OraclePicture *cpu, uint32_t *hist_cpu, uint32_t *hist_gpu,
                                          const char *gpu_library_path) {
    const int pic_w = (pic->full.width + 7) & ~7, pic_h = (pic->full.height + 7) & ~7;
    PictureParentControlSet *pcs = calloc(1, sizeof *pcs);
    if (!pcs)
        return -1;
    pcs->picture_number = pic->picture_number;
    pcs->aligned_width  = (uint16_t)pic_w;
    pcs->aligned_height = (uint16_t)pic_h;
    EbPictureBufferDesc f, q, s16;
    desc_from_plane(&f, &pic->full);
    desc_from_plane(&q, &pic->quarter);
    desc_from_plane(&s16, &pic->sixteenth);
    int rc = svt_me_gpu_hook_init(gpu_library_path, pic_w, pic_h, 4, 3, 4) == EB_ErrorNone ? 0 : -10;
    if (!rc && svt_me_gpu_pyramid_for_analysis(pcs, &f, &q, &s16) != EB_ErrorNone)
        rc = -13;
    if (!rc) {
        uint64_t sum = 0;
        memset(hist_cpu, 0, 256 * sizeof(uint32_t));
        memset(hist_gpu, 0, 256 * sizeof(uint32_t));
        calculate_histogram((uint8_t *)cpu->sixteenth.buf + cpu->sixteenth.org_y * cpu->sixteenth.stride + cpu->sixteenth.org_x,
                            (uint32_t)cpu->sixteenth.width, (uint32_t)cpu->sixteenth.height, (uint32_t)cpu->sixteenth.stride, 1, hist_cpu, &sum);
        calculate_histogram(s16.buffer_y + s16.org_y * s16.stride_y + s16.org_x, s16.width, s16.height, s16.stride_y, 1, hist_gpu, &sum);
        svt_me_gpu_release_picture(pic->picture_number);
    }
    svt_me_gpu_hook_shutdown();
    free(pcs);
    return rc;
}

/* static in the reference (temporal_filtering.c:2760): level 0 copies hme_l0_sa_default_tf into hme_l0_sa */
static void harness_set_hme_search_params_mctf0(MeContext *m) { m->hme_l0_sa = m->hme_l0_sa_default_tf; }

/* The reference's svt_aom_sig_deriv_me_tf (Codec/enc_mode_config.c:834) on a pcs that carries the TfControls fields
 * it reads; the tf_enable_hme_* switch of svt_aom_sig_deriv_multi_processes (:1646-1668, inside a function that
 * needs a fully built encoder) is restated. */
REF_API int ref_derive_params_tf(const MeGpuTfConfig *cfg, MeGpuParams *out) {
    PictureParentControlSet *pcs = calloc(1, sizeof *pcs);
    SequenceControlSet      *scs = calloc(1, sizeof *scs);
    MeContext               *m   = calloc(1, sizeof *m);
    if (!pcs || !scs || !m)
        return -1;
    pcs->scs                   = scs;
    pcs->tf_ctrls.hme_me_level = cfg->hme_me_level;
    pcs->tf_ctrls.qp_opt       = cfg->qp_opt;
    pcs->tf_ctrls.me_exit_th   = cfg->me_exit_th;
    scs->static_config.qp      = cfg->qp;
    svt_aom_derive_input_resolution(&scs->input_resolution, (uint32_t)cfg->width * cfg->height);
    pcs->input_resolution          = scs->input_resolution;
    pcs->tf_enable_hme_flag        = 1;
    pcs->tf_enable_hme_level0_flag = 1;
    pcs->tf_enable_hme_level1_flag = cfg->hme_me_level <= 2;
    pcs->tf_enable_hme_level2_flag = cfg->hme_me_level == 0;
    svt_aom_sig_deriv_me_tf(pcs, m);
    harness_set_hme_search_params_mctf0(m);
    memset(out, 0, sizeof *out);
    params_from_ctx(out, m);
    out->me_type                  = ME_GPU_ME_MCTF;
    out->tf_me_exit_th            = pcs->tf_ctrls.me_exit_th;
    out->picture_number           = cfg->picture_number;
    out->input_resolution         = (uint8_t)scs->input_resolution;
    out->enable_me_8x8 = out->enable_me_16x16 = 1;
    out->max_number_of_pus_per_sb = SQUARE_PU_COUNT;
    out->pcs_temporal_layer_index = out->temporal_layer_index = cfg->temporal_layer_index;
    out->is_ref                   = cfg->is_ref;
    out->num_of_list_to_search    = 1;
    out->num_of_ref_pic_to_search[0] = 1;
    out->max_refs = out->max_cand = out->max_l0 = 1;
    free(m); free(scs); free(pcs);
    return 0;
}

/* Dynamic-GOP detector: the reference's dg_detector_hme_level0 (Codec/pd_process.c:492) run as ONE segment
 * covering the picture, on a PictureParentControlSet that holds only what the function reads; early_hme's
 * metric reset (:569-573) is done here. The mutex and the semaphore are the reference's own (svt_threads.c). */
void dg_detector_hme_level0(struct PictureParentControlSet *ppcs, uint32_t seg_idx);

REF_API int ref_dg_detector(const OraclePicture *src, const OraclePicture *ref, uint8_t input_resolution, MeGpuDgMetrics *out,
                            int simd) {
    set_rtcd(simd);
    PictureParentControlSet *pcs = calloc(1, sizeof *pcs), *rpcs = calloc(1, sizeof *rpcs);
    SequenceControlSet      *scs = calloc(1, sizeof *scs);
    DGDetectorSeg           *dg  = calloc(1, sizeof *dg);
    EbObjectWrapper         *sw = calloc(1, sizeof *sw), *rw = calloc(1, sizeof *rw);
    EbPaReferenceObject     *so = calloc(1, sizeof *so), *ro = calloc(1, sizeof *ro);
    if (!pcs || !rpcs || !scs || !dg || !sw || !rw || !so || !ro)
        return -1;
    EbPictureBufferDesc s16, r16;
    desc_from_plane(&s16, &src->sixteenth);
    desc_from_plane(&r16, &ref->sixteenth);
    so->sixteenth_downsampled_picture_ptr = &s16;
    ro->sixteenth_downsampled_picture_ptr = &r16;
    sw->object_ptr = so;
    rw->object_ptr = ro;
    pcs->pa_ref_pic_wrapper  = sw;
    rpcs->pa_ref_pic_wrapper = rw;
    pcs->scs                 = scs;
    scs->b64_size            = 64;
    pcs->input_resolution    = (EbInputResolution)input_resolution;
    pcs->aligned_width       = (uint16_t)((src->full.width + 7) & ~7);
    pcs->aligned_height      = (uint16_t)((src->full.height + 7) & ~7);
    pcs->me_segments_column_count = pcs->me_segments_row_count = 1;
    pcs->dg_detector         = dg;
    dg->ref_pic              = rpcs;
    dg->metrics_mutex        = svt_create_mutex();
    dg->frame_done_sem       = svt_create_semaphore(0, 1);
    if (!dg->metrics_mutex || !dg->frame_done_sem)
        return -1;
    dg_detector_hme_level0(pcs, 0);
    svt_block_on_semaphore(dg->frame_done_sem); /* posted by the last (only) segment */
    out->tot_dist       = dg->metrics.tot_dist;
    out->tot_cplx       = dg->metrics.tot_cplx;
    out->tot_active     = dg->metrics.tot_active;
    out->sum_in_vectors = dg->metrics.sum_in_vectors;
    svt_destroy_mutex(dg->metrics_mutex);
    svt_destroy_semaphore(dg->frame_done_sem);
    free(ro); free(so); free(rw); free(sw); free(dg); free(scs); free(rpcs); free(pcs);
    return 0;
}

/* ------------------------------------------------------------------------------------------------------------------
 * Host logic of the binding on the CPU (tests/test_hook_host_logic.py): the hook bound to tests/emul/libme_fake_engine.so.
 * Returns 0, or the (negative) number of the first step whose expectation failed. report[0..7] = the fake engine's counters
 * at the end, report[8] = pictures it still holds.
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct HostLogicPic {
    PictureParentControlSet pcs;
    EbObjectWrapper         wrap;
    EbPictureBufferDesc     full, quarter, sixteenth;
    uint8_t                 luma[64], q[64], s[64];
} HostLogicPic;

static void host_logic_pic(HostLogicPic *p, EbPaReferenceObject *obj, SequenceControlSet *scs, uint64_t poc) {
    memset(p, 0, sizeof *p);
    p->wrap.object_ptr         = obj;
    p->pcs.pa_ref_pic_wrapper  = &p->wrap;
    p->pcs.scs                 = scs;
    p->pcs.picture_number      = poc;
    p->pcs.aligned_width       = 128;
    p->pcs.aligned_height      = 64;
    p->pcs.tf_segments_total_count = 8;
    obj->picture_number        = poc;
    p->full.buffer_y = p->luma; p->full.stride_y = 8;
    p->quarter.buffer_y = p->q; p->quarter.stride_y = 8;
    p->sixteenth.buffer_y = p->s; p->sixteenth.stride_y = 8;
    obj->input_padded_pic                  = &p->full;
    obj->quarter_downsampled_picture_ptr   = &p->quarter;
    obj->sixteenth_downsampled_picture_ptr = &p->sixteenth;
}

typedef struct HostLogicTf {
    HostLogicPic *centre, *nb[2];
    int           tid, bad;
    uint32_t      horz, vert;
} HostLogicTf;

static void *host_logic_tf_thread(void *arg) {
    HostLogicTf *a = (HostLogicTf *)arg;
    MeContext   *m = calloc(1, sizeof *m);
    const uint32_t nb64 = 2; /* 128 x 64 */
    for (uint32_t b = 0; b < nb64; b++)
        for (int f = 0; f < 2; f++) {
            if (!svt_me_gpu_tf_ready(&a->centre->pcs, &a->nb[f]->pcs)) { a->bad = 1; continue; }
            m->tf_use_pred_64x64_only_th = 1;
            if (svt_me_gpu_tf_block(&a->centre->pcs, m, &a->nb[f]->pcs, 2 * f, nb64, b) != EB_ErrorNone) { a->bad = 2; continue; }
            /* the fake engine's pattern: (registration count of the neighbour) * 100000 + block * 100 + PU */
            const uint32_t v = m->p_sb_best_sad[0][0][3];
            if (v % 100000u != b * 100u + 3u || v / 100000u == 0) a->bad = 3;
            if (m->p_best_sad_32x32 != &m->p_sb_best_sad[0][0][1] || m->search_results[0][0].hme_sc_x != (int16_t)b) a->bad = 4;
            if ((m->tf_use_pred_64x64_only_th == (uint8_t)~0) != (int)(b & 1)) a->bad = 5;
        }
    a->horz = m->tf_tot_horz_blks;
    a->vert = m->tf_tot_vert_blks;
    svt_me_gpu_tf_centre_done(&a->centre->pcs);
    free(m);
    return NULL;
}

typedef struct HostLogicSeg {
    PictureParentControlSet *pcs;
    EbErrorType              rc;
} HostLogicSeg;

static void *host_logic_seg_thread(void *arg) { /* one TASK_PAME segment item of me_process.c:215-262 */
    HostLogicSeg *a = (HostLogicSeg *)arg;
    MeContext    *m = calloc(1, sizeof *m);
    m->me_type                     = ME_OPEN_LOOP;
    m->num_of_list_to_search       = 2;
    m->num_of_ref_pic_to_search[0] = m->num_of_ref_pic_to_search[1] = 1;
    for (int l = 0; l < 2; l++)
        m->me_ds_ref_array[l][0].picture_number = ((EbPaReferenceObject *)a->pcs->ref_pa_pic_ptr_array[l][0]->object_ptr)->picture_number;
    a->rc = svt_me_gpu_segment(a->pcs, m);
    free(m);
    return NULL;
}

REF_API int ref_hook_host_logic(const char *fake_library_path, uint64_t *report) {
    typedef int (*CountsFn)(uint64_t *);
    uint64_t            c[8];
    SequenceControlSet *scs = calloc(1, sizeof *scs);
    EbPaReferenceObject obj[3];
    HostLogicPic       *p = calloc(5, sizeof *p);
    DGDetectorSeg       dg;
    int                 rc = 0;
    memset(obj, 0, sizeof obj);
    memset(&dg, 0, sizeof dg);
    if (svt_me_gpu_hook_init(fake_library_path, 128, 64, 4, 3, 8) != EB_ErrorNone)
        return -1;
    void    *h = dlopen(fake_library_path, RTLD_NOW | RTLD_NOLOAD);
    CountsFn counts = h ? (CountsFn)dlsym(h, "fake_engine_counts") : NULL;
    if (!counts)
        return -2;
#define STEP(n, cond) do { if (!rc && !(cond)) rc = -(n); } while (0)
    /* 10: picture analysis registers a picture by its PA reference object and gets the two planes back */
    host_logic_pic(&p[0], &obj[0], scs, 0);
    STEP(10, svt_me_gpu_downsample(&p[0].pcs, &p[0].full, &p[0].quarter, &p[0].sixteenth) == EB_ErrorNone);
    counts(c);
    STEP(11, c[0] == 1 && c[1] == 0 && c[5] == 2 && p[0].q[0] == 1 * 16 + 1 && p[0].s[0] == 1 * 16 + 2);
    /* 20: the temporal filter rewrote the picture: same object again -> the first version is released, a new identity */
    STEP(20, svt_me_gpu_downsample(&p[0].pcs, &p[0].full, &p[0].quarter, &p[0].sixteenth) == EB_ErrorNone);
    counts(c);
    STEP(21, c[0] == 2 && c[1] == 1 && p[0].q[0] == 2 * 16 + 1 && svt_me_gpu_downsample_count() == 2);
    /* 30: two more pictures; then the pool hands object 0 to a new picture (POC 7): the old pyramid goes */
    host_logic_pic(&p[1], &obj[1], scs, 1);
    host_logic_pic(&p[2], &obj[2], scs, 2);
    STEP(30, svt_me_gpu_downsample(&p[1].pcs, &p[1].full, &p[1].quarter, &p[1].sixteenth) == EB_ErrorNone);
    STEP(31, svt_me_gpu_downsample(&p[2].pcs, &p[2].full, &p[2].quarter, &p[2].sixteenth) == EB_ErrorNone);
    host_logic_pic(&p[3], &obj[0], scs, 7);
    STEP(32, svt_me_gpu_downsample(&p[3].pcs, &p[3].full, &p[3].quarter, &p[3].sixteenth) == EB_ErrorNone);
    STEP(33, counts(c) == 3 && c[0] == 5 && c[1] == 2);
    /* 40: the detector gets the identities of the two OBJECTS (registration counts 3 and 4), not the picture numbers */
    EbPaReferenceObject stranger;
    memset(&stranger, 0, sizeof stranger);
    host_logic_pic(&p[4], &stranger, scs, 9); /* never downsampled */
    STEP(40, svt_me_gpu_dg_ready(&p[1].pcs, &p[2].pcs) && !svt_me_gpu_dg_ready(&p[1].pcs, &p[4].pcs));
    p[1].pcs.dg_detector = &dg;
    dg.ref_pic           = &p[2].pcs;
    STEP(41, svt_me_gpu_dg_detector(&p[1].pcs) == EB_ErrorNone && dg.metrics.tot_dist == 3 && dg.metrics.tot_cplx == 4 &&
                 dg.metrics.tot_active == 1 && dg.metrics.sum_in_vectors == 2);
    /* 50: temporal filter of centre 1 with neighbours 2 and 7 (object 0): eight segment threads, two blocks, two pairs ->
       exactly two jobs; every thread sees the pair's results; the totals are handed out once per pair */
    pthread_t   th[8];
    HostLogicTf a[8];
    for (int i = 0; i < 8; i++) {
        a[i] = (HostLogicTf){&p[1], {&p[2], &p[3]}, i, 0, 0, 0};
        pthread_create(&th[i], NULL, host_logic_tf_thread, &a[i]);
    }
    uint32_t horz = 0, vert = 0;
    int      bad  = 0;
    for (int i = 0; i < 8; i++) {
        pthread_join(th[i], NULL);
        horz += a[i].horz; vert += a[i].vert; bad |= a[i].bad;
    }
    counts(c);
    STEP(50, !bad);
    STEP(51, c[2] == 2 && c[4] == 2 && horz == 10 && vert == 14);
    /* 60: all eight segments are through: the pairs are gone, the same pair asked again is a new job */
    a[0] = (HostLogicTf){&p[1], {&p[2], &p[3]}, 0, 0, 0, 0};
    p[1].pcs.tf_segments_total_count = 1;
    host_logic_tf_thread(&a[0]);
    counts(c);
    STEP(60, !a[0].bad && c[2] == 4 && c[4] == 4);
    /* 70: open-loop ME of picture 1 with references 2 (list 0) and 7 (list 1): its six segment items arrive on six threads;
       ONE submit and ONE wait, every thread returns after the results are in the pcs; nothing is uploaded again */
    {
        PictureParentControlSet *pcs = &p[1].pcs;
        MotionEstimationData     md;
        MeSbResults              res[2], *rp[2] = {&res[0], &res[1]};
        uint8_t                  idx[2][85], cand[2][85 * 3 * sizeof(MeCandidate)];
        MvCandidate              mv[2][85 * 2];
        uint32_t                 arr[6][2];
        uint8_t                  flags[2][2];
        memset(&md, 0, sizeof md);
        for (int b = 0; b < 2; b++) {
            res[b].total_me_candidate_index = idx[b];
            res[b].me_candidate_array       = (MeCandidate *)cand[b];
            res[b].me_mv_array              = mv[b];
        }
        md.me_results = rp; md.max_cand = 3; md.max_refs = 2; md.max_l0 = 1;
        pcs->pa_me_data = &md;
        pcs->enable_me_8x8 = pcs->enable_me_16x16 = 1;
        pcs->max_number_of_pus_per_sb = 85;
        pcs->me_64x64_distortion = arr[0]; pcs->me_32x32_distortion = arr[1]; pcs->me_16x16_distortion = arr[2];
        pcs->me_8x8_distortion = arr[3]; pcs->me_8x8_cost_variance = arr[4]; pcs->rc_me_distortion = arr[5];
        pcs->stationary_block_present_sb = flags[0]; pcs->rc_me_allow_gm = flags[1];
        pcs->me_segments_total_count = 6;
        pcs->ref_pa_pic_ptr_array[0][0] = &p[2].wrap;
        pcs->ref_pa_pic_ptr_array[1][0] = &p[3].wrap;
        uint64_t     before[8];
        pthread_t    ts[6];
        HostLogicSeg sa[6];
        counts(before);
        for (int i = 0; i < 6; i++) {
            sa[i].pcs = pcs; sa[i].rc = EB_ErrorUndefined;
            pthread_create(&ts[i], NULL, host_logic_seg_thread, &sa[i]);
        }
        int ok = 1;
        for (int i = 0; i < 6; i++) {
            pthread_join(ts[i], NULL);
            ok &= sa[i].rc == EB_ErrorNone;
        }
        counts(c);
        STEP(70, ok);
        STEP(71, c[2] == before[2] + 1 && c[3] == before[3] + 1 && c[0] == before[0]);
        STEP(72, idx[1][84] == (uint8_t)(85 + 84) && idx[0][3] == 3); /* scattered into the per-b64 MeSbResults */
    }
    STEP(61, c[7] == 0); /* nothing the engine had to reject */
#undef STEP
    report[8] = (uint64_t)counts(report);
    svt_me_gpu_hook_shutdown();
    free(p); free(scs);
    return rc;
}
