/*
 * me_derive.c - host-side derivation of the ME controls (product host code, plain C).
 *
 * Mirrors, for the open-loop ME path only, what the encoder computes before the ME kernel runs:
 *   svt_aom_sig_deriv_me               Codec/enc_mode_config.c:681-828
 *   set_me_search_params               Codec/enc_mode_config.c:223-350
 *   set_hme_search_params              Codec/enc_mode_config.c:138-218
 *   svt_aom_set_me_hme_ref_prune_ctrls Codec/enc_mode_config.c:351-430
 *   svt_aom_set_mv_based_sa_ctrls      Codec/enc_mode_config.c:432-451
 *   svt_aom_set_me_sr_adjustment_ctrls Codec/enc_mode_config.c:453-520
 *   svt_aom_set_me_8x8_var_ctrls       Codec/enc_mode_config.c:521-540
 *   svt_aom_set_prehme_ctrls           Codec/enc_mode_config.c:542-593
 *   HME level flags                    Codec/enc_mode_config.c:1633-1645
 *   svt_aom_get_enable_me_8x8          Codec/enc_mode_config.c:76-96
 *   svt_aom_get_gm_core_level / svt_aom_set_gm_controls  Codec/enc_mode_config.c:103-116,1864-1936
 *   svt_aom_derive_input_resolution    Codec/sequence_control_set.c:112-130
 *   mrp level + set_mrp_ctrl           Globals/enc_handle.c:4193-4230,3397-3670
 *   svt_aom_get_max_allocated_me_refs  Codec/pcs.c:91-96, pd_process.c:3500-3519
 * In an integrated encoder these values come from the encoder's own structures; this function exists so
 * that a standalone host (bench, tests, other front ends) produces identical controls. It is pinned
 * against the reference's own svt_aom_sig_deriv_me by tests/test_derive_params.py.
 */
#include <string.h>

#include "me_gpu.h"

#define ENC_MRS (-3)
#define ENC_MRP (-2)
#define ENC_MR (-1)
#define ENC_M0 0
#define ENC_M1 1
#define ENC_M2 2
#define ENC_M3 3
#define ENC_M4 4
#define ENC_M5 5
#define ENC_M6 6
#define ENC_M7 7
#define ENC_M8 8
#define ENC_M9 9

enum { RES_240p, RES_360p, RES_480p, RES_720p, RES_1080p, RES_4K, RES_8K };

#define DMAX(a, b) ((a) > (b) ? (a) : (b))
#define DCLIP3(lo, hi, v) ((v) < (lo) ? (lo) : (v) > (hi) ? (hi) : (v))

static MeGpuSearchArea sa(uint16_t w, uint16_t h) {
    MeGpuSearchArea a;
    a.width  = w;
    a.height = h;
    return a;
}

static uint8_t input_resolution(uint32_t size) {
    if (size < 0x28500) return RES_240p;
    if (size < 0x4CE00) return RES_360p;
    if (size < 0xA1400) return RES_480p;
    if (size < 0x16DA00) return RES_720p;
    if (size < 0x535200) return RES_1080p;
    if (size < 0x140A000) return RES_4K;
    return RES_8K;
}

typedef struct Mrp {
    uint8_t base0, base1, nbase0, nbase1, sc_base0, sc_base1, sc_nbase0, sc_nbase1;
    uint8_t safe_limit_nref;
    uint32_t safe_limit_zz_th;
    uint8_t only_l_bwd;
} Mrp;

static Mrp mrp_ctrl(int level) {
    /* base0 base1 nbase0 nbase1 | sc: base0 base1 nbase0 nbase1 | safe_nref zz_th only_l_bwd */
    static const Mrp t[13] = {
        {1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0},     {4, 3, 4, 3, 2, 2, 2, 2, 0, 0, 0},     {4, 3, 4, 3, 2, 2, 2, 2, 0, 0, 1},
        {4, 3, 4, 3, 2, 2, 2, 2, 0, 0, 1},     {4, 3, 4, 3, 2, 2, 1, 1, 0, 0, 1},     {4, 3, 4, 3, 2, 2, 1, 1, 0, 0, 1},
        {4, 3, 4, 3, 2, 2, 1, 1, 1, 60000, 1}, {3, 2, 3, 2, 2, 2, 1, 1, 1, 60000, 1}, {3, 2, 3, 2, 2, 2, 1, 1, 2, 0, 1},
        {3, 2, 3, 2, 2, 2, 1, 1, 2, 0, 1},     {3, 2, 2, 2, 2, 2, 1, 1, 2, 0, 1},     {3, 2, 1, 1, 2, 2, 1, 1, 2, 0, 1},
        {2, 2, 1, 1, 2, 2, 1, 1, 2, 0, 1}};
    return t[level < 0 || level > 12 ? 0 : level];
}

static int mrp_level(const MeGpuEncoderConfig *c, uint8_t res) {
    if (c->rtc_tune)
        return c->enc_mode <= ENC_M9 ? 10 : 11;
    const int fd_hi_res = !(res <= RES_360p) && !(c->fast_decode <= 1);
    if (c->enc_mode <= ENC_M2)
        return fd_hi_res ? 9 : 2;
    if (c->enc_mode <= ENC_M4)
        return fd_hi_res ? 9 : 5;
    if (!c->rate_control_vbr)
        return c->enc_mode <= ENC_M7 ? 9 : c->enc_mode <= ENC_M9 ? 10 : 0;
    return 12;
}

static void set_me_search_params(const MeGpuEncoderConfig *c, uint8_t res, MeGpuParams *p) {
    const int m = c->enc_mode;
    uint8_t   q_mult = 0;
    MeGpuSearchArea mn, mx;
    if (c->rtc_tune) {
        if (c->sc_class1) {
            if (m <= ENC_M7) { mn = sa(32, 32); mx = sa(96, 96); }
            else if (res < RES_1080p) { mn = sa(16, 16); mx = sa(32, 16); }
            else if (m <= ENC_M9) { mn = sa(24, 24); mx = sa(24, 24); }
            else { mn = sa(16, 6); mx = sa(16, 9); }
        } else if (m <= ENC_M8) {
            if (res < RES_1080p) { mn = sa(16, 16); mx = sa(32, 16); }
            else { mn = sa(16, 6); mx = sa(16, 9); }
        } else {
            if (res < RES_720p) { mn = sa(8, 3); mx = sa(16, 9); }
            else if (res < RES_1080p) { mn = sa(8, 1); mx = sa(16, 7); }
            else if (res < RES_4K) { mn = sa(8, 1); mx = sa(8, 7); }
            else { mn = sa(8, 1); mx = sa(8, 1); }
        }
    } else if (c->sc_class1) {
        if (m <= ENC_M1) { mn = sa(175, 175); mx = sa(750, 750); }
        else if (m <= ENC_M6) { mn = sa(48, 48); mx = sa(224, 224); }
        else if (m <= ENC_M7) { mn = sa(32, 32); mx = sa(164, 164); }
        else if (m <= ENC_M8) { mn = sa(32, 32); mx = sa(96, 96); }
        else if (m <= ENC_M9) { mn = sa(16, 16); mx = sa(96, 96); }
        else { mn = sa(8, 8); mx = sa(32, 32); }
    } else if (m <= ENC_M1) { mn = sa(64, 64); mx = sa(256, 256); }
    else if (m <= ENC_M2) { mn = sa(32, 32); mx = sa(128, 128); }
    else if (m <= ENC_M4) { mn = sa(24, 24); mx = sa(104, 104); }
    else if (m <= ENC_M6) { mn = sa(16, 16); mx = sa(64, 32); q_mult = 7; }
    else if (m <= ENC_M9) {
        if (c->hierarchical_levels <= 3) {
            if (res < RES_4K) { mn = sa(8, 5); mx = sa(16, 9); }
            else { mn = sa(8, 1); mx = sa(8, 1); }
        } else if (res < RES_1080p) { mn = sa(16, 16); mx = sa(32, 16); }
        else { mn = sa(16, 6); mx = sa(16, 9); }
        q_mult = 7;
    } else { mn = sa(16, 6); mx = sa(16, 6); q_mult = 6; }
    if (q_mult) {
        uint16_t qw = (uint16_t)DCLIP3(500, 1000, ((int)(q_mult * ((31 * c->qp) - 700)) >> 3));
        mn.width  = (uint16_t)DMAX(8, (mn.width * qw) / 1000);
        mn.height = (uint16_t)DMAX(3, (mn.height * qw) / 1000);
        mx.width  = (uint16_t)DMAX(8, (mx.width * qw) / 1000);
        mx.height = (uint16_t)DMAX(3, (mx.height * qw) / 1000);
    }
    if (c->low_frame_rate) {
        mn.width  = (uint16_t)((mn.width * 3) >> 1);
        mn.height = (uint16_t)((mn.height * 3) >> 1);
    }
    p->me_sa.sa_min = mn;
    p->me_sa.sa_max = mx;
}

static void set_hme_search_params(const MeGpuEncoderConfig *c, uint8_t res, MeGpuParams *p) {
    const int m = c->enc_mode;
    uint8_t   q_mult = 0;
    MeGpuSearchArea mn, mx = sa(192, 192);
    p->num_hme_sa_w = p->num_hme_sa_h = 2;
    if (m <= ENC_MRS) {
        if (res < RES_4K) { mn = sa(128, 128); mx = sa(256, 256); }
        else { mn = sa(240, 240); mx = sa(480, 480); }
    } else if (m <= ENC_M1) {
        if (res < RES_4K) mn = sa(32, 32);
        else { mn = sa(240, 240); mx = sa(480, 480); }
    } else if (m <= ENC_M3) mn = sa(32, 32);
    else if (m <= ENC_M6) { mn = sa(32, 32); q_mult = 3; }
    else if (!c->rtc_tune && m <= ENC_M7) {
        mn     = (c->sc_class1 || res >= RES_4K) ? sa(32, 32) : sa(16, 16);
        q_mult = 3;
    } else if ((!c->rtc_tune && m <= ENC_M9) || (c->rtc_tune && m <= ENC_M7)) {
        mn     = c->sc_class1 ? sa(32, 32) : sa(16, 16);
        q_mult = 3;
    } else {
        if (c->sc_class1) mn = sa(32, 32);
        else if (res < RES_4K) { mn = sa(8, 8); mx = sa(96, 96); }
        else { mn = sa(16, 16); mx = sa(96, 96); }
        q_mult = 3;
    }
    if (q_mult) {
        uint16_t qw = (uint16_t)DCLIP3(500, 1000, ((int)(q_mult * ((8 * c->qp) - 125))));
        mn.width  = (uint16_t)DMAX(8, (mn.width * qw) / 1000);
        mn.height = (uint16_t)DMAX(8, (mn.height * qw) / 1000);
        mx.width  = (uint16_t)DMAX(96, (mx.width * qw) / 1000);
        mx.height = (uint16_t)DMAX(96, (mx.height * qw) / 1000);
    }
    p->hme_l0_sa.sa_min = mn;
    p->hme_l0_sa.sa_max = mx;
    p->hme_l1_sa = p->hme_l2_sa = (m <= ENC_MR) ? sa(16, 16) : sa(8, 3);
}

static void set_ref_prune(MeGpuParams *p, int level) {
    static const uint16_t hme_dev[7] = {0xFFFF, 80, 50, 30, 15, 5, 5};
    p->enable_me_hme_ref_pruning               = level != 0;
    p->prune_ref_if_hme_sad_dev_bigger_than_th = hme_dev[level];
    p->prune_ref_if_me_sad_dev_bigger_than_th  = level >= 4 ? 60 : 0xFFFF;
    p->zz_sad_th    = level == 6 ? 20 * 64 * 64 : 0;
    p->zz_sad_pct   = level == 6 ? 5 : 0;
    p->phme_sad_th  = level == 6 ? 10 * 64 * 64 : 0;
    p->phme_sad_pct = level == 6 ? 5 : 0;
}

static void set_sr_adjustment(MeGpuParams *p, int level) {
    /* enable, mv_length_th, stationary_sad_th, low_sad_th, distance_based */
    static const uint16_t t[6][5] = {{0, 0, 0, 0, 0},         {1, 4, 12000, 6000, 0},   {1, 4, 12000, 6000, 1},
                                     {1, 4, 12000, 12000, 1}, {2, 16, 20000, 20000, 1}, {2, 20, 24000, 24000, 1}};
    p->enable_me_sr_adjustment = (uint8_t)t[level][0];
    if (level) {
        p->reduce_me_sr_based_on_mv_length_th   = t[level][1];
        p->stationary_hme_sad_abs_th            = t[level][2];
        p->stationary_me_sr_divisor             = 8;
        p->reduce_me_sr_based_on_hme_sad_abs_th = t[level][3];
        p->me_sr_divisor_for_low_hme_sad        = 8;
        p->distance_based_hme_resizing          = (uint8_t)t[level][4];
    }
    if (!p->enable_hme_level2_flag) {
        const int d = p->enable_hme_level1_flag ? 4 : 16;
        p->stationary_hme_sad_abs_th            = (uint16_t)(p->stationary_hme_sad_abs_th / d);
        p->reduce_me_sr_based_on_hme_sad_abs_th = (uint16_t)(p->reduce_me_sr_based_on_hme_sad_abs_th / d);
    }
}

static void set_prehme(MeGpuParams *p, int level) {
    p->prehme_enable = level != 0;
    if (!level)
        return;
    static const uint16_t t[5][8] = {{0},
                                     {8, 144, 8, 496, 144, 3, 496, 3},
                                     {8, 100, 8, 400, 96, 3, 384, 3},
                                     {8, 100, 8, 350, 32, 7, 200, 7},
                                     {8, 100, 8, 350, 32, 7, 128, 7}};
    p->prehme_sa_cfg[0].sa_min = sa(t[level][0], t[level][1]);
    p->prehme_sa_cfg[0].sa_max = sa(t[level][2], t[level][3]);
    p->prehme_sa_cfg[1].sa_min = sa(t[level][4], t[level][5]);
    p->prehme_sa_cfg[1].sa_max = sa(t[level][6], t[level][7]);
    p->prehme_skip_search_line = level >= 3;
    p->prehme_l1_early_exit    = level == 4;
}

int me_gpu_derive_params(const MeGpuEncoderConfig *c, MeGpuParams *p) {
    if (!c || !p || (c->width & 7) || (c->height & 7) || !c->width || !c->height)
        return ME_GPU_ERR_BAD_ARG;
    memset(p, 0, sizeof *p);
    const int     m       = c->enc_mode;
    const uint8_t res     = input_resolution((uint32_t)c->width * c->height);
    const int     is_base = c->temporal_layer_index == 0;
    const Mrp     mrp     = mrp_ctrl(mrp_level(c, res));

    /* ---- pcs / scs level ---- */
    p->picture_number           = c->picture_number;
    p->input_resolution         = res;
    p->enable_me_16x16          = 1;
    p->enable_me_8x8            = c->rtc_tune ? (m <= ENC_M7) : (m <= ENC_M5 ? 1 : m <= ENC_M8 ? res <= RES_720p : 0);
    p->max_number_of_pus_per_sb = ME_GPU_SQUARE_PU_COUNT;
    /* gm level: 0 for I slices (never reach ME); MRP 1, MR 2, <=M4 3, else off. Only `enabled` and
     * use_distance_based_active_th (0 at every level) reach the ME path. */
    p->gm_enabled                      = m <= ENC_M4;
    p->gm_use_distance_based_active_th = 0;
    p->hierarchical_levels             = c->hierarchical_levels;
    p->pcs_temporal_layer_index        = c->temporal_layer_index;
    p->similar_brightness_refs         = c->similar_brightness_refs;
    p->only_l_bwd                      = mrp.only_l_bwd;
    {
        uint8_t l0 = DMAX(DMAX(mrp.sc_base0, mrp.base0), DMAX(mrp.sc_nbase0, mrp.nbase0));
        uint8_t l1 = DMAX(DMAX(mrp.sc_base1, mrp.base1), DMAX(mrp.sc_nbase1, mrp.nbase1));
        uint8_t n0 = c->sc_class1 ? (is_base ? mrp.sc_base0 : mrp.sc_nbase0) : (is_base ? mrp.base0 : mrp.nbase0);
        uint8_t n1 = c->sc_class1 ? (is_base ? mrp.sc_base1 : mrp.sc_nbase1) : (is_base ? mrp.base1 : mrp.nbase1);
        if (c->force_ref_l0 > 0) {
            n0 = (uint8_t)c->force_ref_l0;
            n1 = (uint8_t)DMAX(0, c->force_ref_l1);
            /* a sequence configured with these counts sizes its ME result pool by them (pcs.c:91-96 from the mrp maxima) */
            l0 = n0;
            l1 = n1;
        }
        p->max_refs = (uint8_t)(l0 + l1);
        p->max_cand = (uint8_t)(l0 + l1 + l0 * l1 + (l0 - 1) + (l1 == 3 ? 1 : 0));
        p->max_l0   = l0;
        p->num_of_list_to_search       = c->slice_type == 1 /*P_SLICE*/ ? 1 : 2;
        p->num_of_ref_pic_to_search[0] = n0;
        p->num_of_ref_pic_to_search[1] = p->num_of_list_to_search == 2 ? n1 : 0;
    }
    p->temporal_layer_index = c->temporal_layer_index;
    p->is_ref               = c->is_ref;

    /* ---- svt_aom_sig_deriv_me ---- */
    set_me_search_params(c, res, p);
    set_hme_search_params(c, res, p);
    p->enable_hme_flag        = 1;
    p->enable_hme_level0_flag = 1;
    p->enable_hme_level1_flag = 1;
    p->enable_hme_level2_flag = (c->sc_class1 || m <= ENC_M6) ? 1 : 0;
    p->hme_search_method      = ME_GPU_SUB_SAD_SEARCH;
    p->me_search_method       = ME_GPU_SUB_SAD_SEARCH;
    if (c->rtc_tune && c->sc_class1) {
        p->reduce_hme_l0_sr_th_min = 8;
        p->reduce_hme_l0_sr_th_max = 100;
    } else if (c->rtc_tune && m > ENC_M7) {
        p->reduce_hme_l0_sr_th_min = 8;
        p->reduce_hme_l0_sr_th_max = 200;
    }
    int prehme = 0;
    if (m <= ENC_MRS) prehme = 1;
    else if (c->sc_class1) prehme = c->rtc_tune ? 1 : 2;
    else if (c->rtc_tune) prehme = m <= ENC_M9 ? 4 : 0;
    else prehme = m <= ENC_M7 ? 2 : 4;
    if (!p->enable_hme_level1_flag) prehme = 0;
    set_prehme(p, prehme);

    int prune = 0;
    if (c->sc_class1) prune = m <= ENC_MRS ? 0 : m <= ENC_M2 ? 1 : m <= ENC_M7 ? 3 : 6;
    else if (m <= ENC_MRS) prune = 0;
    else if (m <= ENC_MR) prune = 1;
    else if (m <= ENC_M0) prune = is_base ? 1 : 2;
    else if (m <= ENC_M1) prune = is_base ? 1 : 4;
    else if (m <= ENC_M3) prune = is_base ? 1 : 5;
    else if (m <= ENC_M9) prune = is_base ? 1 : 6;
    else prune = 6;
    set_ref_prune(p, prune);

    int sr = 0;
    if (c->sc_class1) sr = m <= ENC_M7 ? 4 : 5;
    else if (m <= ENC_MR) sr = 0;
    else if (m <= ENC_M0) sr = 1;
    else sr = 3;
    set_sr_adjustment(p, sr);

    const int mv_adj = m <= ENC_MRS ? 1 : m <= ENC_M3 ? 2 : 0;
    if (mv_adj) {
        p->mv_sa_adj_enabled          = 1;
        p->mv_sa_adj_nearest_ref_only = mv_adj == 2;
        p->mv_sa_adj_mv_size_th       = 25;
        p->mv_sa_adj_sa_multiplier    = 2;
    }
    p->me_8x8_var_enabled = 1; /* level 2 */
    p->me_sr_div4_th      = 80000;
    p->me_sr_div2_th      = 150000;
    p->me_sr_mult2_th     = 0xFFFFFFFFu;
    p->prune_me_candidates_th     = m <= ENC_M6 ? 0 : 65;
    p->use_best_unipred_cand_only = m <= ENC_M3 ? 0 : 1;
    if (c->rtc_tune)
        p->me_early_exit_th = c->sc_class1 ? 64 * 64 : m <= ENC_M9 ? 64 * 64 * 8 : 64 * 64 * 9;
    else
        p->me_early_exit_th = m <= ENC_M4 ? 0 : 64 * 64 * 8;
    p->me_safe_limit_zz_th         = mrp.safe_limit_nref == 1 ? mrp.safe_limit_zz_th : 0;
    p->prev_me_stage_based_exit_th = (c->rtc_tune && c->sc_class1) ? 64 * 64 * 4 : 0;
    p->me_type                     = ME_GPU_ME_OPEN_LOOP; /* me_process.c:215 */
    return ME_GPU_OK;
}

/* svt_aom_sig_deriv_me_tf (enc_mode_config.c:834-874) with tf_set_me_hme_params_oq (:598-675), the tf_enable_hme_*
 * switch of svt_aom_sig_deriv_multi_processes (:1646-1668), set_hme_search_params_mctf(ctx, 0)
 * (temporal_filtering.c:2760-2768) and the per-call fields of produce_temporally_filtered_pic (:3144-3181). */
int me_gpu_derive_params_tf(const MeGpuTfConfig *c, MeGpuParams *p) {
    if (!c || !p || (c->width & 7) || (c->height & 7) || !c->width || !c->height || c->hme_me_level > 4)
        return ME_GPU_ERR_BAD_ARG;
    memset(p, 0, sizeof *p);
    const uint8_t res = input_resolution((uint32_t)c->width * c->height);
    const int     lvl = c->hme_me_level;
    p->me_type          = ME_GPU_ME_MCTF;
    p->tf_me_exit_th    = c->me_exit_th;
    p->picture_number   = c->picture_number;
    p->input_resolution = res;
    /* the centre picture's pcs is handed to the call; candidate construction never runs, so the export switches only
     * keep their defaults */
    p->enable_me_16x16 = p->enable_me_8x8 = 1;
    p->max_number_of_pus_per_sb           = ME_GPU_SQUARE_PU_COUNT;
    p->pcs_temporal_layer_index = p->temporal_layer_index = c->temporal_layer_index;
    p->is_ref                      = c->is_ref;
    p->num_of_list_to_search       = 1;
    p->num_of_ref_pic_to_search[0] = 1;
    p->max_refs = p->max_cand = p->max_l0 = 1;
    p->num_hme_sa_w = p->num_hme_sa_h = 2;
    /* hme_l0_sa = hme_l0_sa_default_tf, hme_l1_sa, hme_l2_sa, me_sa per level */
    static const uint16_t t[5][6] = {/* l0 min, l0 max, l1, l2, me min, me max */
                                     {30, 60, 16, 16, 60, 120}, {16, 32, 16, 16, 16, 32}, {16, 32, 16, 16, 8, 8},
                                     {8, 8, 8, 8, 8, 8},        {4, 4, 8, 8, 8, 8}};
    uint16_t l0min = t[lvl][0], l0max = t[lvl][1], l1 = t[lvl][2];
    if (lvl == 2 && res <= RES_360p) {
        l0min = l0max = l1 = 8;
    } else if (lvl == 2 && res <= RES_480p) {
        l0min = 8;
        l0max = 16;
        l1    = 8;
    }
    p->hme_l0_sa.sa_min = sa(l0min, l0min);
    p->hme_l0_sa.sa_max = sa(l0max, l0max);
    p->hme_l1_sa        = sa(l1, l1);
    p->hme_l2_sa        = sa(t[lvl][3], t[lvl][3]);
    p->me_sa.sa_min     = sa(t[lvl][4], t[lvl][4]);
    p->me_sa.sa_max     = sa(t[lvl][5], t[lvl][5]);
    if (c->qp_opt) { /* enc_mode_config.c:667-674 */
        int q = 8 * c->qp - 125;
        q     = q < 250 ? 250 : q > 1000 ? 1000 : q;
        const uint16_t qw = (uint16_t)q;
        p->me_sa.sa_min.width  = (uint16_t)DMAX(8, (p->me_sa.sa_min.width * qw) / 1000);
        p->me_sa.sa_min.height = (uint16_t)DMAX(8, (p->me_sa.sa_min.height * qw) / 1000);
        p->me_sa.sa_max.width  = (uint16_t)DMAX(8, (p->me_sa.sa_max.width * qw) / 1000);
        p->me_sa.sa_max.height = (uint16_t)DMAX(8, (p->me_sa.sa_max.height * qw) / 1000);
    }
    p->enable_hme_flag        = 1;
    p->enable_hme_level0_flag = 1;
    p->enable_hme_level1_flag = lvl <= 2;
    p->enable_hme_level2_flag = lvl == 0;
    p->hme_search_method = p->me_search_method = lvl <= 2 ? ME_GPU_FULL_SAD_SEARCH : ME_GPU_SUB_SAD_SEARCH;
    set_prehme(p, 0);
    set_ref_prune(p, 0);
    set_sr_adjustment(p, 0);
    p->me_early_exit_th            = lvl <= 1 ? 0 : 64 * 64 * 4;
    p->prev_me_stage_based_exit_th = lvl <= 1 ? 0 : 64 * 64 * 4;
    return ME_GPU_OK;
}

/* exported for tests: the mrp inputs of svt_aom_sig_deriv_me this restatement used */
void me_gpu_derive_mrp_inputs(const MeGpuEncoderConfig *c, uint8_t *safe_limit_nref, uint32_t *safe_limit_zz_th) {
    const Mrp mrp     = mrp_ctrl(mrp_level(c, input_resolution((uint32_t)c->width * c->height)));
    *safe_limit_nref  = mrp.safe_limit_nref;
    *safe_limit_zz_th = mrp.safe_limit_zz_th;
}
