/*
 * svt_me_gpu_hook.c - see svt_me_gpu_hook.h. Plain C, no CUDA: the engine is a dlopen'ed C-ABI library.
 */
#include "svt_me_gpu_hook.h"

#include <dlfcn.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#include "reference_object.h"

int (*svt_me_picture_register)(uint64_t, const uint8_t *, int, int, int, int, int)                                   = NULL;
int (*svt_me_picture_submit)(uint64_t, const MeGpuParams *, const uint64_t[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS])        = NULL;
int (*svt_me_picture_wait)(uint64_t, MeGpuOutputs *)                                                                 = NULL;
int (*svt_me_picture_release)(uint64_t)                                                                              = NULL;
static int (*hook_cancel)(uint64_t) = NULL;
static int (*hook_read_plane)(uint64_t, int, uint8_t *, int) = NULL;
static int (*hook_wait_pair)(uint64_t, uint8_t, uint64_t, MeGpuOutputs *) = NULL;
static const char *(*hook_last_error)(void) = NULL;
static void (*hook_shutdown)(void) = NULL;
static int (*hook_dg)(uint64_t, uint64_t, uint8_t, MeGpuDgMetrics *) = NULL;
static void *hook_lib              = NULL;
static int   hook_log_lists        = 0;
static uint64_t hook_n_dg_read(void);
static void     hook_reset_tables(void);
static uint64_t hook_n_tf_read(void);
static int      hook_id_of(const void *obj, uint64_t *id);

#include <stdio.h>
static EbErrorType map_error(int rc) {
    if (rc != ME_GPU_OK) /* the engine's own words, once, where the failure happens: the caller only sees an EbErrorType */
        fprintf(stderr, "svt_me_gpu_hook: engine error %d: %s\n", rc, hook_last_error ? hook_last_error() : "?");
    return rc == ME_GPU_OK ? EB_ErrorNone : rc == ME_GPU_ERR_NO_MEMORY ? EB_ErrorInsufficientResources : EB_ErrorUndefined;
}

EbErrorType svt_me_gpu_hook_init(const char *library_path, int max_width, int max_height, int max_refs_l0, int max_refs_l1,
                                 int pictures_in_flight) {
    typedef int (*InitFn)(int, const int *, int, int, int, int, int);
    void *h = dlopen(library_path ? library_path : "libme_gpu.so", RTLD_NOW | RTLD_LOCAL);
    if (!h)
        return EB_ErrorUndefined;
    InitFn init = (InitFn)dlsym(h, "me_gpu_init");
    if (!init || init(0, NULL, max_width, max_height, max_refs_l0, max_refs_l1, pictures_in_flight) != ME_GPU_OK) {
        dlclose(h);
        return EB_ErrorInsufficientResources;
    }
    hook_log_lists          = getenv("SVT_ME_GPU_LOG_LISTS") != NULL; /* diagnostics only: changes nothing that is computed */
    svt_me_picture_register = (int (*)(uint64_t, const uint8_t *, int, int, int, int, int))dlsym(h, "me_gpu_register_picture");
    svt_me_picture_submit   = (int (*)(uint64_t, const MeGpuParams *, const uint64_t[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS]))dlsym(h, "me_gpu_submit");
    svt_me_picture_wait     = (int (*)(uint64_t, MeGpuOutputs *))dlsym(h, "me_gpu_wait");
    svt_me_picture_release  = (int (*)(uint64_t))dlsym(h, "me_gpu_release_picture");
    hook_cancel             = (int (*)(uint64_t))dlsym(h, "me_gpu_cancel");
    hook_read_plane         = (int (*)(uint64_t, int, uint8_t *, int))dlsym(h, "me_gpu_read_plane");
    hook_wait_pair          = (int (*)(uint64_t, uint8_t, uint64_t, MeGpuOutputs *))dlsym(h, "me_gpu_wait_pair");
    hook_last_error         = (const char *(*)(void))dlsym(h, "me_gpu_last_error");
    hook_shutdown           = (void (*)(void))dlsym(h, "me_gpu_shutdown");
    hook_dg                 = (int (*)(uint64_t, uint64_t, uint8_t, MeGpuDgMetrics *))dlsym(h, "me_gpu_dg_detector");
    hook_lib                = h;
    if (!svt_me_picture_register || !svt_me_picture_submit || !svt_me_picture_wait || !svt_me_picture_release || !hook_shutdown) {
        svt_me_gpu_hook_shutdown();
        return EB_ErrorUndefined;
    }
    return EB_ErrorNone;
}

void svt_me_gpu_hook_shutdown(void) {
    if (hook_shutdown) {
        uint64_t reg, rel, pics;
        svt_me_gpu_stats(&reg, &rel, &pics);
        fprintf(stderr, "svt_me_gpu_hook: %llu pictures searched, %llu pyramids built on the device (%llu for picture analysis / "
                        "the temporal filter / overlays), %llu released, %llu dynamic-GOP detector calls, %llu temporal-filter pairs\n",
                (unsigned long long)pics, (unsigned long long)reg, (unsigned long long)svt_me_gpu_downsample_count(),
                (unsigned long long)rel, (unsigned long long)hook_n_dg_read(), (unsigned long long)hook_n_tf_read());
        hook_shutdown();
        hook_reset_tables(); /* a later svt_me_gpu_hook_init (another encoder instance of the process) starts clean */
    }
    svt_me_picture_register = NULL;
    svt_me_picture_submit   = NULL;
    svt_me_picture_wait     = NULL;
    svt_me_picture_release  = NULL;
    hook_shutdown           = NULL;
    hook_cancel             = NULL;
    hook_read_plane         = NULL;
    hook_wait_pair          = NULL;
    hook_dg                 = NULL;
    if (hook_lib)
        dlclose(hook_lib);
    hook_lib = NULL;
}

void svt_me_gpu_params_from_me_ctx(MeGpuParams *p, const MeContext *m) {
    p->enable_hme_flag            = m->enable_hme_flag;
    p->enable_hme_level0_flag     = m->enable_hme_level0_flag;
    p->enable_hme_level1_flag     = m->enable_hme_level1_flag;
    p->enable_hme_level2_flag     = m->enable_hme_level2_flag;
    p->hme_search_method          = m->hme_search_method;
    p->me_search_method           = m->me_search_method;
    p->use_best_unipred_cand_only = m->use_best_unipred_cand_only;
    p->reduce_hme_l0_sr_th_min    = m->reduce_hme_l0_sr_th_min;
    p->reduce_hme_l0_sr_th_max    = m->reduce_hme_l0_sr_th_max;
    p->num_hme_sa_w               = m->num_hme_sa_w;
    p->num_hme_sa_h               = m->num_hme_sa_h;
    memcpy(&p->me_sa, &m->me_sa, sizeof p->me_sa); /* SearchAreaMinMax and MeGpuSearchAreaMinMax share their layout */
    memcpy(&p->hme_l0_sa, &m->hme_l0_sa, sizeof p->hme_l0_sa);
    memcpy(&p->hme_l1_sa, &m->hme_l1_sa, sizeof p->hme_l1_sa);
    memcpy(&p->hme_l2_sa, &m->hme_l2_sa, sizeof p->hme_l2_sa);
    p->enable_me_hme_ref_pruning               = m->me_hme_prune_ctrls.enable_me_hme_ref_pruning;
    p->prune_ref_if_hme_sad_dev_bigger_than_th = m->me_hme_prune_ctrls.prune_ref_if_hme_sad_dev_bigger_than_th;
    p->prune_ref_if_me_sad_dev_bigger_than_th  = m->me_hme_prune_ctrls.prune_ref_if_me_sad_dev_bigger_than_th;
    p->zz_sad_th                               = m->me_hme_prune_ctrls.zz_sad_th;
    p->zz_sad_pct                              = m->me_hme_prune_ctrls.zz_sad_pct;
    p->phme_sad_th                             = m->me_hme_prune_ctrls.phme_sad_th;
    p->phme_sad_pct                            = m->me_hme_prune_ctrls.phme_sad_pct;
    p->enable_me_sr_adjustment              = m->me_sr_adjustment_ctrls.enable_me_sr_adjustment;
    p->reduce_me_sr_based_on_mv_length_th   = m->me_sr_adjustment_ctrls.reduce_me_sr_based_on_mv_length_th;
    p->stationary_hme_sad_abs_th            = m->me_sr_adjustment_ctrls.stationary_hme_sad_abs_th;
    p->stationary_me_sr_divisor             = m->me_sr_adjustment_ctrls.stationary_me_sr_divisor;
    p->reduce_me_sr_based_on_hme_sad_abs_th = m->me_sr_adjustment_ctrls.reduce_me_sr_based_on_hme_sad_abs_th;
    p->me_sr_divisor_for_low_hme_sad        = m->me_sr_adjustment_ctrls.me_sr_divisor_for_low_hme_sad;
    p->distance_based_hme_resizing          = m->me_sr_adjustment_ctrls.distance_based_hme_resizing;
    p->me_8x8_var_enabled = m->me_8x8_var_ctrls.enabled;
    p->me_sr_div4_th      = m->me_8x8_var_ctrls.me_sr_div4_th;
    p->me_sr_div2_th      = m->me_8x8_var_ctrls.me_sr_div2_th;
    p->me_sr_mult2_th     = m->me_8x8_var_ctrls.me_sr_mult2_th;
    p->mv_sa_adj_enabled          = m->mv_based_sa_adj.enabled;
    p->mv_sa_adj_nearest_ref_only = m->mv_based_sa_adj.nearest_ref_only;
    p->mv_sa_adj_mv_size_th       = m->mv_based_sa_adj.mv_size_th;
    p->mv_sa_adj_sa_multiplier    = m->mv_based_sa_adj.sa_multiplier;
    p->prehme_enable           = m->prehme_ctrl.enable;
    p->prehme_skip_search_line = m->prehme_ctrl.skip_search_line;
    p->prehme_l1_early_exit    = m->prehme_ctrl.l1_early_exit;
    memcpy(p->prehme_sa_cfg, m->prehme_ctrl.prehme_sa_cfg, sizeof p->prehme_sa_cfg);
    p->prune_me_candidates_th      = m->prune_me_candidates_th;
    p->me_early_exit_th            = m->me_early_exit_th;
    p->me_safe_limit_zz_th         = m->me_safe_limit_zz_th;
    p->prev_me_stage_based_exit_th = m->prev_me_stage_based_exit_th;
}

void svt_me_gpu_params_from_ctx(MeGpuParams *p, const PictureParentControlSet *pcs, const MeContext *m) {
    const SequenceControlSet *scs = pcs->scs;
    memset(p, 0, sizeof *p);
    svt_me_gpu_params_from_me_ctx(p, m);
    /* pcs / scs scalars read by svt_aom_motion_estimation_b64 and its callees */
    p->picture_number                  = pcs->picture_number;
    p->input_resolution                = (uint8_t)scs->input_resolution;
    p->enable_me_8x8                   = pcs->enable_me_8x8;
    p->enable_me_16x16                 = pcs->enable_me_16x16;
    p->max_number_of_pus_per_sb        = pcs->max_number_of_pus_per_sb;
    p->gm_enabled                      = pcs->gm_ctrls.enabled;
    p->gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
    p->hierarchical_levels             = pcs->hierarchical_levels;
    p->pcs_temporal_layer_index        = pcs->temporal_layer_index;
    p->similar_brightness_refs         = pcs->similar_brightness_refs;
    p->only_l_bwd                      = scs->mrp_ctrls.only_l_bwd;
    /* the temporal filter runs before the picture's pa_me_data is attached; its job writes none of the candidate arrays */
    p->max_cand                        = pcs->pa_me_data ? pcs->pa_me_data->max_cand : 1;
    p->max_refs                        = pcs->pa_me_data ? pcs->pa_me_data->max_refs : 1;
    p->max_l0                          = pcs->pa_me_data ? pcs->pa_me_data->max_l0 : 1;
    /* per-call MeContext fields, me_process.c:215-262 */
    p->me_type                     = (uint8_t)m->me_type;
    p->tf_me_exit_th               = m->tf_me_exit_th;
    p->num_of_list_to_search       = m->num_of_list_to_search;
    p->num_of_ref_pic_to_search[0] = m->num_of_ref_pic_to_search[0];
    p->num_of_ref_pic_to_search[1] = m->num_of_ref_pic_to_search[1];
    p->temporal_layer_index        = m->temporal_layer_index;
    p->is_ref                      = m->is_ref;
}

EbErrorType svt_me_gpu_register_picture(const PictureParentControlSet *pcs, const EbPictureBufferDesc *input_padded_pic) {
    if (!svt_me_picture_register)
        return EB_ErrorUndefined;
    const EbPictureBufferDesc *d = input_padded_pic; /* y8b luma with its 68-pixel border, reference_object.c:251-253 */
    return map_error(svt_me_picture_register(pcs->picture_number, d->buffer_y, d->stride_y, d->org_x, d->org_y,
                                             pcs->aligned_width, pcs->aligned_height));
}

EbErrorType svt_me_gpu_pyramid_for_analysis(const PictureParentControlSet *pcs, const EbPictureBufferDesc *input_padded_pic,
                                            EbPictureBufferDesc *quarter_pic, EbPictureBufferDesc *sixteenth_pic) {
    if (!hook_read_plane)
        return EB_ErrorUndefined;
    EbErrorType rc = svt_me_gpu_register_picture(pcs, input_padded_pic);
    if (rc != EB_ErrorNone)
        return rc;
    /* the descriptors' buffers start at the first padded byte; org_x = org_y = the plane's padding (32 / 16) */
    int e = hook_read_plane(pcs->picture_number, 1, quarter_pic->buffer_y, quarter_pic->stride_y);
    if (e == ME_GPU_OK)
        e = hook_read_plane(pcs->picture_number, 2, sixteenth_pic->buffer_y, sixteenth_pic->stride_y);
    return map_error(e);
}

EbErrorType svt_me_gpu_release_picture(uint64_t picture_number) {
    return svt_me_picture_release ? map_error(svt_me_picture_release(picture_number)) : EB_ErrorUndefined;
}

/* src_id / ref_ids: the numbers the pictures were registered under (their picture numbers, possibly tagged: see
 * svt_me_gpu_segment); NULL ref_ids = the plain picture numbers of me_ctx->me_ds_ref_array */
static EbErrorType hook_picture(PictureParentControlSet *pcs, const MeContext *me_ctx, uint64_t src_id,
                                const uint64_t ref_ids[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS]) {
    if (!svt_me_picture_submit || !svt_me_picture_wait)
        return EB_ErrorUndefined;
    MeGpuParams gp;
    uint64_t    refs[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS] = {{0}};
    svt_me_gpu_params_from_ctx(&gp, pcs, me_ctx);
    gp.picture_number = src_id; /* ME only reads differences of picture numbers, truncated to 16 bits (motion_estimation.c:1232) */
    for (int l = 0; l < me_ctx->num_of_list_to_search; l++)
        for (int r = 0; r < me_ctx->num_of_ref_pic_to_search[l]; r++)
            refs[l][r] = ref_ids ? ref_ids[l][r] : me_ctx->me_ds_ref_array[l][r].picture_number; /* me_process.c:248-262 */
    if (hook_log_lists) { /* SVT_ME_GPU_LOG_LISTS: the reference lists the encoder's own prediction structure gives (pd_process.c:3675-3693) */
        char line[256];
        int  n = snprintf(line, sizeof line, "svt_me_gpu_hook: lists poc %llu tl %u:", (unsigned long long)pcs->picture_number,
                          (unsigned)pcs->temporal_layer_index);
        for (int l = 0; l < me_ctx->num_of_list_to_search; l++) {
            n += snprintf(line + n, sizeof line - (size_t)n, " L%d", l);
            for (int r = 0; r < me_ctx->num_of_ref_pic_to_search[l]; r++)
                n += snprintf(line + n, sizeof line - (size_t)n, " %llu", (unsigned long long)me_ctx->me_ds_ref_array[l][r].picture_number);
        }
        fprintf(stderr, "%s\n", line);
    }
    const uint64_t job = src_id;
    const int rc = svt_me_picture_submit(job, &gp, refs);
    if (rc != ME_GPU_OK)
        return map_error(rc);

    /* MeGpuOutputs wants one slab per MeSbResults array; the reference allocates me_results[b64] one by one
       (pcs.c:1425-1444), so the three arrays come back into staging slabs and are copied out per b64. The eight
       per-b64 pcs arrays are contiguous already and are written in place. */
    const uint32_t b64_w = (uint32_t)(pcs->aligned_width + 63) / 64, b64_h = (uint32_t)(pcs->aligned_height + 63) / 64;
    const uint32_t nb = b64_w * b64_h;
    const uint32_t P  = pcs->enable_me_16x16 ? (pcs->enable_me_8x8 ? pcs->max_number_of_pus_per_sb : MAX_SB64_PU_COUNT_NO_8X8)
                                             : MAX_SB64_PU_COUNT_WO_16X16; /* pcs.c:108-117 */
    const size_t n_idx = (size_t)nb * P, n_cand = n_idx * gp.max_cand, n_mv = n_idx * gp.max_refs;
    uint8_t     *slab  = (uint8_t *)malloc(n_idx + n_cand + n_mv * sizeof(uint32_t) + 16);
    if (!slab) {
        if (hook_cancel)
            hook_cancel(job); /* retire the job: nothing is delivered */
        return EB_ErrorInsufficientResources;
    }
    MeGpuOutputs out;
    memset(&out, 0, sizeof out);
    out.b64_count                   = nb;
    out.me_mv_array                 = (uint32_t *)slab; /* first: 4-byte aligned */
    out.total_me_candidate_index    = slab + n_mv * sizeof(uint32_t);
    out.me_candidate_array          = out.total_me_candidate_index + n_idx;
    out.me_64x64_distortion         = pcs->me_64x64_distortion;
    out.me_32x32_distortion         = pcs->me_32x32_distortion;
    out.me_16x16_distortion         = pcs->me_16x16_distortion;
    out.me_8x8_distortion           = pcs->me_8x8_distortion;
    out.me_8x8_cost_variance        = pcs->me_8x8_cost_variance;
    out.rc_me_distortion            = pcs->rc_me_distortion;
    out.stationary_block_present_sb = pcs->stationary_block_present_sb;
    out.rc_me_allow_gm              = pcs->rc_me_allow_gm;
    const int rw = svt_me_picture_wait(job, &out);
    if (rw == ME_GPU_OK && out.pu_count == P)
        for (uint32_t b = 0; b < nb; b++) {
            MeSbResults *res = pcs->pa_me_data->me_results[b];
            memcpy(res->total_me_candidate_index, out.total_me_candidate_index + (size_t)b * P, P);
            memcpy(res->me_candidate_array, out.me_candidate_array + (size_t)b * P * gp.max_cand, (size_t)P * gp.max_cand);
            memcpy(res->me_mv_array, out.me_mv_array + (size_t)b * P * gp.max_refs, (size_t)P * gp.max_refs * sizeof(uint32_t));
        }
    free(slab);
    return rw != ME_GPU_OK ? map_error(rw) : out.pu_count == P ? EB_ErrorNone : EB_ErrorUndefined;
}

EbErrorType svt_me_gpu_picture(PictureParentControlSet *pcs, const MeContext *me_ctx) {
    return hook_picture(pcs, me_ctx, pcs->picture_number, NULL);
}

/* ---- per-picture once-logic and lazy registration ------------------------------------------------------------- */
#define HOOK_MAX_REGS 512 /* PA reference objects alive at once (pa_reference_picture_buffer_init_count is < 100) */
#define HOOK_MAX_PICS 128 /* pictures whose segment items are in flight */
typedef struct HookReg {
    const void *obj; /* EbPaReferenceObject the pyramid was uploaded from */
    uint64_t    number; /* its picture number at the time */
    uint64_t    id; /* what the engine knows it as: (registration count << 32) | low 32 bits of the picture number. An alt-ref
                       picture and its overlay share a picture number but not their pixels: identities must not collide, and
                       ME only ever reads 16-bit differences of the numbers */
    int         used;
    int         by_downsample; /* registered by svt_me_gpu_downsample: current by construction (every version of the object's
                                  picture passes through it), whatever obj->picture_number says - the overlay path never
                                  sets it (pd_process.c:2691-2740) */
} HookReg;
typedef struct HookPic {
    const void *pcs; /* the picture's control set: an alt-ref and its overlay are two pictures with one number */
    uint64_t    number;
    uint64_t    src_id, ref_ids[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS];
    int         used, done;
    uint32_t    arrivals;
    EbErrorType rc;
} HookPic;
static HookReg         hook_regs[HOOK_MAX_REGS];
static HookPic         hook_pics[HOOK_MAX_PICS];
static pthread_mutex_t hook_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t  hook_cv = PTHREAD_COND_INITIALIZER;
static uint64_t        hook_n_reg, hook_n_rel, hook_n_pic, hook_n_ds, hook_n_dg;

int svt_me_gpu_live(void) { return svt_me_picture_submit != NULL; }

void svt_me_gpu_stats(uint64_t *registered, uint64_t *released, uint64_t *pictures) {
    pthread_mutex_lock(&hook_mu);
    if (registered) *registered = hook_n_reg;
    if (released) *released = hook_n_rel;
    if (pictures) *pictures = hook_n_pic;
    pthread_mutex_unlock(&hook_mu);
}

/* hook_mu held. The device holds the pyramid of `obj`'s current picture after this; *id = its engine identity. */
static EbErrorType hook_ensure_registered(const PictureParentControlSet *pcs, const EbPaReferenceObject *obj, uint64_t *id) {
    HookReg *slot = NULL, *free_slot = NULL;
    for (int i = 0; i < HOOK_MAX_REGS; i++) {
        if (hook_regs[i].used && hook_regs[i].obj == (const void *)obj)
            slot = &hook_regs[i];
        else if (!hook_regs[i].used && !free_slot)
            free_slot = &hook_regs[i];
    }
    if (slot && (slot->by_downsample || slot->number == obj->picture_number)) {
        *id = slot->id;
        return EB_ErrorNone;
    }
    if (slot) { /* the pool handed the object to a new picture: the old one is done with (reference_object.c:406) */
        svt_me_picture_release(slot->id);
        hook_n_rel++;
        slot->used = 0;
        if (!free_slot)
            free_slot = slot;
    }
    if (!free_slot)
        return EB_ErrorInsufficientResources;
    const EbPictureBufferDesc *d  = (const EbPictureBufferDesc *)obj->input_padded_pic;
    const uint64_t             nid = ((hook_n_reg + 1) << 32) | (obj->picture_number & 0xFFFFFFFFull);
    const int rc = svt_me_picture_register(nid, d->buffer_y, d->stride_y, d->org_x, d->org_y, pcs->aligned_width, pcs->aligned_height);
    if (rc != ME_GPU_OK)
        return map_error(rc);
    free_slot->obj    = obj;
    free_slot->number = obj->picture_number;
    free_slot->id     = nid;
    free_slot->used   = 1;
    free_slot->by_downsample = 0;
    hook_n_reg++;
    *id = nid;
    return EB_ErrorNone;
}

/* Every caller of svt_aom_downsample_filtering_input_picture (picture analysis, pic_analysis_process.c:2345; the temporal
 * filter once it has rewritten the centre picture, temporal_filtering.c:3956; the overlay picture, pd_process.c:2735)
 * lands here when the hook is live: the device gets the luma plane as it is NOW, under a fresh identity (an earlier
 * version of the same object, or the picture the pool object held before, is released), builds the pyramid, and the two
 * CPU planes are filled from it. svt_me_gpu_segment finds the object registered and uploads nothing. */
EbErrorType svt_me_gpu_downsample(const PictureParentControlSet *pcs, const EbPictureBufferDesc *pic, EbPictureBufferDesc *quarter_pic,
                                  EbPictureBufferDesc *sixteenth_pic) {
    if (!svt_me_gpu_live() || !hook_read_plane)
        return EB_ErrorUndefined;
    const EbPaReferenceObject *obj  = (const EbPaReferenceObject *)pcs->pa_ref_pic_wrapper->object_ptr;
    HookReg                   *slot = NULL, *free_slot = NULL;
    pthread_mutex_lock(&hook_mu);
    for (int i = 0; i < HOOK_MAX_REGS; i++) {
        if (hook_regs[i].used && hook_regs[i].obj == (const void *)obj)
            slot = &hook_regs[i];
        else if (!hook_regs[i].used && !free_slot)
            free_slot = &hook_regs[i];
    }
    if (slot) {
        svt_me_picture_release(slot->id);
        hook_n_rel++;
        slot->used = 0;
    } else
        slot = free_slot;
    if (!slot) {
        pthread_mutex_unlock(&hook_mu);
        return EB_ErrorInsufficientResources;
    }
    const uint64_t nid = ((hook_n_reg + 1) << 32) | (pcs->picture_number & 0xFFFFFFFFull);
    int            e   = svt_me_picture_register(nid, pic->buffer_y, pic->stride_y, pic->org_x, pic->org_y, pcs->aligned_width,
                                                 pcs->aligned_height);
    if (e == ME_GPU_OK) {
        slot->obj    = obj;
        slot->number = pcs->picture_number;
        slot->id     = nid;
        slot->used   = 1;
        slot->by_downsample = 1;
        hook_n_reg++;
        hook_n_ds++;
    }
    pthread_mutex_unlock(&hook_mu);
    /* the read-backs wait for the upload and the pyramid kernel on the device; other pictures' calls are not held up */
    if (e == ME_GPU_OK)
        e = hook_read_plane(nid, 1, quarter_pic->buffer_y, quarter_pic->stride_y);
    if (e == ME_GPU_OK)
        e = hook_read_plane(nid, 2, sixteenth_pic->buffer_y, sixteenth_pic->stride_y);
    return map_error(e);
}

uint64_t svt_me_gpu_downsample_count(void) {
    pthread_mutex_lock(&hook_mu);
    const uint64_t n = hook_n_ds;
    pthread_mutex_unlock(&hook_mu);
    return n;
}

EbErrorType svt_me_gpu_segment(PictureParentControlSet *pcs, MeContext *me_ctx) {
    if (!svt_me_gpu_live())
        return EB_ErrorUndefined;
    HookPic *e = NULL;
    int      leader = 0;
    pthread_mutex_lock(&hook_mu);
    for (int i = 0; i < HOOK_MAX_PICS; i++)
        if (hook_pics[i].used && hook_pics[i].pcs == (const void *)pcs && hook_pics[i].number == pcs->picture_number)
            e = &hook_pics[i];
    if (!e) {
        for (int i = 0; i < HOOK_MAX_PICS && !e; i++)
            if (!hook_pics[i].used)
                e = &hook_pics[i];
        if (!e) {
            pthread_mutex_unlock(&hook_mu);
            return EB_ErrorInsufficientResources;
        }
        e->used = 1; e->done = 0; e->arrivals = 0; e->rc = EB_ErrorNone; e->number = pcs->picture_number; e->pcs = pcs;
        leader  = 1;
        hook_n_pic++;
        /* the pyramids this picture reads: its own and its references' (me_process.c:248-262) */
        EbErrorType rc = hook_ensure_registered(pcs, (const EbPaReferenceObject *)pcs->pa_ref_pic_wrapper->object_ptr, &e->src_id);
        for (int l = 0; rc == EB_ErrorNone && l < me_ctx->num_of_list_to_search; l++)
            for (int r = 0; rc == EB_ErrorNone && r < me_ctx->num_of_ref_pic_to_search[l]; r++)
                rc = hook_ensure_registered(pcs, (const EbPaReferenceObject *)pcs->ref_pa_pic_ptr_array[l][r]->object_ptr,
                                            &e->ref_ids[l][r]);
        e->rc = rc;
    }
    pthread_mutex_unlock(&hook_mu);
    if (leader) {
        EbErrorType rc = e->rc == EB_ErrorNone ? hook_picture(pcs, me_ctx, e->src_id, e->ref_ids) : e->rc;
        pthread_mutex_lock(&hook_mu);
        e->rc   = rc;
        e->done = 1;
        pthread_cond_broadcast(&hook_cv);
    } else {
        pthread_mutex_lock(&hook_mu);
        while (!e->done) pthread_cond_wait(&hook_cv, &hook_mu);
    }
    const EbErrorType rc = e->rc;
    if (++e->arrivals == pcs->me_segments_total_count)
        e->used = 0; /* TASK_SUPERRES_RE_ME of the same picture later starts a fresh entry */
    pthread_mutex_unlock(&hook_mu);
    return rc;
}

EbErrorType svt_me_gpu_tf_pair(const PictureParentControlSet *centre_pcs, MeContext *me_ctx, uint64_t neighbour_picture_number,
                               SvtMeGpuTfPair *pair) {
    if (!svt_me_picture_submit || !svt_me_picture_wait)
        return EB_ErrorUndefined;
    MeGpuParams gp;
    uint64_t    refs[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS] = {{0}};
    svt_me_gpu_params_from_ctx(&gp, centre_pcs, me_ctx);
    gp.me_type                     = ME_GPU_ME_MCTF;
    gp.num_of_list_to_search       = 1; /* temporal_filtering.c:3150-3152 */
    gp.num_of_ref_pic_to_search[0] = 1;
    gp.num_of_ref_pic_to_search[1] = 0;
    refs[0][0]                     = neighbour_picture_number;
    int rc = svt_me_picture_submit(centre_pcs->picture_number, &gp, refs);
    if (rc != ME_GPU_OK)
        return map_error(rc);
    MeGpuOutputs out;
    memset(&out, 0, sizeof out);
    out.b64_count     = pair->b64_count;
    out.tf_best_sad   = pair->best_sad;
    out.tf_best_mv    = pair->best_mv;
    out.tf_hme_centre = pair->hme_centre;
    out.tf_hme_sad    = pair->hme_sad;
    out.tf_early_exit = pair->early_exit;
    rc = svt_me_picture_wait(centre_pcs->picture_number, &out);
    if (rc != ME_GPU_OK)
        return map_error(rc);
    me_ctx->tf_tot_horz_blks += out.tf_tot_horz_blks; /* motion_estimation.c:2469-2474 */
    me_ctx->tf_tot_vert_blks += out.tf_tot_vert_blks;
    return EB_ErrorNone;
}

void svt_me_gpu_tf_load_block(MeContext *me_ctx, const SvtMeGpuTfPair *pair, uint32_t b) {
    memcpy(me_ctx->p_sb_best_sad[0][0], pair->best_sad + (size_t)SQUARE_PU_COUNT * b, SQUARE_PU_COUNT * sizeof(uint32_t));
    memcpy(me_ctx->p_sb_best_mv[0][0], pair->best_mv + (size_t)SQUARE_PU_COUNT * b, SQUARE_PU_COUNT * sizeof(uint32_t));
    /* the per-size views the integer search leaves pointing into slot (0, 0) (motion_estimation.c:1368-1384): a thread that never
       ran the CPU search has them unset */
    me_ctx->p_best_sad_64x64 = &me_ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_64x64];
    me_ctx->p_best_sad_32x32 = &me_ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_32x32_0];
    me_ctx->p_best_sad_16x16 = &me_ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_16x16_0];
    me_ctx->p_best_sad_8x8   = &me_ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_8x8_0];
    me_ctx->p_best_mv64x64   = &me_ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_64x64];
    me_ctx->p_best_mv32x32   = &me_ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_32x32_0];
    me_ctx->p_best_mv16x16   = &me_ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_16x16_0];
    me_ctx->p_best_mv8x8     = &me_ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_8x8_0];
    me_ctx->search_results[0][0].hme_sc_x = pair->hme_centre[2 * b];
    me_ctx->search_results[0][0].hme_sc_y = pair->hme_centre[2 * b + 1];
    me_ctx->search_results[0][0].hme_sad  = pair->hme_sad[b];
    if (pair->early_exit[b])
        me_ctx->tf_use_pred_64x64_only_th = (uint8_t)~0; /* motion_estimation.c:3111 */
}

/* ---- the temporal filter inside the encoder ---------------------------------------------------------------------- */
/* produce_temporally_filtered_pic walks blocks outside and neighbour pictures inside, on several segment threads per
 * centre picture. The first thread that needs the ME of a (centre, neighbour) pair runs it for ALL blocks on the device
 * (one ME_MCTF job) while the other threads that need the same pair wait; every later block of any segment loads its
 * results from the pair's arrays. The arrays live until every segment of the centre has called svt_me_gpu_tf_centre_done. */
#define HOOK_MAX_TF 64
typedef struct HookTf {
    const void    *centre;
    uint64_t       number;
    int            frame_index, state; /* 0 free, 1 running, 2 done */
    EbErrorType    rc;
    SvtMeGpuTfPair pair;
    uint32_t       horz, vert; /* the pair's tf_tot_horz/vert_blks, handed to the thread that ran it */
} HookTf;
typedef struct HookTfCentre {
    const void *centre;
    uint64_t    number;
    uint32_t    arrivals;
    int         used;
} HookTfCentre;
static HookTf       hook_tf[HOOK_MAX_TF];
static HookTfCentre hook_tfc[16];
static uint64_t     hook_n_tf;

int svt_me_gpu_tf_ready(const PictureParentControlSet *centre_pcs, const PictureParentControlSet *neighbour_pcs) {
    uint64_t  a, b;
    const int live = svt_me_gpu_live() && hook_wait_pair && centre_pcs->pa_ref_pic_wrapper && neighbour_pcs->pa_ref_pic_wrapper;
    const int ca = live && hook_id_of(centre_pcs->pa_ref_pic_wrapper->object_ptr, &a);
    const int nb = live && hook_id_of(neighbour_pcs->pa_ref_pic_wrapper->object_ptr, &b);
    if (hook_log_lists && !(ca && nb))
        fprintf(stderr, "svt_me_gpu_hook: tf pair (%llu, %llu) stays on the CPU: live %d wait_pair %d centre on device %d neighbour on device %d\n",
                (unsigned long long)centre_pcs->picture_number, (unsigned long long)neighbour_pcs->picture_number, svt_me_gpu_live(),
                hook_wait_pair != NULL, ca, nb);
    return ca && nb;
}

static EbErrorType hook_tf_run(const PictureParentControlSet *centre_pcs, MeContext *me_ctx, const PictureParentControlSet *neighbour_pcs,
                               HookTf *e, uint32_t b64_count) {
    uint64_t cid, nid;
    if (!hook_id_of(centre_pcs->pa_ref_pic_wrapper->object_ptr, &cid) || !hook_id_of(neighbour_pcs->pa_ref_pic_wrapper->object_ptr, &nid))
        return EB_ErrorUndefined;
    SvtMeGpuTfPair *pr = &e->pair;
    pr->b64_count  = b64_count;
    pr->best_sad   = (uint32_t *)malloc((size_t)b64_count * SQUARE_PU_COUNT * sizeof(uint32_t));
    pr->best_mv    = (uint32_t *)malloc((size_t)b64_count * SQUARE_PU_COUNT * sizeof(uint32_t));
    pr->hme_centre = (int16_t *)malloc((size_t)b64_count * 2 * sizeof(int16_t));
    pr->hme_sad    = (uint32_t *)malloc((size_t)b64_count * sizeof(uint32_t));
    pr->early_exit = (uint8_t *)malloc((size_t)b64_count);
    if (!pr->best_sad || !pr->best_mv || !pr->hme_centre || !pr->hme_sad || !pr->early_exit)
        return EB_ErrorInsufficientResources;
    MeGpuParams gp;
    uint64_t    refs[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS] = {{0}};
    svt_me_gpu_params_from_ctx(&gp, centre_pcs, me_ctx);
    gp.picture_number              = cid; /* distances are not used by ME_MCTF beyond the 16-bit difference of the low bits */
    gp.me_type                     = ME_GPU_ME_MCTF;
    gp.num_of_list_to_search       = 1;
    gp.num_of_ref_pic_to_search[0] = 1;
    gp.num_of_ref_pic_to_search[1] = 0;
    refs[0][0]                     = nid;
    int rc = svt_me_picture_submit(cid, &gp, refs);
    if (rc != ME_GPU_OK)
        return map_error(rc);
    MeGpuOutputs out;
    memset(&out, 0, sizeof out);
    out.b64_count     = b64_count;
    out.tf_best_sad   = pr->best_sad;
    out.tf_best_mv    = pr->best_mv;
    out.tf_hme_centre = pr->hme_centre;
    out.tf_hme_sad    = pr->hme_sad;
    out.tf_early_exit = pr->early_exit;
    rc = hook_wait_pair(cid, ME_GPU_ME_MCTF, nid, &out);
    if (rc != ME_GPU_OK)
        return map_error(rc);
    e->horz = out.tf_tot_horz_blks;
    e->vert = out.tf_tot_vert_blks;
    return EB_ErrorNone;
}

EbErrorType svt_me_gpu_tf_block(const PictureParentControlSet *centre_pcs, MeContext *me_ctx, const PictureParentControlSet *neighbour_pcs,
                                int frame_index, uint32_t b64_count, uint32_t b) {
    HookTf *e = NULL, *free_e = NULL;
    int     runner = 0;
    pthread_mutex_lock(&hook_mu);
    for (int i = 0; i < HOOK_MAX_TF; i++) {
        if (hook_tf[i].state && hook_tf[i].centre == (const void *)centre_pcs && hook_tf[i].number == centre_pcs->picture_number &&
            hook_tf[i].frame_index == frame_index)
            e = &hook_tf[i];
        else if (!hook_tf[i].state && !free_e)
            free_e = &hook_tf[i];
    }
    if (!e) {
        if (!free_e) {
            pthread_mutex_unlock(&hook_mu);
            return EB_ErrorInsufficientResources;
        }
        e = free_e;
        memset(e, 0, sizeof *e);
        e->centre = centre_pcs; e->number = centre_pcs->picture_number; e->frame_index = frame_index; e->state = 1;
        runner = 1;
        hook_n_tf++;
    }
    pthread_mutex_unlock(&hook_mu);
    if (runner) {
        const EbErrorType rc = hook_tf_run(centre_pcs, me_ctx, neighbour_pcs, e, b64_count);
        if (rc == EB_ErrorNone) { /* motion_estimation.c:2469-2474, summed over the segment threads at temporal_filtering.c:4259 */
            me_ctx->tf_tot_horz_blks += e->horz;
            me_ctx->tf_tot_vert_blks += e->vert;
        }
        pthread_mutex_lock(&hook_mu);
        e->rc    = rc;
        e->state = 2;
        pthread_cond_broadcast(&hook_cv);
    } else {
        pthread_mutex_lock(&hook_mu);
        while (e->state != 2) pthread_cond_wait(&hook_cv, &hook_mu);
    }
    const EbErrorType rc = e->rc;
    pthread_mutex_unlock(&hook_mu);
    if (rc == EB_ErrorNone)
        svt_me_gpu_tf_load_block(me_ctx, &e->pair, b);
    return rc;
}

void svt_me_gpu_tf_centre_done(const PictureParentControlSet *centre_pcs) {
    if (!svt_me_gpu_live())
        return;
    pthread_mutex_lock(&hook_mu);
    HookTfCentre *c = NULL, *free_c = NULL;
    for (int i = 0; i < 16; i++) {
        if (hook_tfc[i].used && hook_tfc[i].centre == (const void *)centre_pcs && hook_tfc[i].number == centre_pcs->picture_number)
            c = &hook_tfc[i];
        else if (!hook_tfc[i].used && !free_c)
            free_c = &hook_tfc[i];
    }
    if (!c && free_c) {
        c = free_c;
        c->centre = centre_pcs; c->number = centre_pcs->picture_number; c->arrivals = 0; c->used = 1;
    }
    if (c && ++c->arrivals == (uint32_t)centre_pcs->tf_segments_total_count) {
        c->used = 0;
        for (int i = 0; i < HOOK_MAX_TF; i++)
            if (hook_tf[i].state && hook_tf[i].centre == (const void *)centre_pcs && hook_tf[i].number == centre_pcs->picture_number) {
                free(hook_tf[i].pair.best_sad); free(hook_tf[i].pair.best_mv); free(hook_tf[i].pair.hme_centre);
                free(hook_tf[i].pair.hme_sad); free(hook_tf[i].pair.early_exit);
                hook_tf[i].state = 0;
            }
    }
    pthread_mutex_unlock(&hook_mu);
}

static void hook_reset_tables(void) {
    pthread_mutex_lock(&hook_mu);
    for (int i = 0; i < HOOK_MAX_TF; i++)
        if (hook_tf[i].state) {
            free(hook_tf[i].pair.best_sad); free(hook_tf[i].pair.best_mv); free(hook_tf[i].pair.hme_centre);
            free(hook_tf[i].pair.hme_sad); free(hook_tf[i].pair.early_exit);
        }
    memset(hook_tf, 0, sizeof hook_tf);
    memset(hook_tfc, 0, sizeof hook_tfc);
    memset(hook_regs, 0, sizeof hook_regs);
    memset(hook_pics, 0, sizeof hook_pics);
    hook_n_reg = hook_n_rel = hook_n_pic = hook_n_ds = hook_n_dg = hook_n_tf = 0;
    pthread_mutex_unlock(&hook_mu);
}

static uint64_t hook_n_tf_read(void) {
    pthread_mutex_lock(&hook_mu);
    const uint64_t n = hook_n_tf;
    pthread_mutex_unlock(&hook_mu);
    return n;
}

static uint64_t hook_n_dg_read(void) {
    pthread_mutex_lock(&hook_mu);
    const uint64_t n = hook_n_dg;
    pthread_mutex_unlock(&hook_mu);
    return n;
}

/* hook_mu NOT held. The engine identity of the pyramid registered for `obj` (svt_me_gpu_downsample / svt_me_gpu_segment), or 0 */
static int hook_id_of(const void *obj, uint64_t *id) {
    int found = 0;
    pthread_mutex_lock(&hook_mu);
    for (int i = 0; i < HOOK_MAX_REGS && !found; i++)
        if (hook_regs[i].used && hook_regs[i].obj == obj) {
            *id   = hook_regs[i].id;
            found = 1;
        }
    pthread_mutex_unlock(&hook_mu);
    return found;
}

int svt_me_gpu_dg_ready(const PictureParentControlSet *src_pcs, const PictureParentControlSet *ref_pcs) {
    uint64_t a, b;
    return hook_dg && src_pcs->pa_ref_pic_wrapper && ref_pcs->pa_ref_pic_wrapper && hook_id_of(src_pcs->pa_ref_pic_wrapper->object_ptr, &a) &&
           hook_id_of(ref_pcs->pa_ref_pic_wrapper->object_ptr, &b);
}

EbErrorType svt_me_gpu_dg_detector(PictureParentControlSet *src_pcs) {
    if (!hook_dg)
        return EB_ErrorUndefined;
    MeGpuDgMetrics                 m;
    const PictureParentControlSet *ref_pcs = src_pcs->dg_detector->ref_pic;
    /* inside the encoder the pictures are known to the engine by the identities the hook registered them under (keyed by
       their PA reference objects); a host that registered them itself (svt_me_gpu_register_picture) uses the picture numbers */
    uint64_t a = src_pcs->picture_number, b = ref_pcs->picture_number;
    if (src_pcs->pa_ref_pic_wrapper && ref_pcs->pa_ref_pic_wrapper) {
        uint64_t ia, ib;
        if (hook_id_of(src_pcs->pa_ref_pic_wrapper->object_ptr, &ia) && hook_id_of(ref_pcs->pa_ref_pic_wrapper->object_ptr, &ib)) {
            a = ia;
            b = ib;
        }
    }
    const int rc = hook_dg(a, b, (uint8_t)src_pcs->input_resolution, &m);
    if (rc != ME_GPU_OK)
        return map_error(rc);
    src_pcs->dg_detector->metrics.tot_dist       = m.tot_dist; /* then pd_process.c:595-603 unchanged */
    src_pcs->dg_detector->metrics.tot_cplx       = m.tot_cplx;
    src_pcs->dg_detector->metrics.tot_active     = m.tot_active;
    src_pcs->dg_detector->metrics.sum_in_vectors = m.sum_in_vectors;
    pthread_mutex_lock(&hook_mu);
    hook_n_dg++;
    pthread_mutex_unlock(&hook_mu);
    return EB_ErrorNone;
}
