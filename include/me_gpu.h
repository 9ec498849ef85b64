/*
 * me_gpu.h - C-ABI of the B200 open-loop motion-estimation engine.
 *
 * Drop-in boundary for SVT-AV1-PSYEX's per-picture ME path.
 * Plain C, host pointers and sizes only; no CUDA or torch types.
 *
 * What each entry point replaces in the reference (paths under Source/Lib/Codec):
 *   me_gpu_register_picture  <- svt_aom_downsample_filtering_input_picture
 *                               (pic_analysis_process.c:2139) + svt_aom_generate_padding
 *                               (pic_operators.c:397) + the PA reference object planes
 *                               (reference_object.c:249-295)
 *   me_gpu_submit / _wait    <- the b64 loop of svt_aom_motion_estimation_kernel
 *                               (me_process.c:174-290) calling
 *                               svt_aom_motion_estimation_b64 (motion_estimation.c:3076)
 *   me_gpu_release_picture   <- svt_aom_release_pa_reference_objects (reference_object.c:406)
 *   MeGpuParams              <- the MeContext control fields written by
 *                               svt_aom_sig_deriv_me (enc_mode_config.c:681-828) and
 *                               me_process.c:195-262, plus the pcs/scs scalars that
 *                               motion_estimation.c reads
 *   MeGpuOutputs             <- pcs->pa_me_data->me_results[b64] (me_sb_results.h:28-51) and the
 *                               per-b64 pcs arrays written by compute_distortion /
 *                               perform_gm_detection (motion_estimation.c:2964,2838)
 *   me_gpu_submit with me_type = ME_GPU_ME_MCTF
 *                            <- the svt_aom_motion_estimation_b64 call of the temporal filter, one (centre,
 *                               neighbour) pair per job (temporal_filtering.c:3144-3187); results in the tf_*
 *                               arrays of MeGpuOutputs
 *   me_gpu_dg_detector       <- the TASK_DG_DETECTOR_HME segments of one (source, reference) pair:
 *                               early_hme's dispatch + semaphore wait (pd_process.c:561-625) and
 *                               dg_detector_hme_level0 / early_hme_b64 (pd_process.c:393-560)
 *   MeGpuDgMetrics           <- DGDetectorMetrics (pcs.h:731-737)
 *
 * There is no CPU fallback behind this interface: every call fails with
 * ME_GPU_ERR_NO_DEVICE when no CUDA device is usable.
 */
#ifndef ME_GPU_H
#define ME_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ME_GPU_MAX_LISTS 2 /* MAX_NUM_OF_REF_PIC_LIST, definitions.h:2337 */
#define ME_GPU_MAX_REFS 4 /* MAX_REF_IDX / REF_LIST_MAX_DEPTH, definitions.h:2352 */
#define ME_GPU_SQUARE_PU_COUNT 85 /* SQUARE_PU_COUNT, me_sb_results.h:23 */
#define ME_GPU_SEARCH_REGIONS 2 /* SEARCH_REGION_COUNT, me_context.h:320 */

#define ME_GPU_SUB_SAD_SEARCH 0 /* definitions.h:2071 */
#define ME_GPU_FULL_SAD_SEARCH 1

/* MeGpuParams.me_type (EbMeType, me_context.h:44-50): the two callers of svt_aom_motion_estimation_b64 on this path */
#define ME_GPU_ME_MCTF 1 /* temporal-filter ME: one reference, HME + integer search only, early exit on a low HME SAD */
#define ME_GPU_ME_OPEN_LOOP 3 /* TASK_PAME */

/* error codes (0 = ok); the host maps non-zero to EB_ErrorInsufficientResources / EB_ErrorUndefined */
#define ME_GPU_OK 0
#define ME_GPU_ERR_NO_DEVICE -1
#define ME_GPU_ERR_BAD_ARG -2
#define ME_GPU_ERR_NOT_FOUND -3
#define ME_GPU_ERR_NO_MEMORY -4
#define ME_GPU_ERR_UNSUPPORTED -5
#define ME_GPU_ERR_CUDA -6
#define ME_GPU_ERR_STATE -7

typedef struct MeGpuSearchArea { /* SearchArea, me_context.h:321 */
    uint16_t width;
    uint16_t height;
} MeGpuSearchArea;

typedef struct MeGpuSearchAreaMinMax { /* SearchAreaMinMax, me_context.h:325 */
    MeGpuSearchArea sa_min;
    MeGpuSearchArea sa_max;
} MeGpuSearchAreaMinMax;

/* Flat copy of every input svt_aom_motion_estimation_b64 reads besides pixels. */
typedef struct MeGpuParams {
    /* ---- pcs / scs scalars ---- */
    uint64_t picture_number; /* pcs->picture_number */
    uint8_t  input_resolution; /* scs->input_resolution (EbInputResolution, definitions.h:2079) */
    uint8_t  enable_me_8x8; /* pcs->enable_me_8x8 */
    uint8_t  enable_me_16x16; /* pcs->enable_me_16x16 */
    uint8_t  max_number_of_pus_per_sb; /* pcs->max_number_of_pus_per_sb (85) */
    uint8_t  gm_enabled; /* pcs->gm_ctrls.enabled */
    uint8_t  gm_use_distance_based_active_th; /* pcs->gm_ctrls.use_distance_based_active_th */
    uint8_t  hierarchical_levels; /* pcs->hierarchical_levels */
    uint8_t  pcs_temporal_layer_index; /* pcs->temporal_layer_index */
    uint8_t  similar_brightness_refs; /* pcs->similar_brightness_refs */
    uint8_t  only_l_bwd; /* scs->mrp_ctrls.only_l_bwd */
    uint8_t  max_cand; /* pcs->pa_me_data->max_cand */
    uint8_t  max_refs; /* pcs->pa_me_data->max_refs */
    uint8_t  max_l0; /* pcs->pa_me_data->max_l0 */
    /* ---- per-item MeContext fields, me_process.c:219-262 ---- */
    uint8_t num_of_list_to_search;
    uint8_t num_of_ref_pic_to_search[ME_GPU_MAX_LISTS];
    uint8_t temporal_layer_index; /* me_ctx->temporal_layer_index */
    uint8_t is_ref;
    /* ---- controls written by svt_aom_sig_deriv_me ---- */
    uint8_t  enable_hme_flag;
    uint8_t  enable_hme_level0_flag;
    uint8_t  enable_hme_level1_flag;
    uint8_t  enable_hme_level2_flag;
    uint8_t  hme_search_method; /* ME_GPU_SUB_SAD_SEARCH for TASK_PAME */
    uint8_t  me_search_method;
    uint8_t  use_best_unipred_cand_only;
    uint8_t  reduce_hme_l0_sr_th_min;
    uint8_t  reduce_hme_l0_sr_th_max;
    uint16_t num_hme_sa_w; /* always 2 (get_worst_quadrant asserts it) */
    uint16_t num_hme_sa_h;
    MeGpuSearchAreaMinMax me_sa;
    MeGpuSearchAreaMinMax hme_l0_sa;
    MeGpuSearchArea       hme_l1_sa;
    MeGpuSearchArea       hme_l2_sa;
    /* MeHmeRefPruneCtrls, me_context.h:280 */
    uint8_t  enable_me_hme_ref_pruning;
    uint16_t prune_ref_if_hme_sad_dev_bigger_than_th;
    uint16_t prune_ref_if_me_sad_dev_bigger_than_th;
    uint32_t zz_sad_th;
    uint16_t zz_sad_pct;
    uint32_t phme_sad_th;
    uint16_t phme_sad_pct;
    /* MeSrCtrls, me_context.h:292 */
    uint8_t  enable_me_sr_adjustment;
    uint16_t reduce_me_sr_based_on_mv_length_th;
    uint16_t stationary_hme_sad_abs_th;
    uint16_t stationary_me_sr_divisor;
    uint16_t reduce_me_sr_based_on_hme_sad_abs_th;
    uint16_t me_sr_divisor_for_low_hme_sad;
    uint8_t  distance_based_hme_resizing;
    /* Me8x8VarCtrls, me_context.h:310 */
    uint8_t  me_8x8_var_enabled;
    uint32_t me_sr_div4_th;
    uint32_t me_sr_div2_th;
    uint32_t me_sr_mult2_th;
    /* MvBasedSearchAdj, me_context.h:356 */
    uint8_t  mv_sa_adj_enabled;
    uint8_t  mv_sa_adj_nearest_ref_only;
    uint16_t mv_sa_adj_mv_size_th;
    uint16_t mv_sa_adj_sa_multiplier;
    /* PreHmeCtrls, me_context.h:336 */
    uint8_t               prehme_enable;
    uint8_t               prehme_skip_search_line;
    uint8_t               prehme_l1_early_exit;
    MeGpuSearchAreaMinMax prehme_sa_cfg[ME_GPU_SEARCH_REGIONS];
    /* remaining scalars */
    int32_t  prune_me_candidates_th;
    uint32_t me_early_exit_th;
    uint32_t me_safe_limit_zz_th;
    uint32_t prev_me_stage_based_exit_th;
    /* ---- caller of svt_aom_motion_estimation_b64 ---- */
    uint8_t  me_type; /* ME_GPU_ME_OPEN_LOOP (me_process.c:215) or ME_GPU_ME_MCTF (temporal_filtering.c:3182) */
    uint16_t tf_me_exit_th; /* me_ctx->tf_me_exit_th = tf_ctrls.me_exit_th (temporal_filtering.c:3171); ME_MCTF only */
} MeGpuParams;

/* What the encoder knows before svt_aom_sig_deriv_me runs; input of me_gpu_derive_params(). */
typedef struct MeGpuEncoderConfig {
    int8_t   enc_mode; /* ENC_MRS=-3, ENC_MRP=-2, ENC_MR=-1, ENC_M0=0 ... ENC_M13 (definitions.h EncMode) */
    uint8_t  rtc_tune; /* pred_structure == SVT_AV1_PRED_LOW_DELAY_B */
    uint8_t  sc_class1; /* screen-content class */
    uint8_t  qp; /* static_config.qp (default 35) */
    uint8_t  rate_control_vbr; /* static_config.rate_control_mode == VBR */
    uint8_t  fast_decode; /* static_config.fast_decode */
    uint8_t  hierarchical_levels; /* 4 for a 16-picture mini-GOP */
    uint8_t  temporal_layer_index;
    uint8_t  is_ref;
    uint8_t  slice_type; /* 0 = B_SLICE, 1 = P_SLICE (definitions.h SliceType); I pictures are skipped */
    uint8_t  similar_brightness_refs;
    uint8_t  low_frame_rate; /* scs->frame_rate >> 16 != 0 */
    int8_t   force_ref_l0; /* >0: override the mrp-level reference counts (BASELINE config labels) */
    int8_t   force_ref_l1;
    uint16_t width; /* luma size, multiple of 8 */
    uint16_t height;
    uint64_t picture_number;
} MeGpuEncoderConfig;

/* Per-picture results in caller-owned host memory. Array extents:
 *   b64_count = ceil(w/64)*ceil(h/64); P = number of exported PUs (85 / 21 / 5, pcs.c:108-117). */
typedef struct MeGpuOutputs {
    uint32_t b64_count; /* in: capacity in b64s; out: b64s written */
    uint32_t pu_count; /* out: P */
    /* MeSbResults images, one contiguous slab per array (b64-major) */
    uint8_t  *total_me_candidate_index; /* [b64][P] */
    uint8_t  *me_candidate_array; /* [b64][P*max_cand], MeCandidate bitfield bytes */
    uint32_t *me_mv_array; /* [b64][P*max_refs], MvCandidate.as_int */
    /* per-b64 pcs arrays */
    uint32_t *me_64x64_distortion;
    uint32_t *me_32x32_distortion;
    uint32_t *me_16x16_distortion;
    uint32_t *me_8x8_distortion;
    uint32_t *me_8x8_cost_variance;
    uint32_t *rc_me_distortion;
    uint8_t  *stationary_block_present_sb;
    uint8_t  *rc_me_allow_gm;
    /* optional parity/debug taps (may be NULL): MeContext state after the b64 finished */
    uint32_t *p_sb_best_sad; /* [b64][2][4][85] */
    uint32_t *p_sb_best_mv; /* [b64][2][4][85] */
    int16_t  *hme_search_centre; /* [b64][2][4][2] (hme_sc_x, hme_sc_y) */
    uint64_t *hme_sad; /* [b64][2][4] search_results.hme_sad at exit */
    uint8_t  *do_ref; /* [b64][2][4] */
    int16_t  *search_area; /* [b64][2][4][4] final (origin_x, origin_y, width, height) of the integer search */
    /* ME_GPU_ME_MCTF results (the only product outputs of that mode; the arrays above are not written except the
     * taps): what the temporal filter reads from the MeContext after the call (temporal_filtering.c:1867-1870,
     * 2640-2690, 3190-3192) */
    uint32_t *tf_best_sad; /* [b64][85] p_sb_best_sad[0][0]; MAX_SAD_VALUE where the integer search did not run */
    uint32_t *tf_best_mv; /* [b64][85] p_sb_best_mv[0][0] */
    int16_t  *tf_hme_centre; /* [b64][2] search_results[0][0].hme_sc_x, hme_sc_y */
    uint32_t *tf_hme_sad; /* [b64] search_results[0][0].hme_sad after hme_b64 */
    uint8_t  *tf_early_exit; /* [b64] 1: hme_sad < tf_me_exit_th, tf_use_pred_64x64_only_th was forced to 0xFF
                                (motion_estimation.c:3109-3113) and the integer search skipped */
    uint32_t  tf_tot_horz_blks; /* out: b64s with |hme_sc_x| > |hme_sc_y| (motion_estimation.c:2469-2474) */
    uint32_t  tf_tot_vert_blks; /* out: the others */
} MeGpuOutputs;

/* library / device lifecycle -------------------------------------------------------------
 * pictures_in_flight = capacity of the device picture pool (registered and not yet released pictures,
 * sources and references alike); up to min(pictures_in_flight, 16) pictures may be submitted and not yet
 * waited for per device. device_count <= 0 selects the current CUDA device. */
int me_gpu_init(int device_count, const int *device_ids, int max_width, int max_height, int max_refs_l0,
                int max_refs_l1, int pictures_in_flight);
void me_gpu_shutdown(void);
const char *me_gpu_last_error(void);
int me_gpu_device_count(void);

/* pictures ------------------------------------------------------------------------------- */
/* y8b_padded points at the first byte of the padded luma plane (buffer_y), stride bytes per row,
 * picture origin at (org_x, org_y); the caller has already edge-replicated the padding
 * (svt_aom_pad_input_pictures). The 1/4 and 1/16 planes are built and padded on the device. */
int me_gpu_register_picture(uint64_t picture_number, const uint8_t *y8b_padded, int stride, int org_x, int org_y,
                            int width, int height);
/* same, source plane without padding (stride >= width): padding of 68 is generated on the device */
int me_gpu_register_picture_unpadded(uint64_t picture_number, const uint8_t *y8b, int stride, int width,
                                     int height);
int  me_gpu_release_picture(uint64_t picture_number);
/* copy the device pyramids back (parity tests): planes are full padded images; any pointer may be NULL */
int me_gpu_read_pyramid(uint64_t picture_number, uint8_t *full, uint8_t *quarter, uint8_t *sixteenth);
/* one level into a caller-owned padded plane with the caller's stride (>= plane width + 2 * pad): what picture analysis
 * needs to take the device-built 1/4 and 1/16 planes instead of running svt_aom_downsample_filtering_input_picture
 * (pic_analysis_process.c:2139-2196; its histograms read the 1/16 plane, :1560-1570) - SURVEY 8(f)-4 */
int me_gpu_read_plane(uint64_t picture_number, int level, uint8_t *dst, int dst_stride);
/* plane geometry helper: level 0 = full (pad 68), 1 = quarter (pad 32), 2 = sixteenth (pad 16) */
int me_gpu_plane_geometry(int width, int height, int level, int *plane_w, int *plane_h, int *pad, int *stride,
                          int *rows);

/* one process per GPU ---------------------------------------------------------
 * Every process runs its own engine on its own GPU and owns the pictures it registered. A picture another process needs
 * as a reference is pulled from the owner's pool over NVLink: device-to-device through a CUDA IPC mapping of that pool,
 * ordered by the owner's (interprocess) pyramid-complete event. No collective, no host copy. The encoder's equivalent is
 * nothing (one process, shared memory): pictures are simply visible to all ME threads (me_process.c:248-262).
 *   owner:    me_gpu_ipc_export_pool once; per picture me_gpu_ipc_picture AFTER me_gpu_register_picture returned
 *   consumer: me_gpu_ipc_open_peer once per owner; per picture me_gpu_register_picture_from_peer with what the owner's
 *             me_gpu_ipc_picture returned (carried between the processes by any means: 4 + 64 bytes)
 * The owner must keep the picture registered until the consumer's copy has run (the consumer's me_gpu_wait of any job that
 * reads the picture, or its me_gpu_release_picture, implies it). Both pools must be created with the same maximum size. */
int me_gpu_ipc_export_pool(unsigned char handle[64], uint64_t *slot_bytes);
int me_gpu_ipc_open_peer(int peer, const unsigned char handle[64], uint64_t slot_bytes);
int me_gpu_ipc_picture(uint64_t picture_number, int *slot, unsigned char event_handle[64]);
int me_gpu_register_picture_from_peer(uint64_t picture_number, int peer, int peer_slot, const unsigned char event_handle[64],
                                      int width, int height);
uint64_t me_gpu_ipc_bytes_pulled(void); /* bytes copied from peers' pools so far */

/* more than one engine per process ------------------------------------------------------------
 * The reference library serves several encoder instances in one process (EbComponentType handles, enc_handle.c:1405);
 * so does this one. Every me_gpu_* call goes to the engine BOUND TO THE CALLING THREAD: the process-wide default engine
 * unless the thread has called me_gpu_engine_bind. An instance creates its engine once, and each of its threads binds it
 * before its first call (a thread-local pointer: free). Engines share nothing (pool, jobs, streams, errors are per engine
 * / per thread); the static scratch of me_gpu_dg_detector_noalloc is the one thing every caller of a process shares. */
typedef struct MeGpuEngine MeGpuEngine;
int  me_gpu_engine_create(MeGpuEngine **engine); /* not initialised yet: bind it, then me_gpu_init */
void me_gpu_engine_bind(MeGpuEngine *engine); /* NULL = back to the default engine */
int  me_gpu_engine_destroy(MeGpuEngine *engine); /* after me_gpu_shutdown on it (ME_GPU_ERR_STATE otherwise); the caller's binding
                                                    goes back to the default engine, no OTHER thread may still be bound to it */

/* motion estimation ----------------------------------------------------------------------- */
int me_gpu_submit(uint64_t picture_number, const MeGpuParams *params,
                  const uint64_t ref_picture_numbers[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS]);
/* Blocks until the oldest job in flight of the picture is complete, copies its results into *out and retires the job. ONE
 * thread retires a job: a second me_gpu_wait on it while the first is blocked returns ME_GPU_ERR_STATE (threads that only
 * need to know that the results have been delivered call me_gpu_wait_done). After a successful submit EVERY outcome of the
 * wait retires the job - device errors included - except ME_GPU_ERR_BAD_ARG for a too small *out, which leaves it in
 * flight (come back with a larger buffer, or me_gpu_cancel). */
int me_gpu_wait(uint64_t picture_number, MeGpuOutputs *out);
/* same, for exactly the job submitted with this me_type and first reference (ref_picture_numbers[0][0]): the open-loop ME
 * of a picture and its temporal-filter pairs may be in flight together */
int me_gpu_wait_pair(uint64_t picture_number, uint8_t me_type, uint64_t first_ref_picture_number, MeGpuOutputs *out);
/* waits until the device is done with the picture's oldest job and retires it without delivering anything */
int me_gpu_cancel(uint64_t picture_number);
/* blocks until no job of the picture submitted before this call is in flight any more, i.e. until another thread's
 * me_gpu_wait has the results (the segment threads of svt_aom_motion_estimation_kernel); ME_GPU_OK also when
 * the job was retired already (one of the last 64) */
int me_gpu_wait_done(uint64_t picture_number);
/* Same as me_gpu_submit, with the destination of the results known up front: the device->host copies are queued
 * on the job's stream right behind the last kernel, so that me_gpu_wait(picture_number, out) only has to wait.
 * `out` (the struct and the arrays it points to) must stay valid and unchanged until that wait returns. */
int me_gpu_submit_out(uint64_t picture_number, const MeGpuParams *params,
                      const uint64_t ref_picture_numbers[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS], MeGpuOutputs *out);
/* Byte offsets of the 11 product arrays of MeGpuOutputs (total_me_candidate_index ... rc_me_allow_gm, in struct
 * order) inside the engine's per-picture result slab. A caller that carves its arrays out of ONE (pinned) buffer
 * at these offsets gets them with a single device->host copy; any other arrangement works, with one copy per
 * non-adjacent array. */
int me_gpu_output_layout(int width, int height, const MeGpuParams *params, size_t offsets[11], size_t *total_bytes);

/* What the temporal filter knows before svt_aom_sig_deriv_me_tf runs; input of me_gpu_derive_params_tf(). */
typedef struct MeGpuTfConfig {
    uint8_t  hme_me_level; /* pcs->tf_ctrls.hme_me_level, 0..4 (TfControls, definitions.h:166) */
    uint8_t  qp_opt; /* pcs->tf_ctrls.qp_opt */
    uint8_t  qp; /* scs->static_config.qp */
    uint8_t  is_ref; /* centre_pcs->is_ref */
    uint8_t  temporal_layer_index; /* centre_pcs->temporal_layer_index */
    uint16_t me_exit_th; /* pcs->tf_ctrls.me_exit_th */
    uint16_t width; /* luma size, multiple of 8 */
    uint16_t height;
    uint64_t picture_number; /* of the centre picture */
} MeGpuTfConfig;
/* host-side restatement of svt_aom_sig_deriv_me_tf (enc_mode_config.c:834-874), the tf_enable_hme_* switch
 * (enc_mode_config.c:1646-1668), set_hme_search_params_mctf(ctx, 0) and the per-call MeContext fields of
 * produce_temporally_filtered_pic (temporal_filtering.c:3144-3181): one list, one reference */
int me_gpu_derive_params_tf(const MeGpuTfConfig *cfg, MeGpuParams *params);

/* dynamic-GOP detector ------------------------------------------------------------------- */
typedef struct MeGpuDgMetrics { /* DGDetectorMetrics, pcs.h:731-737 (seg_completed has no meaning here) */
    uint64_t tot_dist; /* sum of the per-b64 HME-L0 SADs */
    uint32_t tot_cplx; /* b64s whose SAD exceeds 16*16*30 */
    uint32_t tot_active; /* b64s with a non-zero vector */
    int32_t  sum_in_vectors; /* inward minus outward pointing vector components */
} MeGpuDgMetrics;
/* One full-SAD svt_sad_loop_kernel search per b64 of the source's sixteenth plane (16x16 block) in the reference's
 * sixteenth plane, search area 16 / 64 / 128 squared by input_resolution (<= 360p / <= 480p / larger,
 * pd_process.c:497-498), reduced to the four picture metrics. Both pictures must be registered and of equal size.
 * Synchronous, like early_hme (which blocks on frame_done_sem): returns when *metrics is written. */
int me_gpu_dg_detector(uint64_t src_picture_number, uint64_t ref_picture_number, uint8_t input_resolution,
                       MeGpuDgMetrics *metrics);

/* The same search with NO allocation of any kind behind the call (no cudaMalloc / cudaFree, no stream or event creation,
 * no engine pool): for callers that cannot allocate - code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc. It has NO counterpart in the reference.
 * Works without me_gpu_init, on the current CUDA device. The caller passes the two PADDED sixteenth planes (host pointers
 * at the first byte of the padded plane, padding 16 on every side, stride >= width / 4 + 32; what me_gpu_read_pyramid
 * returns as `sixteenth`) and the luma size of the pictures (multiples of 8). Scratch - both planes and the accumulator -
 * is a file-scope device array of ME_GPU_DG_NOALLOC_MAX_BYTES that every GPU loading the library holds until exit: pictures
 * up to 3840x2160; anything larger is rejected with ME_GPU_ERR_NO_MEMORY. All calls share that array: calls from one
 * process are serialised, calls that overlap on the GPU from several processes' contexts do not share it, and a caller
 * that drives the GPU work itself must not overlap two of them. me_gpu_dg_detector_noalloc_setup() (optional; once per
 * device, that device current, outside any graph capture) resolves the array's device address and loads the kernel;
 * me_gpu_dg_detector_noalloc_error() is the (static) text of this entry point's last error. */
#define ME_GPU_DG_NOALLOC_MAX_BYTES (1280 * 1024)
int me_gpu_dg_detector_noalloc_setup(void);
int me_gpu_dg_detector_noalloc(const uint8_t *src_sixteenth, int src_stride, const uint8_t *ref_sixteenth, int ref_stride,
                               int width, int height, uint8_t input_resolution, MeGpuDgMetrics *metrics);
const char *me_gpu_dg_detector_noalloc_error(void);

/* host-side restatement of svt_aom_sig_deriv_me and the pcs/scs settings it depends on */
int me_gpu_derive_params(const MeGpuEncoderConfig *cfg, MeGpuParams *params);

/* measurement taps used by bench.py ------------------------------------------------------- */
/* device-only timing of the last waited picture on its stream (ms), per stage */
typedef struct MeGpuTimings {
    float pyramid_ms;
    float hme_ms;
    float search_ms;
    float epilogue_ms;
    float total_ms;
    uint64_t kernel_launches; /* cumulative launches of this library's kernels */
    uint64_t byte_sad_ops; /* algorithmic byte-SAD count of the last waited picture */
    uint64_t byte_sad_ops_fullsearch; /* the part of byte_sad_ops spent in the integer full search */
    /* per-kernel-class device time of the last waited picture; filled only in profiling mode
     * (me_gpu_set_profiling(1)), where each launch is bracketed by events on the job's stream */
    float k_search_ms; /* sum over the k_search launches (zz, pre-HME, HME L0/L1/L2, centre SADs) */
    float k_fullsearch_ms; /* sum over the two k_fullsearch launches (probe round, main round) */
    float k_fullsearch_main_ms; /* the main-round k_fullsearch launch alone */
    float k_ctrl_ms; /* sum over the k_ctrl launches */
    float k_epilogue_ms;
    uint64_t byte_sad_ops_fullsearch_inline; /* the part of byte_sad_ops_fullsearch searched inside k_hme (8x8-variance probe, areas
                                                of <= 256 positions) instead of as k_fullsearch tiles */
} MeGpuTimings;
int me_gpu_get_timings(MeGpuTimings *t);
/* profiling mode: bracket every kernel launch of a job with CUDA events (adds ~30 event records per picture) */
void me_gpu_set_profiling(int on);
/* measured VABSDIFF4 lane-operations per second of the engine's first device (x4 = byte-SADs / s): the
 * denominator of the integer-SAD roofline; 0.0 when the engine is not initialised */
double me_gpu_measure_sad_peak(void);

#ifdef __cplusplus
}
#endif
#endif /* ME_GPU_H */
