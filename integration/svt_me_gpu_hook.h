/*
 * svt_me_gpu_hook.h - the reference-side binding of the B200 ME engine: what a maintainer adds to
 * Source/Lib/Codec. Written against the encoder's own types (PictureParentControlSet, MeContext, EbPictureBufferDesc),
 * so it compiles only inside the encoder tree (or, for the tests, next to it: oracle/Makefile builds it with the
 * reference's include paths into oracle/_ref/libme_ref.so).
 *
 * The engine is reached through RTCD-style pointers that stay NULL unless svt_me_gpu_hook_init() found the library:
 * the CPU path is untouched when they are NULL. There is no fallback once they are set: a failing call fails the encode.
 */
#ifndef SVT_ME_GPU_HOOK_H
#define SVT_ME_GPU_HOOK_H

#include "me_gpu.h"

#include "definitions.h"
#include "me_context.h"
#include "pcs.h"
#include "sequence_control_set.h"

#ifdef __cplusplus
extern "C" {
#endif

/* set once by svt_me_gpu_hook_init (called from svt_av1_enc_init after svt_aom_setup_rtcd_internal, enc_handle.c:1445) */
extern int (*svt_me_picture_register)(uint64_t, const uint8_t *, int, int, int, int, int);
extern int (*svt_me_picture_submit)(uint64_t, const MeGpuParams *, const uint64_t[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS]);
extern int (*svt_me_picture_wait)(uint64_t, MeGpuOutputs *);
extern int (*svt_me_picture_release)(uint64_t);

/* dlopen(library_path) + me_gpu_init; EB_ErrorNone when the hook is live */
EbErrorType svt_me_gpu_hook_init(const char *library_path, int max_width, int max_height, int max_refs_l0, int max_refs_l1,
                                 int pictures_in_flight);
void        svt_me_gpu_hook_shutdown(void);

/* the MeContext fields written by svt_aom_sig_deriv_me (enc_mode_config.c:681-828) -> MeGpuParams */
void svt_me_gpu_params_from_me_ctx(MeGpuParams *p, const MeContext *m);
/* everything svt_aom_motion_estimation_b64 reads: the above plus the pcs / scs scalars and the per-call MeContext fields
 * of me_process.c:195-262 */
void svt_me_gpu_params_from_ctx(MeGpuParams *p, const PictureParentControlSet *pcs, const MeContext *m);

/* picture analysis, after svt_aom_pad_input_pictures (pic_analysis_process.c:2198): upload the padded y8b luma; the
 * quarter and sixteenth planes are built on the device */
EbErrorType svt_me_gpu_register_picture(const PictureParentControlSet *pcs, const EbPictureBufferDesc *input_padded_pic);
/* where the PA reference object is released (reference_object.c:406) */
EbErrorType svt_me_gpu_release_picture(uint64_t picture_number);

/* replaces the b64 loop of svt_aom_motion_estimation_kernel for TASK_PAME (me_process.c:174-290): open-loop ME of the
 * whole picture; fills pcs->pa_me_data->me_results[b64] and the eight per-b64 pcs arrays exactly as
 * svt_aom_motion_estimation_b64 does b64 by b64. me_ctx carries the controls and me_ds_ref_array[l][r].picture_number. */
EbErrorType svt_me_gpu_picture(PictureParentControlSet *pcs, const MeContext *me_ctx);

/* What svt_aom_motion_estimation_kernel calls for one TASK_PAME segment item of a picture in place of its b64 loop
 * (me_process.c:174-290). The picture's 48 segment items arrive on up to 25 ME threads in any order: the FIRST thread to
 * arrive registers whatever pyramids the device does not hold yet (the picture's own PA reference object and those of its
 * references, lazily, keyed by the object: a recycled object releases the picture it held before), submits the whole
 * picture, waits and scatters the results into the pcs; every other segment's thread blocks here until that is done.
 * When this returns EB_ErrorNone the pcs holds the results of ALL b64s; the caller then adds its segment's b64 count to
 * pcs->me_processed_b64_count under me_processed_b64_mutex and the thread that completes the count runs
 * svt_aom_global_motion_estimation, exactly as at me_process.c:274-287. me_ctx: controls by svt_aom_sig_deriv_me, the
 * per-call fields of me_process.c:219-262 filled for the picture. Not for super-res / resize pictures (their references
 * are rescaled planes): the caller keeps the CPU loop for those. */
EbErrorType svt_me_gpu_segment(PictureParentControlSet *pcs, MeContext *me_ctx);
/* non-zero when svt_me_gpu_hook_init succeeded: the ME kernel takes the device path */
int svt_me_gpu_live(void);
/* pictures the hook has registered / released so far (diagnostics) */
void svt_me_gpu_stats(uint64_t *registered, uint64_t *released, uint64_t *pictures);

/* ---- picture analysis takes the device-built pyramid --------------------------------------------- */
/* in svt_aom_picture_analysis_kernel, in place of svt_aom_downsample_filtering_input_picture (pic_analysis_process.c:2345):
 * uploads the padded y8b luma (as svt_me_gpu_register_picture does) and fills the PA reference object's quarter and
 * sixteenth planes, padding included, from the device: byte for byte what downsample_2d + svt_aom_generate_padding leave
 * there, so svt_aom_gathering_picture_statistics (:2356) and every later CPU reader see the same planes. The picture
 * stays registered under pcs->picture_number. */
EbErrorType svt_me_gpu_pyramid_for_analysis(const PictureParentControlSet *pcs, const EbPictureBufferDesc *input_padded_pic,
                                            EbPictureBufferDesc *quarter_pic, EbPictureBufferDesc *sixteenth_pic);

/* the same, for the encoder itself: called at the top of svt_aom_downsample_filtering_input_picture
 * (pic_analysis_process.c:2139), so picture analysis, the temporal filter's pad_and_decimate_filtered_pic
 * (temporal_filtering.c:3956) and the overlay path (pd_process.c:2735) all take the device's planes. Keyed by the pcs's
 * PA reference object like svt_me_gpu_segment's lazy registration, which then finds the picture already on the device;
 * a second call for the same object (the temporal filter rewrote the picture) replaces the first version. */
EbErrorType svt_me_gpu_downsample(const PictureParentControlSet *pcs, const EbPictureBufferDesc *pic, EbPictureBufferDesc *quarter_pic,
                                  EbPictureBufferDesc *sixteenth_pic);
uint64_t    svt_me_gpu_downsample_count(void);

/* ---- temporal filter (ME_MCTF, temporal_filtering.c:3144-3187) ------------------------------------------------ */
/* ME results of one (centre, neighbour) pair over all blocks: what the block loop of produce_temporally_filtered_pic
 * reads from the MeContext after its svt_aom_motion_estimation_b64 call. Arrays are caller-owned, b64_count long. */
typedef struct SvtMeGpuTfPair {
    uint32_t  b64_count;
    uint32_t *best_sad; /* [b64][85] */
    uint32_t *best_mv; /* [b64][85] */
    int16_t  *hme_centre; /* [b64][2] */
    uint32_t *hme_sad; /* [b64] */
    uint8_t  *early_exit; /* [b64] */
} SvtMeGpuTfPair;
/* once per (centre, neighbour), before the block loop; me_ctx set up as at temporal_filtering.c:3144-3181 (controls by
 * svt_aom_sig_deriv_me_tf, me_type = ME_MCTF, tf_me_exit_th). Adds the pair's tf_tot_horz/vert_blks to me_ctx. */
EbErrorType svt_me_gpu_tf_pair(const PictureParentControlSet *centre_pcs, MeContext *me_ctx, uint64_t neighbour_picture_number,
                               SvtMeGpuTfPair *pair);
/* in the block loop, instead of svt_aom_motion_estimation_b64 (:3182), for block b: puts the block's results where the
 * filter's readers expect them (p_sb_best_sad/mv[0][0], search_results[0][0], the 64x64-only switch) */
void svt_me_gpu_tf_load_block(MeContext *me_ctx, const SvtMeGpuTfPair *pair, uint32_t b);

/* inside the encoder: at the svt_aom_motion_estimation_b64 call of produce_temporally_filtered_pic (temporal_filtering.c:3182),
 * on whichever segment thread gets there first for a (centre, neighbour) pair: ME_MCTF of ALL blocks of the pair on the
 * device, once; every block of every segment then loads its own results (svt_me_gpu_tf_load_block). me_ctx set up as at
 * :3144-3181. svt_me_gpu_tf_centre_done: once per segment thread after the block loop; the last one frees the pairs. */
int         svt_me_gpu_tf_ready(const PictureParentControlSet *centre_pcs, const PictureParentControlSet *neighbour_pcs);
EbErrorType svt_me_gpu_tf_block(const PictureParentControlSet *centre_pcs, MeContext *me_ctx, const PictureParentControlSet *neighbour_pcs,
                                int frame_index, uint32_t b64_count, uint32_t b);
void        svt_me_gpu_tf_centre_done(const PictureParentControlSet *centre_pcs);

/* ---- dynamic-GOP detector (early_hme, pd_process.c:561-625) ---------------------------------------------------- */
/* replaces the TASK_DG_DETECTOR_HME segment dispatch + semaphore wait: fills src_pcs->dg_detector->metrics for
 * (src_pcs, src_pcs->dg_detector->ref_pic); both pictures registered */
EbErrorType svt_me_gpu_dg_detector(PictureParentControlSet *src_pcs);
/* non-zero when the hook is live and the device holds the pyramids of both pictures (they passed through
 * svt_me_gpu_downsample): early_hme then calls svt_me_gpu_dg_detector instead of posting its segments */
int svt_me_gpu_dg_ready(const PictureParentControlSet *src_pcs, const PictureParentControlSet *ref_pcs);

#ifdef __cplusplus
}
#endif
#endif
