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

#ifdef __cplusplus
}
#endif
#endif
