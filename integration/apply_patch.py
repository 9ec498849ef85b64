#!/usr/bin/env python
"""Applies the B200 ME binding to a COPY of the encoder's source tree (never to /root/reference itself).

    python integration/apply_patch.py <copy of the svt-av1-psyex tree>

What it changes (the same edits integration/encoder_me_gpu.patch shows as a unified diff):
  Source/Lib/Codec/svt_me_gpu_hook.[ch], me_gpu.h   added (copied from integration/ and include/)
  Source/Lib/Globals/enc_handle.c    svt_av1_enc_init: after svt_aom_setup_rtcd_internal, when the environment names the engine
                                     library (SVT_ME_GPU_LIB), svt_me_gpu_hook_init - the hook's pointers are set like RTCD
                                     pointers, once; svt_av1_enc_deinit: svt_me_gpu_hook_shutdown
  Source/Lib/Codec/me_process.c      svt_aom_motion_estimation_kernel, TASK_PAME: the b64 loop is replaced by
                                     svt_me_gpu_segment (whole picture on the device, once per picture) when the hook is live
  Source/Lib/Codec/pic_analysis_process.c  svt_aom_downsample_filtering_input_picture (picture analysis, the temporal filter's
                                     re-decimation, the overlay picture): svt_me_gpu_downsample - upload once, pyramid on the
                                     device, the quarter / sixteenth planes read back for the CPU stages that still use them
  Source/Lib/Codec/pd_process.c      early_hme (dynamic-GOP detector): svt_me_gpu_dg_detector in place of the TASK_DG_DETECTOR_HME
                                     segments and the semaphore wait, when the device holds both pyramids
  Source/Lib/Codec/temporal_filtering.c  produce_temporally_filtered_pic: svt_me_gpu_tf_block in place of the per-block
                                     svt_aom_motion_estimation_b64 call (one ME_MCTF job per (centre, neighbour) pair)
  Source/Lib/Codec/CMakeLists.txt    lists the new files; Source/Lib/CMakeLists.txt links libdl
Every edit is an exact-match replacement: the script fails if the tree is not the one it was written against.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def replace_once(path, old, new):
    s = open(path).read()
    if s.count(old) != 1:
        raise SystemExit(f"{path}: expected exactly one match of the anchor, found {s.count(old)}:\n{old[:200]}")
    open(path, "w").write(s.replace(old, new))


def main():
    root = sys.argv[1]
    codec = os.path.join(root, "Source", "Lib", "Codec")
    for f in ("svt_me_gpu_hook.c", "svt_me_gpu_hook.h"):
        shutil.copy(os.path.join(HERE, f), codec)
    shutil.copy(os.path.join(HERE, "..", "include", "me_gpu.h"), codec)

    # ---- init / deinit --------------------------------------------------------------------------------------------
    eh = os.path.join(root, "Source", "Lib", "Globals", "enc_handle.c")
    replace_once(eh, '''    svt_aom_setup_rtcd_internal(enc_handle_ptr->scs_instance_array[0]->scs->static_config.use_cpu_flags);

    svt_aom_asm_set_convolve_asm_table();
''', '''    svt_aom_setup_rtcd_internal(enc_handle_ptr->scs_instance_array[0]->scs->static_config.use_cpu_flags);
    {
        /* B200 open-loop ME: the hook's pointers are set here, once, like the RTCD pointers above; NULL = CPU path */
        const char *me_gpu_lib = getenv("SVT_ME_GPU_LIB");
        if (me_gpu_lib && *me_gpu_lib) {
            SequenceControlSet *s0 = enc_handle_ptr->scs_instance_array[0]->scs;
            if (svt_me_gpu_hook_init(me_gpu_lib, (s0->max_input_luma_width + 7) & ~7, (s0->max_input_luma_height + 7) & ~7, 4, 3,
                                     (int)s0->pa_reference_picture_buffer_init_count + 8) != EB_ErrorNone) {
                SVT_ERROR("SVT_ME_GPU_LIB is set but the B200 ME engine could not be brought up (no CPU fallback)\\n");
                return EB_ErrorInsufficientResources;
            }
            SVT_INFO("open-loop ME on the B200 engine: %s\\n", me_gpu_lib);
        }
    }

    svt_aom_asm_set_convolve_asm_table();
''')
    replace_once(eh, '''EB_API EbErrorType svt_av1_enc_deinit(EbComponentType *svt_enc_component) {
    if (!svt_enc_component || !svt_enc_component->p_component_private)
        return EB_ErrorBadParameter;
''', '''EB_API EbErrorType svt_av1_enc_deinit(EbComponentType *svt_enc_component) {
    if (!svt_enc_component || !svt_enc_component->p_component_private)
        return EB_ErrorBadParameter;
    svt_me_gpu_deinit_pending = 1;
''')
    # the engine goes down after the pipeline has drained: at the end of deinit
    s = open(eh).read()
    a = s.index("EB_API EbErrorType svt_av1_enc_deinit(EbComponentType *svt_enc_component) {")
    b = s.index("\n}\n", a)
    body = s[a:b]
    last_ret = body.rindex("    return ")
    body = body[:last_ret] + "    if (svt_me_gpu_deinit_pending && svt_me_gpu_live())\n        svt_me_gpu_hook_shutdown();\n" + body[last_ret:]
    s = s[:a] + body + s[b:]
    s = s.replace('EB_API EbErrorType svt_av1_enc_init(EbComponentType *svt_enc_component)\n{',
                  '#include "svt_me_gpu_hook.h"\nstatic int svt_me_gpu_deinit_pending = 0;\n'
                  'EB_API EbErrorType svt_av1_enc_init(EbComponentType *svt_enc_component)\n{', 1)
    open(eh, "w").write(s)

    # ---- the ME kernel ----------------------------------------------------------------------------------------------
    mp = os.path.join(codec, "me_process.c")
    replace_once(mp, '''                    // 64x64 Block Loop
                    for (uint32_t y_b64_index = y_b64_start_index; y_b64_index < y_b64_end_index; ++y_b64_index) {''',
                 '''                    if (svt_me_gpu_live() && !pcs->frame_superres_enabled && !pcs->frame_resize_enabled) {
                        /* B200 path: the whole picture on the device, once; this segment's thread blocks until the pcs
                           holds every b64's results, then accounts for its own b64s like the loop below does one by one */
                        MeContext *mc = me_context_ptr->me_ctx;
                        mc->me_type = ME_OPEN_LOOP;
                        mc->num_of_list_to_search = (pcs->slice_type == P_SLICE) ? 1 : 2;
                        mc->num_of_ref_pic_to_search[0] = pcs->ref_list0_count_try;
                        if (pcs->slice_type == B_SLICE)
                            mc->num_of_ref_pic_to_search[1] = pcs->ref_list1_count_try;
                        mc->temporal_layer_index = pcs->temporal_layer_index;
                        mc->is_ref = pcs->is_ref;
                        for (int i = 0; i < mc->num_of_list_to_search; i++)
                            for (int j = 0; j < mc->num_of_ref_pic_to_search[i]; j++) {
                                EbPaReferenceObject *ref_object = (EbPaReferenceObject *)pcs->ref_pa_pic_ptr_array[i][j]->object_ptr;
                                mc->me_ds_ref_array[i][j].picture_ptr = ref_object->input_padded_pic;
                                mc->me_ds_ref_array[i][j].quarter_picture_ptr = ref_object->quarter_downsampled_picture_ptr;
                                mc->me_ds_ref_array[i][j].sixteenth_picture_ptr = ref_object->sixteenth_downsampled_picture_ptr;
                                mc->me_ds_ref_array[i][j].picture_number = ref_object->picture_number;
                            }
                        if (svt_me_gpu_segment(pcs, mc) != EB_ErrorNone) {
                            SVT_ERROR("B200 ME engine failed on picture %llu (no CPU fallback)\\n", (unsigned long long)pcs->picture_number);
                            abort();
                        }
                        svt_block_on_mutex(pcs->me_processed_b64_mutex);
                        pcs->me_processed_b64_count += (x_b64_end_index - x_b64_start_index) * (y_b64_end_index - y_b64_start_index);
                        if (pcs->me_processed_b64_count == pcs->b64_total_count) {
                            if (pcs->gm_ctrls.enabled && (!pcs->gm_ctrls.pp_enabled || pcs->gm_pp_detected))
                                svt_aom_global_motion_estimation(pcs, input_pic);
                            else
                                memset(pcs->is_global_motion, false, MAX_NUM_OF_REF_PIC_LIST * REF_LIST_MAX_DEPTH);
                        }
                        svt_release_mutex(pcs->me_processed_b64_mutex);
                    } else
                    // 64x64 Block Loop
                    for (uint32_t y_b64_index = y_b64_start_index; y_b64_index < y_b64_end_index; ++y_b64_index) {''')
    s = open(mp).read()
    first_inc = s.index("#include")
    s = s[:first_inc] + '#include "svt_me_gpu_hook.h"\n#include "svt_log.h"\n' + s[first_inc:]
    open(mp, "w").write(s)

    # ---- the pyramid: every caller of the downsample takes the device's planes -------------------------------------
    pa = os.path.join(codec, "pic_analysis_process.c")
    replace_once(pa, '''                                                EbPictureBufferDesc *sixteenth_picture_ptr) {
    // Downsample input picture for HME L0 and L1
    if (pcs->enable_hme_flag || pcs->tf_enable_hme_flag) {''',
                 '''                                                EbPictureBufferDesc *sixteenth_picture_ptr) {
    if (svt_me_gpu_live() && pcs->pa_ref_pic_wrapper && (pcs->enable_hme_flag || pcs->tf_enable_hme_flag) &&
        (pcs->enable_hme_level0_flag || pcs->tf_enable_hme_level0_flag) &&
        (pcs->enable_hme_level1_flag || pcs->tf_enable_hme_level1_flag)) {
        /* B200 path: the device builds the pyramid ME will search (once per picture version) and hands the two planes
           back; only for the pcs's own PA reference object at the picture's own size (not the rescaled copies of
           super-res / resize, which stay on the CPU like their ME does) */
        const EbPaReferenceObject *own = (const EbPaReferenceObject *)pcs->pa_ref_pic_wrapper->object_ptr;
        if (quarter_picture_ptr == own->quarter_downsampled_picture_ptr &&
            sixteenth_picture_ptr == own->sixteenth_downsampled_picture_ptr && input_padded_pic->width == pcs->aligned_width &&
            input_padded_pic->height == pcs->aligned_height && !(pcs->aligned_width & 7) && !(pcs->aligned_height & 7)) {
            if (svt_me_gpu_downsample(pcs, input_padded_pic, quarter_picture_ptr, sixteenth_picture_ptr) != EB_ErrorNone) {
                SVT_ERROR("B200 ME engine failed to build the pyramid of picture %llu (no CPU fallback)\\n",
                          (unsigned long long)pcs->picture_number);
                abort();
            }
            return;
        }
    }
    // Downsample input picture for HME L0 and L1
    if (pcs->enable_hme_flag || pcs->tf_enable_hme_flag) {''')
    s = open(pa).read()
    first_inc = s.index("#include")
    s = s[:first_inc] + '#include "svt_me_gpu_hook.h"\n#include "svt_log.h"\n' + s[first_inc:]
    open(pa, "w").write(s)

    # ---- dynamic-GOP detector: one device call instead of the TASK_DG_DETECTOR_HME segments -------------------------
    pd = os.path.join(codec, "pd_process.c")
    replace_once(pd, '''    // create segments for the dg detector and send them to the motion estimation kernel
    for (uint16_t seg_idx = 0; seg_idx < dg_detector_seg_total_count; ++seg_idx) {
''', '''    if (svt_me_gpu_live() && svt_me_gpu_dg_ready(src_pcs, ref_pcs)) {
        /* B200 path: HME level 0 of the whole picture pair in one launch over the two device-resident sixteenth planes */
        if (svt_me_gpu_dg_detector(src_pcs) != EB_ErrorNone) {
            SVT_ERROR("B200 ME engine failed in the dynamic-GOP detector (no CPU fallback)\\n");
            abort();
        }
    } else {
    // create segments for the dg detector and send them to the motion estimation kernel
    for (uint16_t seg_idx = 0; seg_idx < dg_detector_seg_total_count; ++seg_idx) {
''')
    replace_once(pd, '''    // wait for all segments to complete before the frame based calculations can be performed using the dg metrics
    svt_block_on_semaphore(src_pcs->dg_detector->frame_done_sem);
''', '''    // wait for all segments to complete before the frame based calculations can be performed using the dg metrics
    svt_block_on_semaphore(src_pcs->dg_detector->frame_done_sem);
    }
''')
    s = open(pd).read()
    first_inc = s.index("#include")
    s = s[:first_inc] + '#include "svt_me_gpu_hook.h"\n#include "svt_log.h"\n' + s[first_inc:]
    open(pd, "w").write(s)

    # ---- temporal filter: the ME of a (centre, neighbour) pair on the device, once, for all blocks -------------------
    tf = os.path.join(codec, "temporal_filtering.c")
    replace_once(tf, '''                    svt_aom_motion_estimation_b64(centre_pcs,
                        (uint32_t)blk_row * blk_cols + blk_col,
                        (uint32_t)blk_col * BW, // x block
                        (uint32_t)blk_row * BH, // y block
                        ctx,
                        input_picture_ptr_central); // source picture
''', '''                    if (svt_me_gpu_tf_ready(centre_pcs, pcs_list[frame_index])) { /* 8- and 10-bit alike: ME reads the 8-bit y8b planes */
                        /* B200 path: the first segment thread that needs this (centre, neighbour) pair runs its ME for all
                           blocks in one job; this block's results are loaded where the filter reads them */
                        if (svt_me_gpu_tf_block(centre_pcs, ctx, pcs_list[frame_index], frame_index, blk_rows * blk_cols,
                                                (uint32_t)blk_row * blk_cols + blk_col) != EB_ErrorNone) {
                            SVT_ERROR("B200 ME engine failed in the temporal filter (no CPU fallback)\\n");
                            abort();
                        }
                    } else
                    svt_aom_motion_estimation_b64(centre_pcs,
                        (uint32_t)blk_row * blk_cols + blk_col,
                        (uint32_t)blk_col * BW, // x block
                        (uint32_t)blk_row * BH, // y block
                        ctx,
                        input_picture_ptr_central); // source picture
''')
    replace_once(tf, '''    svt_block_on_mutex(centre_pcs->temp_filt_mutex);
    centre_pcs->temp_filt_seg_acc++;
''', '''    svt_me_gpu_tf_centre_done(centre_pcs); /* the last segment frees the pairs' results */
    svt_block_on_mutex(centre_pcs->temp_filt_mutex);
    centre_pcs->temp_filt_seg_acc++;
''')
    s = open(tf).read()
    first_inc = s.index("#include")
    s = s[:first_inc] + '#include "svt_me_gpu_hook.h"\n#include "svt_log.h"\n' + s[first_inc:]
    open(tf, "w").write(s)

    # ---- the new source joins the CODEC object library ----
    replace_once(os.path.join(codec, "CMakeLists.txt"), "set(all_files\n        adaptive_mv_pred.c\n",
                 "set(all_files\n        svt_me_gpu_hook.c\n        svt_me_gpu_hook.h\n        me_gpu.h\n        adaptive_mv_pred.c\n")

    # ---- link libdl --------------------------------------------------------------------------------------------------
    cm = os.path.join(root, "Source", "Lib", "CMakeLists.txt")
    replace_once(cm, "target_link_libraries(SvtAv1Enc PUBLIC ${PLATFORM_LIBS})",
                 "target_link_libraries(SvtAv1Enc PUBLIC ${PLATFORM_LIBS} ${CMAKE_DL_LIBS})")
    print("patched", root)


if __name__ == "__main__":
    main()
