/*
 * me_oracle.h - CPU restatement of SVT-AV1-PSYEX open-loop motion estimation.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is part of the product path: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library,
 * and only as the checker. The CUDA engine (svt-av1-psyex_b200/csrc) never links or calls it.
 *
 * Parity status: PINNED. The restatement is checked bit-for-bit against the reference's own C code
 * compiled from /root/reference (oracle/_ref/libme_ref.so, built by oracle/Makefile) by
 * tests/test_oracle_vs_ref.py, and against golden vectors that script froze into tests/golden/.
 *
 * Every function cites the reference lines it restates (paths relative to Source/Lib).
 */
#ifndef ME_ORACLE_H
#define ME_ORACLE_H

#include <stdint.h>
#include "me_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

/* One padded 8-bit luma plane: sample (x,y) at buf[(org_y + y) * stride + org_x + x]
 * (EbPictureBufferDesc, Codec/pic_buffer_desc.h:30-65). */
typedef struct OraclePlane {
    uint8_t *buf;
    int32_t  stride;
    int32_t  org_x;
    int32_t  org_y;
    int32_t  width;
    int32_t  height;
} OraclePlane;

/* EbPaReferenceObject planes (Codec/reference_object.h:70-86) */
typedef struct OraclePicture {
    OraclePlane full; /* input_padded_pic, pad 68 */
    OraclePlane quarter; /* quarter_downsampled_picture_ptr, pad 32 */
    OraclePlane sixteenth; /* sixteenth_downsampled_picture_ptr, pad 16 */
    uint64_t    picture_number;
} OraclePicture;

/* leaf kernels -------------------------------------------------------------------------- */
void     oracle_downsample_2d(const uint8_t *in, uint32_t in_stride, uint32_t in_w, uint32_t in_h, uint8_t *out,
                              uint32_t out_stride, uint32_t step);
void     oracle_generate_padding(uint8_t *buf, uint32_t stride, uint32_t width, uint32_t height, uint32_t pad_w,
                                 uint32_t pad_h);
void     oracle_build_pyramid(OraclePicture *pic);
uint32_t oracle_nxm_sad(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                        uint32_t height, uint32_t width);
void     oracle_sad_loop(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                         uint32_t block_height, uint32_t block_width, uint64_t *best_sad, int16_t *x_best,
                         int16_t *y_best, uint32_t ref_stride_raw, uint8_t skip_search_line, int16_t sa_width,
                         int16_t sa_height);

/* whole path ---------------------------------------------------------------------------- */
/* ME of every b64 of `src` against refs[list][idx] (NULL for unused slots); fills `out`
 * (all arrays caller-allocated and zero-initialised). Returns 0, or <0 on bad arguments. */
int oracle_me_picture(const OraclePicture *src, const OraclePicture *const refs[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS],
                      const MeGpuParams *params, MeGpuOutputs *out);
/* Same for b64 indices [first, first+count) only (bounded CPU-baseline samples). Output slabs are still
 * indexed by absolute b64 index. *byte_sad_ops accumulates the algorithmic byte-SAD count. */
int oracle_me_b64_range(const OraclePicture *src,
                        const OraclePicture *const refs[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS],
                        const MeGpuParams *params, MeGpuOutputs *out, uint32_t first, uint32_t count,
                        uint64_t *byte_sad_ops);

/* dynamic-GOP detector (me_oracle_dg.c): HME level 0 of every b64 of src against ref on the sixteenth planes,
 * reduced to the DGDetectorMetrics of the pair */
int oracle_dg_detector(const OraclePicture *src, const OraclePicture *ref, uint8_t input_resolution, MeGpuDgMetrics *out);

#ifdef __cplusplus
}
#endif
#endif
