/*
 * me_oracle.c - CPU restatement of SVT-AV1-PSYEX open-loop ME (pyramid, pre-HME, HME L0/L1/L2,
 * integer full search, candidate construction, distortion statistics).
 *
 * TEST INFRASTRUCTURE ONLY - see me_oracle.h. Parity PINNED against oracle/_ref (the reference's own
 * C sources compiled in place) by tests/test_oracle_vs_ref.py and the frozen tests/golden vectors.
 *
 * File:line citations are into /root/reference/Source/Lib (ME = Codec/motion_estimation.c).
 */
#include "me_oracle.h"

#include <stdlib.h>
#include <string.h>

#define OR_MIN(a, b) ((a) < (b) ? (a) : (b))
#define OR_MAX(a, b) ((a) > (b) ? (a) : (b))
#define OR_ABS(a) ((a) < 0 ? -(a) : (a))
#define OR_MAX_SAD_VALUE (128 * 128 * 255) /* Codec/motion_estimation.h:85 */
#define OR_MAX_U32 0xFFFFFFFFu

/* ------------------------------------------------------------------------------------------------
 * Stage 1: pyramid
 * ---------------------------------------------------------------------------------------------- */

/* Codec/pic_analysis_process.c:130-160 (svt_aom_downsample_2d_c): 2x2 box average sampled every
 * `step` pixels starting at step/2, rounding (sum+2)>>2. */
void oracle_downsample_2d(const uint8_t *in, uint32_t in_stride, uint32_t in_w, uint32_t in_h, uint8_t *out,
                          uint32_t out_stride, uint32_t step) {
    const uint32_t half = step >> 1;
    uint32_t       oy   = 0;
    for (uint32_t y = half; y < in_h; y += step, oy++) {
        const uint8_t *cur  = in + (size_t)y * in_stride;
        const uint8_t *prev = cur - in_stride;
        uint32_t       ox   = 0;
        for (uint32_t x = half; x < in_w; x += step, ox++) {
            uint32_t s = (uint32_t)prev[x - 1] + prev[x] + cur[x - 1] + cur[x];
            out[(size_t)oy * out_stride + ox] = (uint8_t)((s + 2) >> 2);
        }
    }
}

/* Codec/pic_operators.c:397-443 (svt_aom_generate_padding): replicate left/right edge samples, then
 * copy the first/last padded row upwards/downwards (so corners hold the corner sample). */
void oracle_generate_padding(uint8_t *buf, uint32_t stride, uint32_t width, uint32_t height, uint32_t pad_w,
                             uint32_t pad_h) {
    uint8_t *row = buf + pad_w + (size_t)pad_h * stride;
    for (uint32_t y = 0; y < height; y++, row += stride) {
        memset(row - pad_w, row[0], pad_w);
        memset(row + width, row[width - 1], pad_w);
    }
    uint8_t *top = buf + (size_t)pad_h * stride;
    uint8_t *bot = buf + (size_t)(pad_h + height - 1) * stride;
    for (uint32_t i = 1; i <= pad_h; i++) {
        memcpy(top - (size_t)i * stride, top, stride);
        memcpy(bot + (size_t)i * stride, bot, stride);
    }
}

/* Codec/pic_analysis_process.c:2139-2196 (svt_aom_downsample_filtering_input_picture) with HME L0 and
 * L1 enabled: quarter from full (step 2), sixteenth from quarter (step 2), each padded afterwards.
 * The destination origin is org_x + org_x*stride as in the reference (:2151,:2172); pads are square. */
void oracle_build_pyramid(OraclePicture *pic) {
    OraclePlane *f = &pic->full, *q = &pic->quarter, *s = &pic->sixteenth;
    oracle_downsample_2d(f->buf + f->org_x + (size_t)f->org_y * f->stride, f->stride, f->width, f->height,
                         q->buf + q->org_x + (size_t)q->org_x * q->stride, q->stride, 2);
    oracle_generate_padding(q->buf, q->stride, q->width, q->height, q->org_x, q->org_y);
    oracle_downsample_2d(q->buf + q->org_x + (size_t)q->org_y * q->stride, q->stride, q->width, q->height,
                         s->buf + s->org_x + (size_t)s->org_x * s->stride, s->stride, 2);
    oracle_generate_padding(s->buf, s->stride, s->width, s->height, s->org_x, s->org_y);
}

/* ------------------------------------------------------------------------------------------------
 * Leaf SAD kernels
 * ---------------------------------------------------------------------------------------------- */

/* C_DEFAULT/compute_sad_c.c:20-37 (svt_fast_loop_nxm_sad_kernel) */
uint32_t oracle_nxm_sad(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                        uint32_t height, uint32_t width) {
    uint32_t sad = 0;
    for (uint32_t y = 0; y < height; y++) {
        for (uint32_t x = 0; x < width; x++) sad += (uint32_t)abs((int)src[x] - (int)ref[x]);
        src += src_stride;
        ref += ref_stride;
    }
    return sad;
}

/* C_DEFAULT/compute_sad_c.c:58-101 (svt_sad_loop_kernel_c): raster scan of the search area, strict '<'
 * keeps the first minimum; best starts at 0xffffff and the outputs are untouched when nothing beats it.
 * Even search rows are skipped when skip_search_line and the block is 16 wide and at most 16 high. */
void oracle_sad_loop(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                     uint32_t block_height, uint32_t block_width, uint64_t *best_sad, int16_t *x_best, int16_t *y_best,
                     uint32_t ref_stride_raw, uint8_t skip_search_line, int16_t sa_width, int16_t sa_height) {
    *best_sad = 0xffffff;
    const int skip = (block_width == 16 && block_height <= 16 && skip_search_line);
    for (int y = 0; y < sa_height; y++, ref += ref_stride_raw) {
        if (skip && (y & 1) == 0)
            continue;
        for (int x = 0; x < sa_width; x++) {
            uint32_t sad = oracle_nxm_sad(src, src_stride, ref + x, ref_stride, block_height, block_width);
            if (sad < *best_sad) {
                *best_sad = sad;
                *x_best   = (int16_t)x;
                *y_best   = (int16_t)y;
            }
        }
    }
}
