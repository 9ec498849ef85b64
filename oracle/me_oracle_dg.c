/*
 * me_oracle_dg.c - CPU restatement of the dynamic-GOP detector's HME level 0 (TASK_DG_DETECTOR_HME).
 *
 * TEST INFRASTRUCTURE ONLY - see me_oracle.h. Parity PINNED against oracle/_ref (pd_process.c compiled in
 * place, called through ref_dg_detector of ref_harness.c) by tests/test_oracle.py and tests/golden.
 *
 * File:line citations are into /root/reference/Source/Lib/Codec/pd_process.c.
 */
#include "me_oracle.h"

#define DG_MAX(a, b) ((a) > (b) ? (a) : (b))

/* :393-484 (early_hme_b64): HME-L0 style clamp of a centred search area on the sixteenth plane, one
 * svt_sad_loop_kernel over it, vector scaled by 4. All window arithmetic is int16_t as in the reference;
 * the left / top corrections move the origin and leave the size as it was (the subtracted term is zero
 * once the origin has been replaced). */
static void dg_early_hme_b64(const uint8_t *blk, uint32_t blk_stride, int full_sad, int16_t org_x, int16_t org_y,
                             uint32_t block_width, uint32_t block_height, int16_t sa_width, int16_t sa_height,
                             const OraclePlane *ref, uint64_t *best_sad, int16_t *mv_col, int16_t *mv_row) {
    sa_width           = (int16_t)((sa_width + 7) & ~0x07);
    int16_t pad_width  = (int16_t)(ref->org_x) - 1;
    int16_t pad_height = (int16_t)(ref->org_y) - 1;
    int16_t sa_origin_x = -(int16_t)(sa_width >> 1);
    int16_t sa_origin_y = -(int16_t)(sa_height >> 1);

    if ((org_x + sa_origin_x) < -pad_width) {
        sa_origin_x = -pad_width - org_x;
        sa_width    = sa_width - (-pad_width - (org_x + sa_origin_x));
    }
    if ((org_x + sa_origin_x) > (int16_t)ref->width - 1)
        sa_origin_x = sa_origin_x - ((org_x + sa_origin_x) - ((int16_t)ref->width - 1));
    if ((org_x + sa_origin_x + sa_width) > (int16_t)ref->width)
        sa_width = DG_MAX(1, sa_width - ((org_x + sa_origin_x + sa_width) - (int16_t)ref->width));
    sa_width = (sa_width < 8) ? sa_width : sa_width & ~0x07;
    if ((org_y + sa_origin_y) < -pad_height) {
        sa_origin_y = -pad_height - org_y;
        sa_height   = sa_height - (-pad_height - (org_y + sa_origin_y));
    }
    if ((org_y + sa_origin_y) > (int16_t)ref->height - 1)
        sa_origin_y = sa_origin_y - ((org_y + sa_origin_y) - ((int16_t)ref->height - 1));
    if ((org_y + sa_origin_y + sa_height) > (int16_t)ref->height)
        sa_height = DG_MAX(1, sa_height - ((org_y + sa_origin_y + sa_height) - (int16_t)ref->height));

    int16_t  x_tl = ((int16_t)ref->org_x + org_x) + sa_origin_x;
    int16_t  y_tl = ((int16_t)ref->org_y + org_y) + sa_origin_y;
    uint32_t idx  = (uint32_t)(x_tl + y_tl * ref->stride);

    oracle_sad_loop(blk, full_sad ? blk_stride : blk_stride * 2, ref->buf + idx, full_sad ? ref->stride : ref->stride * 2,
                    full_sad ? block_height : block_height >> 1, block_width, best_sad, mv_col, mv_row, (uint32_t)ref->stride,
                    0, sa_width, sa_height);
    *best_sad = full_sad ? *best_sad : *best_sad * 2;
    *mv_col   = (int16_t)((*mv_col + sa_origin_x) * 4);
    *mv_row   = (int16_t)((*mv_row + sa_origin_y) * 4);
}

/* :492-560 (dg_detector_hme_level0) over all segments of the picture, with early_hme's reset of the
 * metrics (:569-573). The best SAD and the vector are locals that live across b64s in the reference (a
 * search that finds nothing below 0xffffff would leave the previous vector in place); kept the same. */
int oracle_dg_detector(const OraclePicture *src, const OraclePicture *ref, uint8_t input_resolution,
                       MeGpuDgMetrics *out) {
    if (!src || !ref || !out)
        return -1;
    const OraclePlane *ss = &src->sixteenth, *rs = &ref->sixteenth;
    const int16_t      sa = input_resolution <= 1 ? 16 : input_resolution <= 2 ? 64 : 128; /* :497-498 */
    const uint32_t     aligned_w = (uint32_t)(src->full.width + 7) & ~7u, aligned_h = (uint32_t)(src->full.height + 7) & ~7u;
    const uint32_t     bw = (aligned_w + 63) / 64, bh = (aligned_h + 63) / 64;
    uint64_t           sad = (uint64_t)~0;
    int16_t            col = 0, row = 0;
    out->tot_dist = 0;
    out->tot_cplx = out->tot_active = 0;
    out->sum_in_vectors = 0;
    for (uint32_t y = 0; y < bh; y++)
        for (uint32_t x = 0; x < bw; x++) {
            const uint32_t ox = x * 64, oy = y * 64;
            const uint8_t *blk = ss->buf + (size_t)(ss->org_y + (oy >> 2)) * ss->stride + ss->org_x + (ox >> 2);
            dg_early_hme_b64(blk, (uint32_t)ss->stride, 1, (int16_t)ox >> 2, (int16_t)oy >> 2, 16, 16, sa, sa, rs, &sad, &col,
                             &row);
            out->tot_dist += sad;
            out->tot_cplx += (sad > (16 * 16 * 30));
            out->tot_active += (col != 0 || row != 0);
            /* :530-551: does the vector point towards the picture centre or away from it */
            if (y < bh / 2)
                out->sum_in_vectors += row > 0 ? -1 : row < 0 ? 1 : 0;
            else if (y > bh / 2)
                out->sum_in_vectors += row > 0 ? 1 : row < 0 ? -1 : 0;
            if (x < bw / 2)
                out->sum_in_vectors += col > 0 ? -1 : col < 0 ? 1 : 0;
            else if (x > bw / 2)
                out->sum_in_vectors += col > 0 ? 1 : col < 0 ? -1 : 0;
        }
    return 0;
}
