/*
 * me_types.h - device/host shared data layout of the B200 open-loop ME engine.
 *
 * Everything the kernels exchange lives in HBM in these structures:
 *   picture pool   one slab per registered picture: padded full / quarter / sixteenth luma planes
 *                  (EbPaReferenceObject, Codec/reference_object.h:70-86) with 128-byte-aligned strides
 *   B64State       the per-64x64-block part of MeContext (Codec/me_context.h:366-509) that survives
 *                  between stages: HME centres and SADs, search_results, pre-HME data, zz SADs
 *   SearchTask     one svt_sad_loop_kernel call (Codec/aom_dsp_rtcd.h:779) described as data
 *   FsTask/FsTile  one open_loop_me_fullpel_search_sblock call (Codec/motion_estimation.c:781) and its tiles
 *   best keys      p_sb_best_sad / p_sb_best_mv (me_context.h:395-396) as 64-bit (sad << 32 | scan order) keys
 *                  so that a parallel minimum reproduces the reference's first-minimum tie-break
 */
#ifndef ME_TYPES_H
#define ME_TYPES_H

#include <stdint.h>

#include "me_gpu.h"

#if defined(__CUDACC__)
#define ME_HD __host__ __device__ __forceinline__
#else
#define ME_HD static inline
#endif

#define ME_NL ME_GPU_MAX_LISTS
#define ME_NR ME_GPU_MAX_REFS
#define ME_NREF (ME_NL * ME_NR) /* (list, ref) slots per b64 */
#define ME_NPU ME_GPU_SQUARE_PU_COUNT
#define ME_MAX_SAD_VALUE (128u * 128u * 255u) /* Codec/motion_estimation.h:85 */
#define ME_MAX_U32 0xFFFFFFFFu
#define ME_SAD_LOOP_INIT 0xffffffu /* svt_sad_loop_kernel_c initial best, C_DEFAULT/compute_sad_c.c:66 */

/* logical paddings of the three planes (reference_object.c:249-295); the clamps in the HME stages use
 * pad - 1 of the plane they search */
#define ME_PAD_FULL 68
#define ME_PAD_QUARTER 32
#define ME_PAD_SIXTEENTH 16

/* search-task slots per b64 and stage */
#define ME_SLOTS 32 /* 8 (list, ref) x 4 quadrants (HME) ; 8 x 2 regions (pre-HME) ; 8 x 2 (zero / centre SADs) */

typedef struct MePlaneGeom {
    int32_t  width, height; /* interior size */
    int32_t  pad; /* logical = physical padding on all four sides */
    int32_t  stride; /* bytes per row, multiple of 128 */
    int32_t  rows; /* height + 2 * pad */
    uint32_t offset; /* byte offset of the plane (first padded row) inside a pool slot */
} MePlaneGeom;

typedef struct MePicGeom {
    int32_t     width, height; /* aligned luma size (multiple of 8) */
    int32_t     b64_cols, b64_rows, b64_count;
    MePlaneGeom plane[3]; /* 0 full, 1 quarter, 2 sixteenth */
    uint64_t    slot_bytes;
} MePicGeom;

typedef struct MePreHme { /* SearchInfo, Codec/me_context.h:329 */
    uint16_t sa_w, sa_h;
    int16_t  col, row;
    uint32_t sad;
    uint8_t  valid;
    uint8_t  pad_[3];
} MePreHme;

typedef struct B64State {
    /* HME centres and SADs: [list][ref][sr_w][sr_h] (me_context.h:430-454) */
    int16_t  l0x[ME_NL][ME_NR][2][2], l0y[ME_NL][ME_NR][2][2];
    int16_t  l1x[ME_NL][ME_NR][2][2], l1y[ME_NL][ME_NR][2][2];
    int16_t  l2x[ME_NL][ME_NR][2][2], l2y[ME_NL][ME_NR][2][2];
    uint32_t l0sad[ME_NL][ME_NR][2][2], l1sad[ME_NL][ME_NR][2][2], l2sad[ME_NL][ME_NR][2][2];
    /* SearchResults, me_context.h:348 */
    int16_t  sc_x[ME_NL][ME_NR], sc_y[ME_NL][ME_NR];
    uint32_t hme_sad[ME_NL][ME_NR];
    uint8_t  do_ref[ME_NL][ME_NR];
    uint32_t sr_div[ME_NL][ME_NR];
    uint32_t zz_sad[ME_NL][ME_NR];
    MePreHme phme[ME_NL][ME_NR][2];
    uint8_t  performed_phme[ME_NL][ME_NR][2];
    /* integer search: centre after check_00_center, area before / after the 8x8-variance probe */
    int16_t  cx[ME_NL][ME_NR], cy[ME_NL][ME_NR];
    int16_t  sw[ME_NL][ME_NR], sh[ME_NL][ME_NR];
    int16_t  sa_final[ME_NL][ME_NR][4]; /* x0, y0, w, h of the main search */
    uint8_t  probed[ME_NL][ME_NR]; /* the centre position was searched first (scan order 0) */
    uint8_t  searched[ME_NL][ME_NR];
    uint8_t  pend00[ME_NL][ME_NR]; /* check_00_center SADs in flight */
    uint8_t  pending; /* MePending: which task results the next stage has to consume */
    uint8_t  tf_exit; /* ME_MCTF: hme_sad < tf_me_exit_th, the integer search is skipped (ME:3109-3113) */
    uint64_t ops; /* algorithmic byte-SAD count, SURVEY 8(d) */
    uint64_t ops_fs; /* part of ops spent in the integer full search */
    uint64_t ops_fs_inline; /* part of ops_fs searched inside k_hme (probe, areas of <= 256 positions) rather than as k_fullsearch tiles */
} B64State;

/* one svt_sad_loop_kernel call. Coordinates are in the plane's own interior coordinate system (may be
 * negative / beyond the interior: the padding is real memory). */
typedef struct SearchTask {
    uint8_t  valid;
    uint8_t  level; /* 0 full, 1 quarter, 2 sixteenth */
    uint8_t  bw; /* block width in pixels */
    uint8_t  rows; /* block rows that are summed (already halved for sub-sampled SAD) */
    uint8_t  row_step; /* 2 for sub-sampled SAD (src and ref strides doubled), else 1 */
    uint8_t  skip_line; /* skip even search rows (only honoured for 16-wide blocks of <= 16 rows) */
    uint16_t ref_slot;
    int16_t  sx, sy; /* source block origin */
    int16_t  rx, ry; /* reference position of search point (0, 0) */
    int16_t  sa_w, sa_h;
    int16_t  x0, y0; /* search-area origin relative to the block (kept for the consuming stage) */
} SearchTask;

typedef struct SearchResult {
    uint32_t sad; /* ME_SAD_LOOP_INIT when no position was evaluated / beat it */
    int16_t  x, y;
    uint8_t  found;
    uint8_t  pad_[3];
} SearchResult;

/* one open_loop_me_fullpel_search_sblock call on the full-resolution planes */
typedef struct FsTask {
    uint8_t  valid;
    uint8_t  probe; /* single centre position of the 8x8-variance probe: scan order 0 */
    uint16_t ref_slot;
    int16_t  x0, y0; /* search-area origin relative to the b64 origin */
    int16_t  w, h;
} FsTask;

typedef struct FsTile { /* a rectangle of search positions of one FsTask, unit of work of the search kernel */
    uint32_t task; /* b64 * ME_NREF + li * ME_NR + ri */
    uint16_t job;
    uint16_t tx, ty; /* tile origin inside the search area */
    uint16_t ref_slot;
    int16_t  x0, y0, w, h; /* the task's search area (copied: the FsTask slot is reused by the next round) */
    uint8_t  probe, kx /* positions per thread: the tile is 4 * kx positions wide */, pad_[2];
} FsTile;

/* one picture's ME job; an array of these (one per picture of a batch) lives in HBM */
typedef struct MeJob {
    MeGpuParams p;
    MePicGeom   g;
    uint16_t    src_slot;
    uint16_t    ref_slot[ME_NL][ME_NR];
    uint64_t    ref_picture_number[ME_NL][ME_NR];
    uint32_t    pu_count; /* exported PUs per b64: 85 / 21 / 5 */
    /* device buffers of this job */
    B64State           *state; /* [b64_count] */
    SearchTask         *tasks; /* [b64_count][ME_SLOTS] */
    SearchResult       *results; /* [b64_count][ME_SLOTS] */
    FsTask             *fs_tasks; /* [b64_count][ME_NREF] */
    unsigned long long *keys; /* [b64_count][ME_NREF][ME_NPU] */
    /* outputs (device mirrors of MeGpuOutputs) */
    uint8_t  *total_me_candidate_index;
    uint8_t  *me_candidate_array;
    uint32_t *me_mv_array;
    uint32_t *dist; /* [6][b64_count]: 64x64, 32x32, 16x16, 8x8 distortion, 8x8 cost variance, rc_me_distortion */
    uint8_t  *flags; /* [2][b64_count]: stationary_block_present_sb, rc_me_allow_gm */
    uint32_t *dbg_best_sad; /* optional taps, may be NULL */
    uint32_t *dbg_best_mv;
    int16_t  *dbg_centre;
    uint64_t *dbg_hme_sad;
    uint8_t  *dbg_do_ref;
    int16_t  *dbg_area;
    unsigned long long *ops_total; /* [5] per job: all byte-SADs, full-search byte-SADs, tf_tot_horz_blks, tf_tot_vert_blks, full-search byte-SADs done inside k_hme */
    /* ME_MCTF outputs (device mirrors of MeGpuOutputs.tf_*) */
    uint32_t *tf_best_sad; /* [b64_count][ME_NPU] */
    uint32_t *tf_best_mv;
    int16_t  *tf_hme_centre; /* [b64_count][2] */
    uint32_t *tf_hme_sad; /* [b64_count] */
    uint8_t  *tf_early_exit; /* [b64_count] */
} MeJob;

/* geometry of a width x height picture (both multiples of 8) in the pool */
ME_HD void me_make_geom(MePicGeom *g, int width, int height) {
    const int pads[3] = {ME_PAD_FULL, ME_PAD_QUARTER, ME_PAD_SIXTEENTH};
    g->width     = width;
    g->height    = height;
    g->b64_cols  = (width + 63) / 64;
    g->b64_rows  = (height + 63) / 64;
    g->b64_count = g->b64_cols * g->b64_rows;
    uint32_t off = 0;
    for (int l = 0; l < 3; l++) {
        MePlaneGeom *pl = &g->plane[l];
        pl->width       = width >> l;
        pl->height      = height >> l;
        pl->pad         = pads[l];
        pl->stride      = (pl->width + 2 * pl->pad + 127) & ~127;
        pl->rows        = pl->height + 2 * pl->pad;
        pl->offset      = off;
        off += (uint32_t)pl->stride * (uint32_t)pl->rows;
        off = (off + 255u) & ~255u;
    }
    g->slot_bytes = off;
}

#define ME_KEY_INIT (((unsigned long long)ME_MAX_SAD_VALUE << 32) | 0xFFFFFFFFull)

#endif
