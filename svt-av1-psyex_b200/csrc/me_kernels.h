/* me_kernels.h - launch wrappers of me_kernels.cu (internal to libme_gpu.so) */
#ifndef ME_KERNELS_H
#define ME_KERNELS_H

#include <cuda_runtime.h>

#include "me_types.h"

/* full-search tile: (4 * KX) x FS_TH search positions per CTA iteration; warp w takes search rows [8w, 8w + 8): lane =
 * (byte shift s, row), KX positions x = s + 4k per thread */
#define FS_TW 32 /* widest tile (KX = 8); also the row stride of the packed tile-position index */
#ifndef FS_TH
#define FS_TH 32
#endif
#define FS_THREADS (FS_TH * 4)
#define FS_CTAS_PER_SM 3 /* 73 KB of shared memory each (TMA double buffer + the four shifted copies): 170 registers per thread */

/* positions per thread (x = s + 4k, k < KX) of the tiles of a search area `w` positions wide: a warp covers 4 * KX
 * columns; the KX in 5..8 that wastes the fewest columns wins (104 -> 7: 4 x 28 = 112 instead of 4 x 32 = 128) */
__host__ __device__ static inline int fs_choose_kx(int w) {
    int best = 8, cost = ((w + 31) / 32) * 32;
    for (int k = 7; k >= 5; k--) {
        const int c = ((w + 4 * k - 1) / (4 * k)) * 4 * k;
        if (c < cost) {
            cost = c;
            best = k;
        }
    }
    return best;
}

typedef struct FsTileList {
    unsigned int count; /* tiles appended by k_hme */
    unsigned int next; /* next tile to hand out (k_fullsearch) */
    unsigned int cap;
    unsigned int overflow;
    FsTile      *tiles;
} FsTileList;

#ifdef __CUDACC__
/* append the tiles (4 * KX x FS_TH positions) of a large search area to the work list of k_fullsearch */
__device__ static inline void fs_emit_tiles(const FsTask &f, FsTileList *list, uint32_t task, uint16_t job) {
    const uint32_t kx = (uint32_t)fs_choose_kx(f.w), tw = 4 * kx;
    const uint32_t ntx = ((uint32_t)f.w + tw - 1) / tw, nty = (uint32_t)(f.h + FS_TH - 1) / FS_TH;
    const uint32_t n = ntx * nty, base = atomicAdd(&list->count, n);
    if (base + n > list->cap) {
        atomicExch(&list->overflow, 1u);
        return;
    }
    FsTile *t = list->tiles + base;
    for (uint32_t ty = 0; ty < nty; ty++)
        for (uint32_t tx = 0; tx < ntx; tx++, t++) {
            t->task = task;
            t->job  = job;
            t->tx   = (uint16_t)(tx * tw);
            t->ty   = (uint16_t)(ty * FS_TH);
            t->ref_slot = f.ref_slot;
            t->x0 = f.x0; t->y0 = f.y0; t->w = f.w; t->h = f.h;
            t->probe = f.probe;
            t->kx    = (uint8_t)kx;
        }
}
#endif

cudaError_t        me_kernels_setup(void);
unsigned long long me_launch_count(void);
void   me_launch_pad_full(uint8_t *pool, size_t slot_off, const MePlaneGeom *pl, cudaStream_t st);
void   me_launch_repitch(const uint8_t *src, int src_pitch, uint8_t *dst, int dst_pitch, int row_bytes, int rows, cudaStream_t st);
void   me_launch_downsample(uint8_t *pool, size_t slot_off, const MePlaneGeom *src, const MePlaneGeom *dst, cudaStream_t st);
void   me_launch_epilogue(MeJob *jobs, int njobs, int max_b64, cudaStream_t st);
double me_measure_sad_peak(cudaStream_t st);
/* me_hme.cu */
#define HM_NMAPS 7
#define FS_MAP 6 /* tensor map of a k_fullsearch tile's window: full plane, 128 bytes x 96 rows */
/* TMA descriptors (CUtensorMap, 128 bytes each) of the three planes of the pool as (x, row, slot) byte tensors, one per
 * box width that k_hme stages search windows with; passed to the kernel by value (__grid_constant__) */
typedef struct HmeMaps {
    alignas(64) unsigned char m[HM_NMAPS][128];
} HmeMaps;
void   me_launch_fullsearch(MeJob *jobs, FsTileList *list, const uint8_t *pool, const HmeMaps *maps, int sms, cudaStream_t st);
cudaError_t        me_hme_setup(void);
unsigned long long me_hme_launch_count(void);
/* 0 on success; -1 when the driver has no cuTensorMapEncodeTiled, else the CUresult of the failed encode */
int    me_hme_make_maps(HmeMaps *maps, const uint8_t *pool, const MePicGeom *g, int nslots);
/* phase: 0 = the whole per-b64 chain; 1 / 2 = its two halves around the tiled searches of reference (0, 0) (me_hme.cu) */
void   me_launch_hme(MeJob *jobs, int njobs, int max_b64, int phase, FsTileList *list, const uint8_t *pool, const HmeMaps *maps,
                     cudaStream_t st);
/* me_dg.cu: dynamic-GOP detector HME level 0 of one (source, reference) pair, summed into *acc (device memory, zeroed
 * by the caller on the same stream) */
unsigned long long me_dg_launch_count(void);
void   me_launch_dg(const uint8_t *pool, const MePicGeom *g, uint32_t src_slot, uint32_t ref_slot, int sa, MeGpuDgMetrics *acc,
                    cudaStream_t st);

#endif
