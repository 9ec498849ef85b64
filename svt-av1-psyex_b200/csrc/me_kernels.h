/* me_kernels.h - launch wrappers of me_kernels.cu (internal to libme_gpu.so) */
#ifndef ME_KERNELS_H
#define ME_KERNELS_H

#include <cuda_runtime.h>

#include "me_types.h"

#define FS_TW 64 /* full-search tile: 64 x FS_TH search positions per CTA iteration, one warp per 32 x 8 positions */
#ifndef FS_TH
#define FS_TH 16
#endif
#define FS_THREADS (FS_TH * 8)
#define FS_CTAS_PER_SM (512 / FS_THREADS) /* 128 registers per thread */

typedef struct FsTileList {
    unsigned int count; /* tiles appended by k_ctrl */
    unsigned int next; /* next tile to hand out (k_fullsearch) */
    unsigned int cap;
    unsigned int overflow;
    FsTile      *tiles;
} FsTileList;

cudaError_t        me_kernels_setup(void);
unsigned long long me_launch_count(void);
void   me_launch_pad_full(uint8_t *pool, size_t slot_off, const MePlaneGeom *pl, cudaStream_t st);
void   me_launch_repitch(const uint8_t *src, int src_pitch, uint8_t *dst, int dst_pitch, int row_bytes, int rows, cudaStream_t st);
void   me_launch_downsample(uint8_t *pool, size_t slot_off, const MePlaneGeom *src, const MePlaneGeom *dst, cudaStream_t st);
void   me_launch_ctrl(MeJob *jobs, int njobs, int max_b64, uint32_t stage_mask, FsTileList *list, cudaStream_t st);
void   me_launch_search(const MeJob *jobs, int njobs, int max_b64, int slots, const uint8_t *pool, cudaStream_t st);
void   me_launch_fullsearch(MeJob *jobs, FsTileList *list, const uint8_t *pool, int ctas, cudaStream_t st);
void   me_launch_epilogue(MeJob *jobs, int njobs, int max_b64, cudaStream_t st);
double me_measure_sad_peak(cudaStream_t st);
/* me_hme.cu */
#define HM_NMAPS 6
/* TMA descriptors (CUtensorMap, 128 bytes each) of the three planes of the pool as (x, row, slot) byte tensors, one per
 * box width that k_hme stages search windows with; passed to the kernel by value (__grid_constant__) */
typedef struct HmeMaps {
    alignas(64) unsigned char m[HM_NMAPS][128];
} HmeMaps;
cudaError_t        me_hme_setup(void);
unsigned long long me_hme_launch_count(void);
/* 0 on success; -1 when the driver has no cuTensorMapEncodeTiled, else the CUresult of the failed encode */
int    me_hme_make_maps(HmeMaps *maps, const uint8_t *pool, const MePicGeom *g, int nslots);
void   me_launch_hme(MeJob *jobs, int njobs, int max_b64, FsTileList *list, const uint8_t *pool, const HmeMaps *maps,
                     cudaStream_t st);
/* me_dg.cu: dynamic-GOP detector HME level 0 of one (source, reference) pair, summed into *acc (device memory, zeroed
 * by the caller on the same stream) */
unsigned long long me_dg_launch_count(void);
void   me_launch_dg(const uint8_t *pool, const MePicGeom *g, uint32_t src_slot, uint32_t ref_slot, int sa, MeGpuDgMetrics *acc,
                    cudaStream_t st);

#endif
