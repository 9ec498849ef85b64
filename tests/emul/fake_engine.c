/*
 * fake_engine.c - TEST HARNESS only: a stand-in for libme_gpu.so with the entry points integration/svt_me_gpu_hook.c binds,
 * so that the hook's HOST LOGIC (registration keyed by PA reference object, release on re-use, the once-per-pair protocol of
 * the temporal filter among segment threads, the identities it hands to the engine) can be tested on the CPU
 * (tests/test_hook_host_logic.py). It computes nothing: results are patterns of the identities it was called with, and every
 * misuse a real engine would reject (unknown picture, double registration, job not found) is rejected here too.
 */
#include <pthread.h>
#include <stdint.h>
#include <string.h>

#include "me_gpu.h"

#define FAKE_API __attribute__((visibility("default")))
#define MAXP 256
#define MAXJ 64

static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
static uint64_t        pics[MAXP];
static int             npics, inited;
static struct { uint64_t pic, ref; uint8_t type; int used; } jobs[MAXJ];
static uint64_t        cnt[8]; /* registers, releases, submits, waits, wait_pairs, read_planes, dg calls, errors */

static int find(uint64_t id) {
    for (int i = 0; i < npics; i++)
        if (pics[i] == id)
            return i;
    return -1;
}
static int err(int code) { cnt[7]++; return code; }

FAKE_API int me_gpu_init(int n, const int *ids, int w, int h, int r0, int r1, int inflight) {
    (void)n; (void)ids; (void)w; (void)h; (void)r0; (void)r1; (void)inflight;
    pthread_mutex_lock(&mu);
    npics = 0; inited = 1;
    memset(jobs, 0, sizeof jobs);
    memset(cnt, 0, sizeof cnt);
    pthread_mutex_unlock(&mu);
    return ME_GPU_OK;
}
FAKE_API void        me_gpu_shutdown(void) { inited = 0; }
FAKE_API const char *me_gpu_last_error(void) { return "fake engine"; }

FAKE_API int me_gpu_register_picture(uint64_t id, const uint8_t *luma, int stride, int ox, int oy, int w, int h) {
    (void)stride; (void)ox; (void)oy; (void)w; (void)h;
    pthread_mutex_lock(&mu);
    int rc = ME_GPU_OK;
    if (!inited || !luma) rc = err(ME_GPU_ERR_BAD_ARG);
    else if (find(id) >= 0 || npics == MAXP) rc = err(ME_GPU_ERR_STATE);
    else { pics[npics++] = id; cnt[0]++; }
    pthread_mutex_unlock(&mu);
    return rc;
}
FAKE_API int me_gpu_release_picture(uint64_t id) {
    pthread_mutex_lock(&mu);
    const int i = find(id);
    int       rc = ME_GPU_OK;
    if (i < 0) rc = err(ME_GPU_ERR_NOT_FOUND);
    else { pics[i] = pics[--npics]; cnt[1]++; }
    pthread_mutex_unlock(&mu);
    return rc;
}
FAKE_API int me_gpu_read_plane(uint64_t id, int level, uint8_t *dst, int dst_stride) {
    (void)dst_stride;
    pthread_mutex_lock(&mu);
    int rc = ME_GPU_OK;
    if (find(id) < 0 || level < 1 || level > 2 || !dst) rc = err(ME_GPU_ERR_NOT_FOUND);
    else { dst[0] = (uint8_t)((id >> 32) * 16 + (unsigned)level); cnt[5]++; } /* a marker of (registration count, level) */
    pthread_mutex_unlock(&mu);
    return rc;
}
FAKE_API int me_gpu_submit(uint64_t id, const MeGpuParams *p, const uint64_t refs[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS]) {
    pthread_mutex_lock(&mu);
    int rc = ME_GPU_OK, slot = -1;
    if (find(id) < 0) rc = err(ME_GPU_ERR_NOT_FOUND);
    for (int l = 0; !rc && l < p->num_of_list_to_search; l++)
        for (int r = 0; !rc && r < p->num_of_ref_pic_to_search[l]; r++)
            if (find(refs[l][r]) < 0) rc = err(ME_GPU_ERR_NOT_FOUND);
    for (int i = 0; !rc && i < MAXJ; i++) {
        if (jobs[i].used && jobs[i].pic == id && jobs[i].type == p->me_type && jobs[i].ref == refs[0][0]) rc = err(ME_GPU_ERR_STATE);
        if (!jobs[i].used && slot < 0) slot = i;
    }
    if (!rc && slot < 0) rc = err(ME_GPU_ERR_NO_MEMORY);
    if (!rc) { jobs[slot].used = 1; jobs[slot].pic = id; jobs[slot].type = p->me_type; jobs[slot].ref = refs[0][0]; cnt[2]++; }
    pthread_mutex_unlock(&mu);
    return rc;
}
static int retire(uint64_t id, int exact, uint8_t type, uint64_t ref, MeGpuOutputs *out) {
    pthread_mutex_lock(&mu);
    int rc = err(ME_GPU_ERR_NOT_FOUND);
    cnt[7]--; /* provisional */
    for (int i = 0; i < MAXJ; i++)
        if (jobs[i].used && jobs[i].pic == id && (!exact || (jobs[i].type == type && jobs[i].ref == ref))) {
            jobs[i].used = 0;
            rc = ME_GPU_OK;
            if (out && out->tf_best_sad) /* ME_MCTF pattern: every value names the neighbour and the block */
                for (uint32_t b = 0; b < out->b64_count; b++) {
                    for (int k = 0; k < 85; k++) {
                        out->tf_best_sad[(size_t)b * 85 + k] = (uint32_t)(jobs[i].ref >> 32) * 100000u + b * 100u + (uint32_t)k;
                        out->tf_best_mv[(size_t)b * 85 + k]  = (uint32_t)jobs[i].ref;
                    }
                    out->tf_hme_centre[2 * b] = (int16_t)b; out->tf_hme_centre[2 * b + 1] = (int16_t)-(int)b;
                    out->tf_hme_sad[b] = b + 1; out->tf_early_exit[b] = (uint8_t)(b & 1);
                }
            if (out) { out->tf_tot_horz_blks = 5; out->tf_tot_vert_blks = 7; }
            if (out && out->total_me_candidate_index) { /* open-loop pattern: 85 PUs per b64, the index names (b64, PU) */
                out->pu_count = 85;
                for (uint32_t k = 0; k < out->b64_count * 85u; k++) out->total_me_candidate_index[k] = (uint8_t)k;
            }
            break;
        }
    if (rc != ME_GPU_OK) cnt[7]++;
    pthread_mutex_unlock(&mu);
    return rc;
}
FAKE_API int me_gpu_wait(uint64_t id, MeGpuOutputs *out) { const int rc = retire(id, 0, 0, 0, out); if (!rc) { pthread_mutex_lock(&mu); cnt[3]++; pthread_mutex_unlock(&mu); } return rc; }
FAKE_API int me_gpu_wait_pair(uint64_t id, uint8_t type, uint64_t ref, MeGpuOutputs *out) {
    const int rc = retire(id, 1, type, ref, out);
    if (!rc) { pthread_mutex_lock(&mu); cnt[4]++; pthread_mutex_unlock(&mu); }
    return rc;
}
FAKE_API int me_gpu_cancel(uint64_t id) { return retire(id, 0, 0, 0, NULL); }
FAKE_API int me_gpu_dg_detector(uint64_t a, uint64_t b, uint8_t res, MeGpuDgMetrics *m) {
    (void)res;
    pthread_mutex_lock(&mu);
    int rc = ME_GPU_OK;
    if (find(a) < 0 || find(b) < 0) rc = err(ME_GPU_ERR_NOT_FOUND);
    else { m->tot_dist = a >> 32; m->tot_cplx = (uint32_t)(b >> 32); m->tot_active = (uint32_t)(a & 0xffff); m->sum_in_vectors = (int32_t)(b & 0xffff); cnt[6]++; }
    pthread_mutex_unlock(&mu);
    return rc;
}
/* registers, releases, submits, waits, wait_pairs, read_planes, dg calls, rejected calls; returns the pictures still held */
FAKE_API int fake_engine_counts(uint64_t out[8]) {
    pthread_mutex_lock(&mu);
    memcpy(out, cnt, sizeof cnt);
    const int n = npics;
    pthread_mutex_unlock(&mu);
    return n;
}
