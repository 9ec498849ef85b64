/*
 * me_engine.cu - host runtime behind the C-ABI of include/me_gpu.h.
 *
 * Owns, per device: the picture pool (padded full / quarter / sixteenth planes of every registered
 * picture, resident in HBM from register to release), a ring of job contexts (one stream + the per-picture
 * state, task, key and output buffers) and an upload stream. A job is the whole open-loop ME of one
 * picture: ~15 kernel launches on the job's stream, no host round trip until me_gpu_wait copies the
 * MeSbResults images back. Pictures of different jobs overlap on the device through their streams.
 *
 * With several devices in one process every registered pyramid is built once on its owner device and
 * replicated to the others with cudaMemcpyPeerAsync (NVLink); jobs are dealt round-robin by picture.
 * There is no collective on the search and no CPU fallback: without a usable device every call fails.
 */
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <condition_variable>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "me_ctrl.h"
#include "me_kernels.h"

namespace {

constexpr int kMaxJobs = 16;

struct Slot {
    bool        used = false;
    uint64_t    picture_number = 0;
    int         width = 0, height = 0;
    int         refcount = 0; /* jobs in flight that read this slot */
    bool        release_pending = false;
    bool        ready_seen[8] = {}; /* the host has observed ready[dev] complete: no stream wait needed */
    cudaEvent_t ready[8] = {}; /* per device: pyramid complete on that device */
    cudaEvent_t ipc_ready = nullptr; /* interprocess twin of ready[0], created when the picture is first exported to another process
                                        (a stream of another device of THIS process must not wait on an interprocess event:
                                        the driver faults) */
};

/* uploaded in one copy at submit; the tail (list header + counters) comes back in one copy */
struct JobTail {
    FsTileList         list;
    unsigned long long ops[5]; /* byte-SADs, full-search byte-SADs, tf_tot_horz_blks, tf_tot_vert_blks, full-search byte-SADs inside k_hme */
};
struct JobHead {
    MeJob   job;
    JobTail tail;
};

/* product outputs of one picture in one device slab WITHOUT gaps: the 32-bit arrays first, then the byte
 * arrays. off[] / bytes[] are indexed in MeGpuOutputs field order (total, cands, mvs, six distortion arrays,
 * two flag arrays); order[] lists the fields by increasing offset. */
struct OutLayout {
    size_t off[11], bytes[11], total;
    int    order[11];
};
OutLayout output_layout(size_t nb, size_t P, size_t max_cand, size_t max_refs) {
    OutLayout    L;
    const size_t b[11] = {nb * P, nb * P * max_cand, nb * P * max_refs * 4, nb * 4, nb * 4, nb * 4, nb * 4, nb * 4, nb * 4, nb, nb};
    const int    ord[11] = {2, 3, 4, 5, 6, 7, 8, 0, 1, 9, 10};
    size_t       o = 0;
    for (int k = 0; k < 11; k++) {
        const int i = ord[k];
        L.order[k] = i;
        L.off[i]   = o;
        L.bytes[i] = b[i];
        o += b[i];
    }
    L.total = o;
    return L;
}

/* ME_MCTF jobs use the same slab for the five tf_* arrays of MeGpuOutputs, in struct order */
struct TfLayout {
    size_t off[5], bytes[5], total;
};
TfLayout tf_layout(size_t nb) {
    TfLayout     L;
    const size_t b[5] = {nb * ME_NPU * 4, nb * ME_NPU * 4, nb * 2 * sizeof(int16_t), nb * 4, nb};
    size_t       o    = 0;
    for (int i = 0; i < 5; i++) {
        L.off[i]   = o;
        L.bytes[i] = b[i];
        o += b[i];
    }
    L.total = o;
    return L;
}

struct JobCtx {
    bool         busy = false;
    bool         retiring = false; /* a waiter has taken this job and dropped the lock to wait for the device */
    uint64_t     picture_number = 0;
    uint64_t     ticket = 0; /* submission order */
    uint8_t      me_type = 0; /* (picture, me_type, first reference) identifies a job: a picture's open-loop ME and its */
    uint64_t     ref00 = 0; /*  temporal-filter pairs can be in flight together */
    int          device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t  ev[5] = {}; /* start, hme done, search done, end, (spare) */
    cudaEvent_t  pev[64] = {}; /* profiling mode: (begin, end) pairs around every kernel launch */
    cudaStream_t tail = nullptr; /* k_fullsearch, k_epilogue and the result copies: `stream`, or a stream of higher priority */
    cudaEvent_t  ev_mid = nullptr; /* k_hme done (stream -> tail) */
    int          pkind[32] = {}; /* kernel class of each pair: 0 ctrl, 1 search, 2 fullsearch probe round, 3 fullsearch main, 4 epilogue */
    int          npairs = 0;
    bool         prof = false;
    JobHead     *h_head = nullptr; /* pinned: what is uploaded at submit */
    JobHead     *d_head = nullptr;
    MeJob       *h_job = nullptr, *d_job = nullptr; /* = &head->job */
    FsTileList  *d_list = nullptr; /* = &d_head->list */
    JobTail     *h_tail = nullptr; /* pinned: tile-list header + byte-SAD counters read back after the job */
    uint8_t     *out_slab = nullptr; /* device: total | cands | mvs | dist[6] | flags[2], see output_layout() */
    size_t       out_cap = 0;
    bool         d2h_enqueued = false; /* outputs were given at submit: the copies are already on the stream */
    std::vector<int> slots_used;
    /* device buffers */
    B64State *state = nullptr; SearchTask *tasks = nullptr; SearchResult *results = nullptr; FsTask *fs_tasks = nullptr;
    unsigned long long *keys = nullptr; FsTile *tiles = nullptr;
    uint32_t *dbg_sad = nullptr, *dbg_mv = nullptr; int16_t *dbg_centre = nullptr; uint64_t *dbg_hme = nullptr;
    uint8_t *dbg_do_ref = nullptr; int16_t *dbg_area = nullptr;
    size_t cap_cands = 0, cap_mvs = 0;
    uint32_t tile_cap = 0;
};

struct Device {
    int          id = 0;
    uint8_t     *pool = nullptr;
    cudaStream_t upload = nullptr;
    int          sms = 148;
    uint8_t     *stage = nullptr; /* landing buffer of contiguous host->device copies (one picture, reused in stream order) */
    size_t       stage_bytes = 0;
    cudaStream_t    dg_stream = nullptr; /* dynamic-GOP detector: its own stream, a device accumulator, a pinned landing struct */
    MeGpuDgMetrics *dg_acc = nullptr, *dg_host = nullptr;
    std::vector<JobCtx> jobs;
    struct MapSet { int width, height; HmeMaps maps; };
    std::vector<MapSet> mapsets; /* TMA descriptors of the pool per picture size seen (normally one) */
};

struct Engine {
    bool                initialised = false;
    std::vector<Device> dev;
    MePicGeom           gmax{};
    int                 nslots = 0;
    std::vector<Slot>   slots;
    std::map<uint64_t, int> by_number;
    int                 max_l0 = 0, max_l1 = 0;
    uint64_t            submit_counter = 0, register_counter = 0, dg_counter = 0;
    std::mutex          dg_mu; /* one detector call at a time (they share the per-device accumulator) */
    MeGpuTimings        last{};
    std::mutex          mu;
    std::condition_variable cv; /* a job was retired */
    bool                profiling = false;
    uint64_t            ticket = 0; /* submissions so far: orders the jobs of one picture */
    struct Peer { bool open = false; uint8_t *base = nullptr; uint64_t slot_bytes = 0; std::map<int, cudaEvent_t> ev; };
    Peer                peers[16]; /* picture pools of other processes (one per GPU) mapped through CUDA IPC */
    uint64_t            peer_bytes = 0; /* bytes pulled from peers so far (NVLink) */
    uint64_t            done_ring[64] = {}; /* picture numbers of the last retired jobs (me_gpu_wait_done after the retirement) */
    bool                done_valid[64] = {};
    unsigned            done_next = 0;
};

/* The engine the me_gpu_* calls of the calling thread go to: the process-wide default engine unless the thread has bound
 * another one (me_gpu_engine_bind): several encoder instances in one process each get their own pool, jobs and streams. */
static Engine               g_default;
static thread_local Engine *t_eng = &g_default;
static inline Engine &E() { return *t_eng; }
thread_local std::string t_err; /* every calling thread reads back its own last error */

int fail(int code, const std::string &msg) {
    t_err = msg;
    return code;
}
#define CU(x)                                                                                              \
    do {                                                                                                   \
        cudaError_t e_ = (x);                                                                              \
        if (e_ != cudaSuccess)                                                                             \
            return fail(ME_GPU_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_));                \
    } while (0)

template <typename T>
cudaError_t dalloc(T **p, size_t n) {
    return cudaMalloc((void **)p, n * sizeof(T));
}

uint64_t max_tiles_per_b64(const MeGpuParams *p) {
    /* upper bound of the tiles one b64 can emit: every searched reference at sa_max (x1.5 for high 8x8 variance),
       the references the MV-based adjustment applies to (reference 0 of each list when nearest_ref_only) x multiplier */
    auto tiles = [&](uint32_t mult) -> uint64_t {
        uint32_t w = p->me_sa.sa_max.width * mult, h = p->me_sa.sa_max.height * mult;
        if (p->me_sr_mult2_th != 0xFFFFFFFFu) {
            w = w * 3 / 2;
            h = h * 3 / 2;
        }
        w = (w + 7) & ~7u; /* ME:1314 */
        h = h < 3 ? 3 : h;
        const uint32_t tw = 4u * (uint32_t)fs_choose_kx((int)w);
        return (uint64_t)((w + tw - 1) / tw) * ((h + FS_TH - 1) / FS_TH);
    };
    uint32_t nref = 0, nmult = 0;
    for (uint32_t li = 0; li < p->num_of_list_to_search && li < ME_NL; li++) {
        const uint32_t n = p->num_of_ref_pic_to_search[li] < ME_NR ? p->num_of_ref_pic_to_search[li] : ME_NR;
        nref += n;
        if (p->mv_sa_adj_enabled)
            nmult += p->mv_sa_adj_nearest_ref_only ? (n ? 1u : 0u) : n;
    }
    return (nref - nmult) * tiles(1) + nmult * tiles(p->mv_sa_adj_enabled ? p->mv_sa_adj_sa_multiplier : 1);
}

int alloc_job(Device &d, JobCtx &jc, const MePicGeom &gm, int max_l0, int max_l1) {
    const size_t nb = (size_t)gm.b64_count;
    const size_t max_refs = (size_t)max_l0 + max_l1;
    const size_t max_cand = (size_t)max_l0 + max_l1 + (size_t)max_l0 * max_l1 + (max_l0 ? max_l0 - 1 : 0) + (max_l1 == 3 ? 1 : 0);
    jc.device    = d.id;
    jc.cap_cands = nb * ME_NPU * (max_cand ? max_cand : 1);
    jc.cap_mvs   = nb * ME_NPU * (max_refs ? max_refs : 1);
    jc.tile_cap  = (uint32_t)(nb * ME_NREF * 128); /* >= max_tiles_per_b64 of every preset (832 at preset 0) */
    CU(cudaStreamCreateWithFlags(&jc.stream, cudaStreamNonBlocking));
    {
        /* ME_GPU_DEBUG builds, ME_GPU_SPLIT=1 (experiment): the tail of a picture (tiled full search, epilogue, result copies)
           on a stream of higher priority than the k_hme launches of the pictures behind it. Measured: within +-3 % on the
           1080p workloads, -6 % end to end at 4K preset 2 */
#ifdef ME_GPU_DEBUG
        static const int split = getenv("ME_GPU_SPLIT") ? atoi(getenv("ME_GPU_SPLIT")) : 0;
#else
        const int split = 0;
#endif
        int lo = 0, hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        if (split && hi < lo) {
            CU(cudaStreamCreateWithPriority(&jc.tail, cudaStreamNonBlocking, hi));
            CU(cudaEventCreateWithFlags(&jc.ev_mid, cudaEventDisableTiming));
        } else
            jc.tail = jc.stream;
    }
    for (auto &e : jc.ev) CU(cudaEventCreate(&e));
    for (auto &e : jc.pev) CU(cudaEventCreate(&e));
    CU(cudaMallocHost((void **)&jc.h_head, sizeof(JobHead)));
    CU(cudaMallocHost((void **)&jc.h_tail, sizeof(JobTail)));
    CU(dalloc(&jc.d_head, 1));
    jc.h_job  = &jc.h_head->job;
    jc.d_job  = &jc.d_head->job;
    jc.d_list = &jc.d_head->tail.list;
    jc.out_cap = output_layout(nb, ME_NPU, max_cand ? max_cand : 1, max_refs ? max_refs : 1).total;
    if (jc.out_cap < tf_layout(nb).total)
        jc.out_cap = tf_layout(nb).total;
    CU(cudaMalloc((void **)&jc.out_slab, jc.out_cap));
    CU(dalloc(&jc.state, nb));
    CU(dalloc(&jc.tasks, nb * ME_SLOTS));
    CU(dalloc(&jc.results, nb * ME_SLOTS));
    CU(dalloc(&jc.fs_tasks, nb * ME_NREF));
    CU(dalloc(&jc.keys, nb * ME_NREF * ME_NPU));
    CU(dalloc(&jc.tiles, (size_t)jc.tile_cap));
    CU(dalloc(&jc.dbg_sad, nb * ME_NREF * ME_NPU));
    CU(dalloc(&jc.dbg_mv, nb * ME_NREF * ME_NPU));
    CU(dalloc(&jc.dbg_centre, nb * ME_NREF * 2));
    CU(dalloc(&jc.dbg_hme, nb * ME_NREF));
    CU(dalloc(&jc.dbg_do_ref, nb * ME_NREF));
    CU(dalloc(&jc.dbg_area, nb * ME_NREF * 4));
    CU(cudaMemset(jc.tasks, 0, nb * ME_SLOTS * sizeof(SearchTask)));
    CU(cudaMemset(jc.results, 0, nb * ME_SLOTS * sizeof(SearchResult)));
    return ME_GPU_OK;
}

void free_job(JobCtx &jc) {
    if (jc.tail && jc.tail != jc.stream) cudaStreamDestroy(jc.tail);
    if (jc.ev_mid) cudaEventDestroy(jc.ev_mid);
    if (jc.stream) cudaStreamDestroy(jc.stream);
    for (auto &e : jc.ev)
        if (e) cudaEventDestroy(e);
    for (auto &e : jc.pev)
        if (e) cudaEventDestroy(e);
    cudaFreeHost(jc.h_head); cudaFreeHost(jc.h_tail);
    void *ptrs[] = {jc.d_head, jc.state, jc.tasks, jc.results, jc.fs_tasks, jc.keys, jc.tiles, jc.out_slab, jc.dbg_sad, jc.dbg_mv,
                    jc.dbg_centre, jc.dbg_hme, jc.dbg_do_ref, jc.dbg_area};
    for (void *p : ptrs) cudaFree(p);
    jc = JobCtx();
}

void free_slot_if_idle(int s) {
    Slot &sl = E().slots[s];
    if (sl.used && sl.release_pending && sl.refcount == 0) {
        E().by_number.erase(sl.picture_number);
        sl.used = false;
        sl.release_pending = false;
    }
}

/* builds the pyramid of slot s on its owner device and replicates the slot to the other devices */
int build_and_replicate(int s, int owner, const MePicGeom &gm, bool pad_full) {
    Device      &d   = E().dev[owner];
    const size_t off = (size_t)s * E().gmax.slot_bytes;
    CU(cudaSetDevice(d.id));
    if (pad_full)
        me_launch_pad_full(d.pool, off, &gm.plane[0], d.upload);
    me_launch_downsample(d.pool, off, &gm.plane[0], &gm.plane[1], d.upload);
    me_launch_downsample(d.pool, off, &gm.plane[1], &gm.plane[2], d.upload);
    CU(cudaGetLastError());
    CU(cudaEventRecord(E().slots[s].ready[owner], d.upload));
    for (size_t o = 0; o < E().dev.size(); o++) {
        if ((int)o == owner)
            continue;
        Device &p = E().dev[o];
        /* the copy runs on the destination's upload stream, after the owner finished the pyramid */
        CU(cudaSetDevice(p.id));
        CU(cudaStreamWaitEvent(p.upload, E().slots[s].ready[owner], 0));
        CU(cudaMemcpyPeerAsync(p.pool + off, p.id, d.pool + off, d.id, E().gmax.slot_bytes, p.upload));
        CU(cudaEventRecord(E().slots[s].ready[o], p.upload));
    }
    return ME_GPU_OK;
}

int register_common(uint64_t picture_number, const uint8_t *src, int stride, int org_x, int org_y, int width, int height,
                    bool padded) {
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "me_gpu_init has not been called");
    if (!src || width <= 0 || height <= 0 || (width & 7) || (height & 7) || width > E().gmax.width || height > E().gmax.height)
        return fail(ME_GPU_ERR_BAD_ARG, "bad picture geometry (luma size must be a multiple of 8 and within the init maximum)");
    if (padded && (org_x < ME_PAD_FULL || org_y < ME_PAD_FULL || stride < width + org_x + ME_PAD_FULL))
        return fail(ME_GPU_ERR_BAD_ARG, "padded plane needs at least 68 pixels of border on every side");
    if (E().by_number.count(picture_number))
        return fail(ME_GPU_ERR_STATE, "picture already registered");
    int s = -1;
    for (int i = 0; i < E().nslots; i++)
        if (!E().slots[i].used) {
            s = i;
            break;
        }
    if (s < 0)
        return fail(ME_GPU_ERR_NO_MEMORY, "picture pool is full (raise pictures_in_flight or release pictures)");
    MePicGeom gm;
    me_make_geom(&gm, width, height);
    const int    owner = (int)(E().register_counter++ % E().dev.size());
    Device      &d     = E().dev[owner];
    const size_t off   = (size_t)s * E().gmax.slot_bytes + gm.plane[0].offset;
    CU(cudaSetDevice(d.id));
    /* rows of the source to their place in the pool. When the source rows are (nearly) contiguous and 4-byte granular, one
       contiguous copy into the landing buffer plus a re-pitch kernel: a pitched copy of 2 KB rows reaches a third of the
       PCIe rate. Otherwise (wide strides, odd alignment) the pitched copy. */
    const uint8_t *first     = padded ? src + (size_t)(org_y - ME_PAD_FULL) * stride + (org_x - ME_PAD_FULL) : src;
    const int      row_bytes = padded ? width + 2 * ME_PAD_FULL : width, rows = padded ? height + 2 * ME_PAD_FULL : height;
    uint8_t       *dst       = d.pool + off + (padded ? 0 : (size_t)ME_PAD_FULL * gm.plane[0].stride + ME_PAD_FULL);
    const size_t   span      = (size_t)(rows - 1) * (size_t)stride + (size_t)row_bytes;
    if (stride > 0 && !(stride & 3) && !(row_bytes & 3) && !((uintptr_t)first & 3) && span <= d.stage_bytes &&
        (size_t)stride * 4 <= (size_t)row_bytes * 5) {
        CU(cudaMemcpyAsync(d.stage, first, span, cudaMemcpyHostToDevice, d.upload));
        me_launch_repitch(d.stage, stride, dst, gm.plane[0].stride, row_bytes, rows, d.upload);
    } else
        CU(cudaMemcpy2DAsync(dst, gm.plane[0].stride, first, stride, row_bytes, rows, cudaMemcpyHostToDevice, d.upload));
    Slot &sl          = E().slots[s];
    sl.used           = true;
    sl.picture_number = picture_number;
    sl.width          = width;
    sl.height         = height;
    sl.refcount       = 0;
    sl.release_pending = false;
    for (auto &v : sl.ready_seen) v = false;
    E().by_number[picture_number] = s;
    return build_and_replicate(s, owner, gm, !padded);
}

} // namespace

extern "C" {

const char *me_gpu_last_error(void) { return t_err.c_str(); }

int me_gpu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
        return 0;
    return n;
}

void me_gpu_shutdown(void) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!E().initialised)
        return;
    for (auto &d : E().dev) {
        cudaSetDevice(d.id);
        cudaDeviceSynchronize();
        for (auto &jc : d.jobs) free_job(jc);
        if (d.upload) cudaStreamDestroy(d.upload);
        if (d.dg_stream) cudaStreamDestroy(d.dg_stream);
        cudaFree(d.dg_acc);
        cudaFreeHost(d.dg_host);
        cudaFree(d.pool);
        cudaFree(d.stage);
    }
    for (auto &pr : E().peers) {
        for (auto &kv : pr.ev) cudaEventDestroy(kv.second);
        pr.ev.clear();
        if (pr.open) cudaIpcCloseMemHandle(pr.base);
        pr = Engine::Peer();
    }
    for (auto &s : E().slots) {
        for (auto &e : s.ready)
            if (e) cudaEventDestroy(e);
        if (s.ipc_ready) cudaEventDestroy(s.ipc_ready);
    }
    E().dev.clear();
    E().slots.clear();
    E().by_number.clear();
    E().initialised = false;
    E().cv.notify_all();
}

int me_gpu_init(int device_count, const int *device_ids, int max_width, int max_height, int max_refs_l0, int max_refs_l1,
                int pictures_in_flight) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (E().initialised)
        return fail(ME_GPU_ERR_STATE, "already initialised");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(ME_GPU_ERR_NO_DEVICE, "no CUDA device (this engine has no CPU fallback)");
    if (max_width <= 0 || max_height <= 0 || max_refs_l0 < 1 || max_refs_l0 > ME_NR || max_refs_l1 < 0 || max_refs_l1 > ME_NR ||
        pictures_in_flight < 2 || device_count > 8)
        return fail(ME_GPU_ERR_BAD_ARG, "bad init arguments");
    std::vector<int> ids;
    if (device_count <= 0) {
        int cur = 0;
        cudaGetDevice(&cur);
        ids.push_back(cur);
    } else
        for (int i = 0; i < device_count; i++) ids.push_back(device_ids ? device_ids[i] : i);
    for (int id : ids)
        if (id < 0 || id >= ndev)
            return fail(ME_GPU_ERR_BAD_ARG, "device id out of range");
    me_make_geom(&E().gmax, (max_width + 7) & ~7, (max_height + 7) & ~7);
    E().nslots = pictures_in_flight;
    E().max_l0 = max_refs_l0;
    E().max_l1 = max_refs_l1;
    E().slots.assign((size_t)E().nslots, Slot());
    E().dev.resize(ids.size());
    const int njobs = pictures_in_flight < kMaxJobs ? pictures_in_flight : kMaxJobs;
    for (size_t i = 0; i < ids.size(); i++) {
        Device &d = E().dev[i];
        d.id      = ids[i];
        CU(cudaSetDevice(d.id));
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, d.id));
        if (prop.major < 10)
            return fail(ME_GPU_ERR_UNSUPPORTED, "sm_100a device required");
        d.sms = prop.multiProcessorCount;
        CU(me_kernels_setup());
        CU(me_hme_setup());
        CU(cudaStreamCreateWithFlags(&d.upload, cudaStreamNonBlocking));
        CU(cudaMalloc((void **)&d.pool, (size_t)E().nslots * E().gmax.slot_bytes + 65536));
        CU(cudaMemset(d.pool, 0, (size_t)E().nslots * E().gmax.slot_bytes + 65536));
        d.stage_bytes = (size_t)(E().gmax.width + 2 * ME_PAD_FULL) * (size_t)(E().gmax.height + 2 * ME_PAD_FULL) * 5 / 4;
        CU(cudaMalloc((void **)&d.stage, d.stage_bytes));
        CU(cudaStreamCreateWithFlags(&d.dg_stream, cudaStreamNonBlocking));
        CU(cudaMalloc((void **)&d.dg_acc, sizeof(MeGpuDgMetrics)));
        CU(cudaMallocHost((void **)&d.dg_host, sizeof(MeGpuDgMetrics)));
        d.jobs.resize((size_t)njobs);
        for (auto &jc : d.jobs) {
            int rc = alloc_job(d, jc, E().gmax, max_refs_l0, max_refs_l1);
            if (rc)
                return rc;
        }
        for (size_t o = 0; o < ids.size(); o++)
            if (o != i) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, d.id, ids[o]);
                if (can) {
                    cudaError_t e = cudaDeviceEnablePeerAccess(ids[o], 0);
                    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                        return fail(ME_GPU_ERR_CUDA, "cudaDeviceEnablePeerAccess failed");
                    cudaGetLastError();
                }
            }
        CU(cudaDeviceSynchronize());
    }
    for (auto &s : E().slots)
        for (size_t i = 0; i < ids.size(); i++) {
            CU(cudaSetDevice(ids[i]));
            CU(cudaEventCreateWithFlags(&s.ready[i], cudaEventDisableTiming));
        }
    E().initialised = true;
    return ME_GPU_OK;
}

int me_gpu_plane_geometry(int width, int height, int level, int *plane_w, int *plane_h, int *pad, int *stride, int *rows) {
    if (level < 0 || level > 2 || width <= 0 || height <= 0 || (width & 7) || (height & 7))
        return ME_GPU_ERR_BAD_ARG;
    /* geometry of the planes as me_gpu_read_pyramid returns them: reference strides (width + 2 * pad) */
    static const int pads[3] = {ME_PAD_FULL, ME_PAD_QUARTER, ME_PAD_SIXTEENTH};
    if (plane_w) *plane_w = width >> level;
    if (plane_h) *plane_h = height >> level;
    if (pad) *pad = pads[level];
    if (stride) *stride = (width >> level) + 2 * pads[level];
    if (rows) *rows = (height >> level) + 2 * pads[level];
    return ME_GPU_OK;
}

int me_gpu_register_picture(uint64_t picture_number, const uint8_t *y8b_padded, int stride, int org_x, int org_y, int width,
                            int height) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!y8b_padded)
        return fail(ME_GPU_ERR_BAD_ARG, "null plane");
    /* y8b_padded points at the first byte of the padded buffer: the interior starts at (org_x, org_y) */
    return register_common(picture_number, y8b_padded, stride, org_x, org_y, width, height, true);
}

int me_gpu_register_picture_unpadded(uint64_t picture_number, const uint8_t *y8b, int stride, int width, int height) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (stride < width)
        return fail(ME_GPU_ERR_BAD_ARG, "stride < width");
    return register_common(picture_number, y8b, stride, 0, 0, width, height, false);
}

int me_gpu_release_picture(uint64_t picture_number) {
    std::unique_lock<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    auto it = E().by_number.find(picture_number);
    if (it == E().by_number.end())
        return fail(ME_GPU_ERR_NOT_FOUND, "picture not registered");
    const int s = it->second;
    /* uploads / peer copies into this slot must finish before the slot can be handed out again; the wait happens outside
       the lock (other ME threads keep submitting and waiting meanwhile). The slot cannot change hands during it: it is
       still registered, and registering the same number again fails until it is released. */
    cudaEvent_t ev[8];
    int         ids[8];
    size_t      n = 0;
    for (size_t i = 0; i < E().dev.size(); i++)
        if (!E().slots[s].ready_seen[i]) {
            ev[n]    = E().slots[s].ready[i];
            ids[n++] = E().dev[i].id;
        }
    lk.unlock();
    for (size_t i = 0; i < n; i++) {
        cudaSetDevice(ids[i]);
        cudaEventSynchronize(ev[i]);
    }
    lk.lock();
    it = E().by_number.find(picture_number);
    if (!E().initialised || it == E().by_number.end() || it->second != s)
        return fail(ME_GPU_ERR_NOT_FOUND, "picture released twice");
    E().slots[s].release_pending = true;
    free_slot_if_idle(s);
    return ME_GPU_OK;
}

/* one level of a registered picture's pyramid into a caller-owned padded plane (picture analysis reads the
 * device-built sixteenth plane for its histograms, pic_analysis_process.c:1560-1570, instead of building its own) */
int me_gpu_read_plane(uint64_t picture_number, int level, uint8_t *dst, int dst_stride) {
    std::unique_lock<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    auto it = E().by_number.find(picture_number);
    if (it == E().by_number.end())
        return fail(ME_GPU_ERR_NOT_FOUND, "picture not registered");
    if (level < 0 || level > 2 || !dst)
        return fail(ME_GPU_ERR_BAD_ARG, "level must be 0 (full), 1 (quarter) or 2 (sixteenth)");
    const int s = it->second;
    MePicGeom gm;
    me_make_geom(&gm, E().slots[s].width, E().slots[s].height);
    const MePlaneGeom &pl = gm.plane[level];
    const int          pw = pl.width + 2 * pl.pad;
    if (dst_stride < pw)
        return fail(ME_GPU_ERR_BAD_ARG, "destination stride smaller than the padded plane's row");
    Device &d = E().dev[0];
    CU(cudaSetDevice(d.id));
    E().slots[s].refcount++; /* the slot stays while the copy runs outside the lock */
    cudaEvent_t ready = E().slots[s].ready[0];
    const uint8_t *src = d.pool + (size_t)s * E().gmax.slot_bytes + pl.offset;
    lk.unlock();
    cudaError_t e = cudaEventSynchronize(ready);
    if (e == cudaSuccess)
        e = cudaMemcpy2D(dst, (size_t)dst_stride, src, pl.stride, pw, pl.rows, cudaMemcpyDeviceToHost);
    lk.lock();
    E().slots[s].refcount--;
    free_slot_if_idle(s);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(ME_GPU_ERR_CUDA, std::string("me_gpu_read_plane: ") + cudaGetErrorString(e));
    }
    return ME_GPU_OK;
}

/* ---- one process per GPU: pyramids travel between the processes' pools over NVLink (CUDA IPC), no collective ---------- */
int me_gpu_ipc_export_pool(unsigned char handle[64], uint64_t *slot_bytes) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    if (!handle || !slot_bytes)
        return fail(ME_GPU_ERR_BAD_ARG, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64 && sizeof(cudaIpcEventHandle_t) == 64, "IPC handles are 64 bytes");
    CU(cudaSetDevice(E().dev[0].id));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, E().dev[0].pool));
    memcpy(handle, &h, 64);
    *slot_bytes = E().gmax.slot_bytes;
    return ME_GPU_OK;
}

int me_gpu_ipc_open_peer(int peer, const unsigned char handle[64], uint64_t slot_bytes) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    if (peer < 0 || peer >= 16 || !handle)
        return fail(ME_GPU_ERR_BAD_ARG, "peer index out of range");
    if (slot_bytes != E().gmax.slot_bytes)
        return fail(ME_GPU_ERR_BAD_ARG, "the peer's pool was sized for another maximum picture size");
    if (E().peers[peer].open)
        return fail(ME_GPU_ERR_STATE, "peer already open");
    CU(cudaSetDevice(E().dev[0].id));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    void *base = nullptr;
    CU(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    E().peers[peer].open       = true;
    E().peers[peer].base       = (uint8_t *)base;
    E().peers[peer].slot_bytes = slot_bytes;
    return ME_GPU_OK;
}

int me_gpu_ipc_picture(uint64_t picture_number, int *slot, unsigned char event_handle[64]) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    auto it = E().by_number.find(picture_number);
    if (it == E().by_number.end() || !slot || !event_handle)
        return fail(ME_GPU_ERR_NOT_FOUND, "picture not registered");
    CU(cudaSetDevice(E().dev[0].id));
    Slot &sl = E().slots[it->second];
    if (!sl.ipc_ready)
        CU(cudaEventCreateWithFlags(&sl.ipc_ready, cudaEventDisableTiming | cudaEventInterprocess));
    /* everything that fills this device's copy of the slot went through the upload stream: a marker behind it */
    CU(cudaEventRecord(sl.ipc_ready, E().dev[0].upload));
    cudaIpcEventHandle_t h;
    CU(cudaIpcGetEventHandle(&h, sl.ipc_ready));
    memcpy(event_handle, &h, 64);
    *slot = it->second;
    return ME_GPU_OK;
}

int me_gpu_register_picture_from_peer(uint64_t picture_number, int peer, int peer_slot, const unsigned char event_handle[64],
                                      int width, int height) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    if (peer < 0 || peer >= 16 || !E().peers[peer].open || peer_slot < 0 || !event_handle)
        return fail(ME_GPU_ERR_BAD_ARG, "peer not open");
    if (width <= 0 || height <= 0 || (width & 7) || (height & 7) || width > E().gmax.width || height > E().gmax.height)
        return fail(ME_GPU_ERR_BAD_ARG, "bad picture geometry");
    if (E().by_number.count(picture_number))
        return fail(ME_GPU_ERR_STATE, "picture already registered");
    int s = -1;
    for (int i = 0; i < E().nslots; i++)
        if (!E().slots[i].used) {
            s = i;
            break;
        }
    if (s < 0)
        return fail(ME_GPU_ERR_NO_MEMORY, "picture pool is full (raise pictures_in_flight or release pictures)");
    Engine::Peer &pr = E().peers[peer];
    Device       &d  = E().dev[0];
    CU(cudaSetDevice(d.id));
    cudaEvent_t ev = nullptr;
    auto        f  = pr.ev.find(peer_slot);
    if (f == pr.ev.end()) { /* one event per peer slot for the life of the engine */
        cudaIpcEventHandle_t h;
        memcpy(&h, event_handle, 64);
        CU(cudaIpcOpenEventHandle(&ev, h));
        pr.ev[peer_slot] = ev;
    } else
        ev = f->second;
    MePicGeom gm;
    me_make_geom(&gm, width, height);
    /* the owner recorded the event behind its pyramid kernels before it handed out the handle: the copy starts after them */
    CU(cudaStreamWaitEvent(d.upload, ev, 0));
    CU(cudaMemcpyAsync(d.pool + (size_t)s * E().gmax.slot_bytes, pr.base + (size_t)peer_slot * pr.slot_bytes, gm.slot_bytes,
                       cudaMemcpyDeviceToDevice, d.upload));
    E().peer_bytes += gm.slot_bytes;
    Slot &sl          = E().slots[s];
    sl.used           = true;
    sl.picture_number = picture_number;
    sl.width          = width;
    sl.height         = height;
    sl.refcount       = 0;
    sl.release_pending = false;
    for (auto &v : sl.ready_seen) v = false;
    E().by_number[picture_number] = s;
    CU(cudaEventRecord(sl.ready[0], d.upload));
    for (size_t o = 1; o < E().dev.size(); o++) { /* several devices in this process as well: replicate as usual */
        Device &p2 = E().dev[o];
        CU(cudaSetDevice(p2.id));
        CU(cudaStreamWaitEvent(p2.upload, sl.ready[0], 0));
        CU(cudaMemcpyPeerAsync(p2.pool + (size_t)s * E().gmax.slot_bytes, p2.id, d.pool + (size_t)s * E().gmax.slot_bytes, d.id,
                               E().gmax.slot_bytes, p2.upload));
        CU(cudaEventRecord(sl.ready[o], p2.upload));
    }
    return ME_GPU_OK;
}

uint64_t me_gpu_ipc_bytes_pulled(void) {
    std::lock_guard<std::mutex> lk(E().mu);
    return E().peer_bytes;
}

int me_gpu_read_pyramid(uint64_t picture_number, uint8_t *full, uint8_t *quarter, uint8_t *sixteenth) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    auto it = E().by_number.find(picture_number);
    if (it == E().by_number.end())
        return fail(ME_GPU_ERR_NOT_FOUND, "picture not registered");
    const int s = it->second;
    MePicGeom gm;
    me_make_geom(&gm, E().slots[s].width, E().slots[s].height);
    /* read from the last device: proves the replication as well when there are several */
    const size_t di = E().dev.size() - 1;
    Device      &d  = E().dev[di];
    CU(cudaSetDevice(d.id));
    CU(cudaEventSynchronize(E().slots[s].ready[di]));
    uint8_t *dst[3] = {full, quarter, sixteenth};
    for (int l = 0; l < 3; l++) {
        if (!dst[l])
            continue;
        const MePlaneGeom &pl = gm.plane[l];
        const int          pw = pl.width + 2 * pl.pad;
        CU(cudaMemcpy2D(dst[l], pw, d.pool + (size_t)s * E().gmax.slot_bytes + pl.offset, pl.stride, pw, pl.rows,
                        cudaMemcpyDeviceToHost));
    }
    return ME_GPU_OK;
}

/* enqueue the device->host copies of a finished (or running) job on its stream. Product outputs live in one
 * device slab in MeGpuOutputs order: destination arrays that are adjacent in the same relative layout
 * (me_gpu_output_layout) are fetched with a single copy. */
static int enqueue_d2h(JobCtx *jc, const MeGpuOutputs *out) {
    const MeJob    &j  = *jc->h_job;
    const size_t    nb = (size_t)j.g.b64_count;
    const OutLayout L  = output_layout(nb, j.pu_count, j.p.max_cand, j.p.max_refs);
    cudaStream_t    st = jc->tail;
    uint8_t *dst[11] = {out->total_me_candidate_index, out->me_candidate_array, (uint8_t *)out->me_mv_array,
                        (uint8_t *)out->me_64x64_distortion, (uint8_t *)out->me_32x32_distortion,
                        (uint8_t *)out->me_16x16_distortion, (uint8_t *)out->me_8x8_distortion,
                        (uint8_t *)out->me_8x8_cost_variance, (uint8_t *)out->rc_me_distortion,
                        out->stationary_block_present_sb, out->rc_me_allow_gm};
    if (j.p.me_type == ME_GPU_ME_MCTF) { /* the tf_* arrays are the job's results; adjacent destinations share a copy */
        const TfLayout T = tf_layout(nb);
        uint8_t *tdst[5] = {(uint8_t *)out->tf_best_sad, (uint8_t *)out->tf_best_mv, (uint8_t *)out->tf_hme_centre,
                            (uint8_t *)out->tf_hme_sad, out->tf_early_exit};
        for (int k = 0; k < 5;) {
            if (!tdst[k]) {
                k++;
                continue;
            }
            int    e = k;
            size_t n = T.bytes[k];
            while (e + 1 < 5 && tdst[e + 1] && tdst[e + 1] == tdst[k] + n) n += T.bytes[++e];
            CU(cudaMemcpyAsync(tdst[k], jc->out_slab + T.off[k], n, cudaMemcpyDeviceToHost, st));
            k = e + 1;
        }
        for (auto &d : dst) d = nullptr;
    }
    /* runs of destination arrays that are exactly adjacent in slab order come back with one copy */
    for (int k = 0; k < 11;) {
        const int i = L.order[k];
        if (!dst[i]) {
            k++;
            continue;
        }
        int    e = k;
        size_t n = L.bytes[i];
        while (e + 1 < 11 && dst[L.order[e + 1]] && dst[L.order[e + 1]] == dst[i] + n) {
            e++;
            n += L.bytes[L.order[e]];
        }
        CU(cudaMemcpyAsync(dst[i], jc->out_slab + L.off[i], n, cudaMemcpyDeviceToHost, st));
        k = e + 1;
    }
#define D2H(dstp, src, bytes)                                                               \
    do {                                                                                    \
        if (dstp)                                                                           \
            CU(cudaMemcpyAsync((dstp), (src), (bytes), cudaMemcpyDeviceToHost, st));        \
    } while (0)
    D2H(out->p_sb_best_sad, jc->dbg_sad, nb * ME_NREF * ME_NPU * 4);
    D2H(out->p_sb_best_mv, jc->dbg_mv, nb * ME_NREF * ME_NPU * 4);
    D2H(out->hme_search_centre, jc->dbg_centre, nb * ME_NREF * 2 * sizeof(int16_t));
    D2H(out->hme_sad, jc->dbg_hme, nb * ME_NREF * sizeof(uint64_t));
    D2H(out->do_ref, jc->dbg_do_ref, nb * ME_NREF);
    D2H(out->search_area, jc->dbg_area, nb * ME_NREF * 4 * sizeof(int16_t));
#undef D2H
    CU(cudaMemcpyAsync(jc->h_tail, &jc->d_head->tail, sizeof(JobTail), cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(jc->ev[4], st));
    jc->d2h_enqueued = true;
    return ME_GPU_OK;
}

static int submit_locked(uint64_t picture_number, const MeGpuParams *params,
                         const uint64_t ref_picture_numbers[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS], const MeGpuOutputs *out) {
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    if (!params || !ref_picture_numbers)
        return fail(ME_GPU_ERR_BAD_ARG, "null argument");
    const MeGpuParams &p = *params;
    if (p.num_hme_sa_w != 2 || p.num_hme_sa_h != 2) /* get_worst_quadrant asserts it, motion_estimation.c:1875 */
        return fail(ME_GPU_ERR_UNSUPPORTED, "num_hme_sa_w/h must be 2");
    if (p.num_of_list_to_search < 1 || p.num_of_list_to_search > ME_NL || p.num_of_ref_pic_to_search[0] < 1 ||
        p.num_of_ref_pic_to_search[0] > ME_NR || p.num_of_ref_pic_to_search[1] > ME_NR)
        return fail(ME_GPU_ERR_BAD_ARG, "bad reference counts");
    if (p.max_number_of_pus_per_sb != ME_NPU)
        return fail(ME_GPU_ERR_UNSUPPORTED, "max_number_of_pus_per_sb must be 85");
    if (p.me_type != ME_GPU_ME_OPEN_LOOP && p.me_type != ME_GPU_ME_MCTF)
        return fail(ME_GPU_ERR_BAD_ARG, "me_type must be ME_GPU_ME_OPEN_LOOP or ME_GPU_ME_MCTF");
    if (p.me_type == ME_GPU_ME_MCTF && (p.num_of_list_to_search != 1 || p.num_of_ref_pic_to_search[0] != 1))
        return fail(ME_GPU_ERR_BAD_ARG, "temporal-filter ME takes one list with one reference");
    auto it = E().by_number.find(picture_number);
    if (it == E().by_number.end())
        return fail(ME_GPU_ERR_NOT_FOUND, "source picture not registered");
    const int   src_slot = it->second;
    const Slot &src      = E().slots[src_slot];
    /* device: pictures are dealt round-robin */
    const size_t di = (size_t)(E().submit_counter % E().dev.size());
    Device      &d  = E().dev[di];
    JobCtx      *jc = nullptr;
    for (auto &c : d.jobs)
        if (!c.busy) {
            jc = &c;
            break;
        }
    if (!jc)
        return fail(ME_GPU_ERR_STATE, "too many pictures in flight: call me_gpu_wait first");
    for (auto &dv : E().dev)
        for (auto &c : dv.jobs)
            if (c.busy && c.picture_number == picture_number && c.me_type == p.me_type && c.ref00 == ref_picture_numbers[0][0])
                return fail(ME_GPU_ERR_STATE, "picture already submitted (same me_type and first reference)");

    MeJob &j = *jc->h_job;
    memset(jc->h_head, 0, sizeof(JobHead));
    j.p = p;
    me_make_geom(&j.g, src.width, src.height);
    j.g.slot_bytes = E().gmax.slot_bytes;
    j.src_slot     = (uint16_t)src_slot;
    jc->slots_used.clear();
    jc->slots_used.push_back(src_slot);
    for (int li = 0; li < p.num_of_list_to_search; li++)
        for (int ri = 0; ri < p.num_of_ref_pic_to_search[li]; ri++) {
            auto r = E().by_number.find(ref_picture_numbers[li][ri]);
            if (r == E().by_number.end())
                return fail(ME_GPU_ERR_NOT_FOUND, "reference picture not registered");
            const Slot &rs = E().slots[r->second];
            if (rs.width != src.width || rs.height != src.height)
                return fail(ME_GPU_ERR_BAD_ARG, "reference picture size differs from the source (get_me_reference warns and skips)");
            j.ref_slot[li][ri]           = (uint16_t)r->second;
            j.ref_picture_number[li][ri] = ref_picture_numbers[li][ri];
            jc->slots_used.push_back(r->second);
        }
    j.pu_count = (uint32_t)me_exported_pu_count(&p);
    const size_t nb = (size_t)j.g.b64_count;
    if (p.max_cand == 0 || p.max_refs == 0)
        return fail(ME_GPU_ERR_BAD_ARG, "max_cand / max_refs must be set (pcs->pa_me_data)");
    const OutLayout L = output_layout(nb, j.pu_count, p.max_cand, p.max_refs);
    if (L.total > jc->out_cap)
        return fail(ME_GPU_ERR_BAD_ARG, "max_cand / max_refs exceed what me_gpu_init allocated");
    if (max_tiles_per_b64(&p) * nb > jc->tile_cap) {
        /* screen-content presets search up to 750 x 750 (x2 by MV) positions per reference: the job's tile list grows on
           demand (the context is idle: its last job has been retired). Rare, and the one place a submit allocates. */
        const uint64_t want = max_tiles_per_b64(&p) * nb;
        if (want > (1ull << 28))
            return fail(ME_GPU_ERR_UNSUPPORTED, "search area too large for the tile list");
        CU(cudaSetDevice(d.id));
        CU(cudaStreamSynchronize(jc->tail));
        cudaFree(jc->tiles);
        jc->tiles    = nullptr;
        jc->tile_cap = 0;
        if (dalloc(&jc->tiles, (size_t)want) != cudaSuccess) {
            cudaGetLastError();
            return fail(ME_GPU_ERR_NO_MEMORY, "no device memory for the full-search tile list of this search area");
        }
        jc->tile_cap = (uint32_t)want;
    }
    if (out && out->b64_count < nb)
        return fail(ME_GPU_ERR_BAD_ARG, "MeGpuOutputs.b64_count is smaller than the picture's b64 count");
    j.state = jc->state; j.tasks = jc->tasks; j.results = jc->results; j.fs_tasks = jc->fs_tasks; j.keys = jc->keys;
    j.total_me_candidate_index = jc->out_slab + L.off[0];
    j.me_candidate_array       = jc->out_slab + L.off[1];
    j.me_mv_array              = (uint32_t *)(jc->out_slab + L.off[2]);
    j.dist                     = (uint32_t *)(jc->out_slab + L.off[3]);
    j.flags                    = jc->out_slab + L.off[9];
    j.dbg_best_sad = jc->dbg_sad; j.dbg_best_mv = jc->dbg_mv; j.dbg_centre = jc->dbg_centre; j.dbg_hme_sad = jc->dbg_hme;
    j.dbg_do_ref = jc->dbg_do_ref; j.dbg_area = jc->dbg_area;
    if (out) { /* destinations known up front (me_gpu_submit_out): taps nobody asked for are not written at all */
        if (!out->p_sb_best_sad) j.dbg_best_sad = nullptr;
        if (!out->p_sb_best_mv) j.dbg_best_mv = nullptr;
        if (!out->hme_search_centre) j.dbg_centre = nullptr;
        if (!out->hme_sad) j.dbg_hme_sad = nullptr;
        if (!out->do_ref) j.dbg_do_ref = nullptr;
        if (!out->search_area) j.dbg_area = nullptr;
    }
    j.ops_total = jc->d_head->tail.ops;
    {
        const TfLayout T = tf_layout(nb); /* far below the open-loop layout's size */
        j.tf_best_sad    = (uint32_t *)(jc->out_slab + T.off[0]);
        j.tf_best_mv     = (uint32_t *)(jc->out_slab + T.off[1]);
        j.tf_hme_centre  = (int16_t *)(jc->out_slab + T.off[2]);
        j.tf_hme_sad     = (uint32_t *)(jc->out_slab + T.off[3]);
        j.tf_early_exit  = jc->out_slab + T.off[4];
        if (T.total > jc->out_cap)
            return fail(ME_GPU_ERR_BAD_ARG, "result slab too small for the temporal-filter outputs");
    }
    JobTail &tl = jc->h_head->tail;
    tl.list.count = tl.list.next = tl.list.overflow = 0;
    tl.list.cap   = jc->tile_cap;
    tl.list.tiles = jc->tiles;
    tl.ops[0] = tl.ops[1] = tl.ops[2] = tl.ops[3] = tl.ops[4] = 0;

    CU(cudaSetDevice(d.id));
    cudaStream_t st = jc->stream;
    for (int s : jc->slots_used) {
        Slot &sl = E().slots[s];
        if (sl.ready_seen[di])
            continue;
        if (cudaEventQuery(sl.ready[di]) == cudaSuccess)
            sl.ready_seen[di] = true;
        else {
            cudaGetLastError();
            CU(cudaStreamWaitEvent(st, sl.ready[di], 0));
        }
    }
    CU(cudaEventRecord(jc->ev[0], st));
    CU(cudaMemcpyAsync(jc->d_head, jc->h_head, sizeof(JobHead), cudaMemcpyHostToDevice, st));
    /* candidate / MV entries that no PU writes must read 0, as in the zero-initialised MeSbResults */
    CU(cudaMemsetAsync(jc->out_slab, 0, L.total, st));

    const int nbi = (int)nb;
    jc->npairs = 0;
    const bool prof = E().profiling;
    /* profiling mode: a (begin, end) event pair around each launch, tagged with its kernel class */
    auto pbegin = [&](int kind) {
        if (prof && jc->npairs < 32) {
            jc->pkind[jc->npairs] = kind;
            cudaEventRecord(jc->pev[2 * jc->npairs], st);
        }
    };
    auto pend = [&]() {
        if (prof && jc->npairs < 32) {
            cudaEventRecord(jc->pev[2 * jc->npairs + 1], st);
            jc->npairs++;
        }
    };
    {
        /* the whole per-b64 chain in one launch; large search areas come back as tiles */
        const HmeMaps *maps = nullptr;
        for (auto &ms : d.mapsets)
            if (ms.width == j.g.width && ms.height == j.g.height)
                maps = &ms.maps;
        if (!maps) {
            Device::MapSet ms;
            ms.width  = j.g.width;
            ms.height = j.g.height;
            const int rc = me_hme_make_maps(&ms.maps, d.pool, &j.g, E().nslots);
            if (rc)
                return fail(ME_GPU_ERR_CUDA, rc < 0 ? "the driver does not export cuTensorMapEncodeTiled"
                                                    : "cuTensorMapEncodeTiled failed (CUresult " + std::to_string(rc) + ")");
            d.mapsets.reserve(16); /* pointers into the vector stay valid for the launches in flight (by-value copies anyway) */
            d.mapsets.push_back(ms);
            maps = &d.mapsets.back().maps;
        }
        const bool two_pass = mc_fs_two_pass(&p) != 0;
        pbegin(1);
        me_launch_hme(jc->d_job, 1, nbi, two_pass ? 1 : 0, jc->d_list, d.pool, maps, st);
        pend();
        if (prof) CU(cudaEventRecord(jc->ev[1], st));
        if (jc->tail != st) { /* the rest of the picture goes to the higher-priority stream */
            CU(cudaEventRecord(jc->ev_mid, st));
            CU(cudaStreamWaitEvent(jc->tail, jc->ev_mid, 0));
            st = jc->tail;
        }
#ifdef ME_GPU_DEBUG
        static const int skip = getenv("ME_GPU_SKIP") ? atoi(getenv("ME_GPU_SKIP")) : 0; /* timing experiments only: results wrong */
#else
        const int skip = 0;
#endif
        pbegin(3);
        if (!(skip & 1))
            me_launch_fullsearch(jc->d_job, jc->d_list, d.pool, maps, d.sms, st);
        pend();
        if (two_pass) {
            /* enable_me_sr_adjustment == 2 (ME:1355-1364): the other references' search areas depend on the finished
               search of reference (0, 0), tiles included: their integer-search stages run now, then their tiles */
            CU(cudaMemsetAsync(jc->d_list, 0, 2 * sizeof(unsigned int), st)); /* count = next = 0 */
            pbegin(1);
            me_launch_hme(jc->d_job, 1, nbi, 2, jc->d_list, d.pool, maps, st);
            pend();
            pbegin(3);
            me_launch_fullsearch(jc->d_job, jc->d_list, d.pool, maps, d.sms, st);
            pend();
        }
    }
    if (jc->tail != st) {
        CU(cudaEventRecord(jc->ev_mid, st));
        CU(cudaStreamWaitEvent(jc->tail, jc->ev_mid, 0));
        st = jc->tail;
    }
    if (prof) CU(cudaEventRecord(jc->ev[2], st));
    pbegin(4);
#ifdef ME_GPU_DEBUG
    static const int skip_epi = getenv("ME_GPU_SKIP") ? atoi(getenv("ME_GPU_SKIP")) & 2 : 0; /* timing experiments only */
#else
    const int skip_epi = 0;
#endif
    if (!skip_epi)
        me_launch_epilogue(jc->d_job, 1, nbi, st);
    pend();
    if (prof) CU(cudaEventRecord(jc->ev[3], st));
    CU(cudaGetLastError());
    jc->d2h_enqueued = false;
    jc->prof         = prof;
    if (out) {
        int rc = enqueue_d2h(jc, out);
        if (rc)
            return rc;
    }
    jc->busy           = true;
    jc->retiring       = false;
    jc->picture_number = picture_number;
    jc->me_type        = p.me_type;
    jc->ref00          = ref_picture_numbers[0][0];
    jc->ticket         = ++E().ticket;
    for (int s : jc->slots_used) E().slots[s].refcount++;
    E().submit_counter++;
    return ME_GPU_OK;
}

int me_gpu_submit(uint64_t picture_number, const MeGpuParams *params,
                  const uint64_t ref_picture_numbers[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS]) {
    std::lock_guard<std::mutex> lk(E().mu);
    return submit_locked(picture_number, params, ref_picture_numbers, nullptr);
}

int me_gpu_submit_out(uint64_t picture_number, const MeGpuParams *params,
                      const uint64_t ref_picture_numbers[ME_GPU_MAX_LISTS][ME_GPU_MAX_REFS], MeGpuOutputs *out) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!out)
        return fail(ME_GPU_ERR_BAD_ARG, "null outputs");
    return submit_locked(picture_number, params, ref_picture_numbers, out);
}

int me_gpu_output_layout(int width, int height, const MeGpuParams *params, size_t offsets[11], size_t *total_bytes) {
    if (!params || !offsets || !total_bytes || width <= 0 || height <= 0)
        return ME_GPU_ERR_BAD_ARG;
    const size_t    nb = (size_t)((width + 63) / 64) * (size_t)((height + 63) / 64);
    const OutLayout L  = output_layout(nb, (size_t)me_exported_pu_count(params), params->max_cand, params->max_refs);
    for (int i = 0; i < 11; i++) offsets[i] = L.off[i];
    *total_bytes = L.total;
    return ME_GPU_OK;
}

/* the oldest job in flight of a picture (any me_type / reference when !exact) */
static JobCtx *find_job(uint64_t picture_number, bool exact, uint8_t me_type, uint64_t ref00, Device **dv) {
    JobCtx *jc = nullptr;
    for (auto &d : E().dev)
        for (auto &c : d.jobs)
            if (c.busy && c.picture_number == picture_number && (!exact || (c.me_type == me_type && c.ref00 == ref00)) &&
                (!jc || c.ticket < jc->ticket)) {
                jc  = &c;
                *dv = &d;
            }
    return jc;
}

/* a job leaves the engine: on success and on EVERY failure after its submission succeeded */
static void retire_locked(JobCtx *jc) {
    jc->busy     = false;
    jc->retiring = false;
    for (int s : jc->slots_used) {
        E().slots[s].refcount--;
        free_slot_if_idle(s);
    }
    E().done_ring[E().done_next % 64]  = jc->picture_number;
    E().done_valid[E().done_next % 64] = true;
    E().done_next++;
    E().cv.notify_all();
}

/* out == nullptr: cancel (wait for the device to be done with the job, deliver nothing) */
static int wait_impl(uint64_t picture_number, bool exact, uint8_t me_type, uint64_t ref00, MeGpuOutputs *out) {
    std::unique_lock<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    Device *dv = nullptr;
    JobCtx *jc = find_job(picture_number, exact, me_type, ref00, &dv);
    if (!jc)
        return fail(ME_GPU_ERR_NOT_FOUND, "picture was not submitted");
    if (jc->retiring)
        return fail(ME_GPU_ERR_STATE, "another thread is already waiting for this job's results (me_gpu_wait_done blocks until it has them)");
    const MeJob   &j  = *jc->h_job;
    const size_t   nb = (size_t)j.g.b64_count;
    if (out && out->b64_count < nb) /* nothing has been touched: the caller may come back with a large enough buffer, or cancel */
        return fail(ME_GPU_ERR_BAD_ARG, "MeGpuOutputs.b64_count is smaller than the picture's b64 count");
    jc->retiring = true; /* from here on this thread owns the job's retirement, whatever happens */
    cudaError_t e = cudaSetDevice(dv->id);
    int         rc = ME_GPU_OK;
    if (e == cudaSuccess && out && !jc->d2h_enqueued)
        rc = enqueue_d2h(jc, out);
    cudaStream_t tail = jc->tail;
    cudaEvent_t  done = jc->ev[4];
    const bool   have_event = jc->d2h_enqueued;
    lk.unlock();
    if (e == cudaSuccess) /* without result copies (cancel of a plain submit, or a failed enqueue): the stream itself */
        e = have_event && rc == ME_GPU_OK ? cudaEventSynchronize(done) : cudaStreamSynchronize(tail);
    lk.lock();
    if (e != cudaSuccess || rc != ME_GPU_OK) {
        retire_locked(jc);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(ME_GPU_ERR_CUDA, std::string("job failed: ") + cudaGetErrorString(e));
        }
        return rc;
    }
    if (!out) {
        retire_locked(jc);
        return ME_GPU_OK;
    }
    const unsigned overflow = jc->h_tail->list.overflow;
    MeGpuTimings t{};
    cudaEventElapsedTime(&t.total_ms, jc->ev[0], jc->ev[4]);
    if (jc->prof) {
        cudaEventElapsedTime(&t.hme_ms, jc->ev[0], jc->ev[1]);
        cudaEventElapsedTime(&t.search_ms, jc->ev[1], jc->ev[2]);
        cudaEventElapsedTime(&t.epilogue_ms, jc->ev[2], jc->ev[3]);
    }
    for (int i = 0; i < jc->npairs; i++) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, jc->pev[2 * i], jc->pev[2 * i + 1]);
        switch (jc->pkind[i]) {
        case 1: t.k_search_ms += ms; break;
        case 3: t.k_fullsearch_ms += ms; t.k_fullsearch_main_ms += ms; break;
        default: t.k_epilogue_ms += ms; break;
        }
    }
    t.kernel_launches = me_launch_count() + me_hme_launch_count() + me_dg_launch_count();
    t.byte_sad_ops    = jc->h_tail->ops[0];
    t.byte_sad_ops_fullsearch = jc->h_tail->ops[1];
    t.byte_sad_ops_fullsearch_inline = jc->h_tail->ops[4];
    E().last            = t;
    out->b64_count    = (uint32_t)nb;
    out->pu_count     = j.pu_count;
    if (j.p.me_type == ME_GPU_ME_MCTF) {
        out->tf_tot_horz_blks = (uint32_t)jc->h_tail->ops[2];
        out->tf_tot_vert_blks = (uint32_t)jc->h_tail->ops[3];
    }
    retire_locked(jc);
    if (overflow)
        return fail(ME_GPU_ERR_UNSUPPORTED, overflow == 2 ? "a search window exceeds the staging buffer of k_hme"
                                                           : "full-search tile list overflowed");
    return ME_GPU_OK;
}

int me_gpu_wait(uint64_t picture_number, MeGpuOutputs *out) {
    if (!out)
        return fail(ME_GPU_ERR_BAD_ARG, "null outputs");
    return wait_impl(picture_number, false, 0, 0, out);
}

int me_gpu_wait_pair(uint64_t picture_number, uint8_t me_type, uint64_t first_ref_picture_number, MeGpuOutputs *out) {
    if (!out)
        return fail(ME_GPU_ERR_BAD_ARG, "null outputs");
    return wait_impl(picture_number, true, me_type, first_ref_picture_number, out);
}

int me_gpu_cancel(uint64_t picture_number) { return wait_impl(picture_number, false, 0, 0, nullptr); }

int me_gpu_wait_done(uint64_t picture_number) {
    std::unique_lock<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    Device *dv   = nullptr;
    JobCtx *jc   = find_job(picture_number, false, 0, 0, &dv);
    if (!jc) {
        for (int i = 0; i < 64; i++)
            if (E().done_valid[i] && E().done_ring[i] == picture_number)
                return ME_GPU_OK;
        return fail(ME_GPU_ERR_NOT_FOUND, "picture was not submitted (or was retired more than 64 jobs ago)");
    }
    const uint64_t ticket = jc->ticket;
    E().cv.wait(lk, [&] { return !E().initialised || !jc->busy || jc->ticket != ticket; });
    return ME_GPU_OK;
}

/* TASK_DG_DETECTOR_HME of one (source, reference) pair: one launch of k_dg_hme on the detector's stream, the four
 * sums come back in one 24-byte copy. Calls alternate between the devices (every device holds every picture). */
int me_gpu_dg_detector(uint64_t src_picture_number, uint64_t ref_picture_number, uint8_t input_resolution,
                       MeGpuDgMetrics *metrics) {
    std::lock_guard<std::mutex>  dg(E().dg_mu);
    std::unique_lock<std::mutex> lk(E().mu);
    if (!E().initialised)
        return fail(ME_GPU_ERR_STATE, "not initialised");
    if (!metrics)
        return fail(ME_GPU_ERR_BAD_ARG, "null metrics");
    auto si = E().by_number.find(src_picture_number), ri = E().by_number.find(ref_picture_number);
    if (si == E().by_number.end() || ri == E().by_number.end())
        return fail(ME_GPU_ERR_NOT_FOUND, "picture not registered");
    const int ss = si->second, rs = ri->second;
    if (E().slots[ss].width != E().slots[rs].width || E().slots[ss].height != E().slots[rs].height)
        return fail(ME_GPU_ERR_BAD_ARG, "source and reference picture sizes differ");
    const size_t di = (size_t)(E().dg_counter++ % E().dev.size());
    Device      &d  = E().dev[di];
    MePicGeom    gm;
    me_make_geom(&gm, E().slots[ss].width, E().slots[ss].height);
    gm.slot_bytes = E().gmax.slot_bytes;
    /* pd_process.c:497-498: 16 up to 360p, 64 up to 480p, else 128 (EbInputResolution, definitions.h:2079) */
    const int sa = input_resolution <= 1 ? 16 : input_resolution <= 2 ? 64 : 128;
    CU(cudaSetDevice(d.id));
    for (int s : {ss, rs}) {
        Slot &sl = E().slots[s];
        if (!sl.ready_seen[di])
            CU(cudaStreamWaitEvent(d.dg_stream, sl.ready[di], 0));
        sl.refcount++;
    }
    cudaError_t e = cudaMemsetAsync(d.dg_acc, 0, sizeof(MeGpuDgMetrics), d.dg_stream);
    if (e == cudaSuccess) {
        me_launch_dg(d.pool, &gm, (uint32_t)ss, (uint32_t)rs, sa, d.dg_acc, d.dg_stream);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(d.dg_host, d.dg_acc, sizeof(MeGpuDgMetrics), cudaMemcpyDeviceToHost, d.dg_stream);
    lk.unlock();
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(d.dg_stream);
    lk.lock();
    for (int s : {ss, rs}) {
        E().slots[s].refcount--;
        free_slot_if_idle(s);
    }
    if (e != cudaSuccess)
        return fail(ME_GPU_ERR_CUDA, std::string("dynamic-GOP detector failed: ") + cudaGetErrorString(e));
    *metrics = *d.dg_host;
    return ME_GPU_OK;
}

int me_gpu_get_timings(MeGpuTimings *t) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!t)
        return ME_GPU_ERR_BAD_ARG;
    *t                 = E().last;
    t->kernel_launches = me_launch_count() + me_hme_launch_count() + me_dg_launch_count();
    return ME_GPU_OK;
}

void me_gpu_set_profiling(int on) {
    std::lock_guard<std::mutex> lk(E().mu);
    E().profiling = on != 0;
}

/* VABSDIFF4 lane-operations per second of device 0 of the engine (x4 = byte-SADs per second): the measured
 * denominator of the integer-SAD roofline */
double me_gpu_measure_sad_peak(void) {
    std::lock_guard<std::mutex> lk(E().mu);
    if (!E().initialised)
        return 0.0;
    cudaSetDevice(E().dev[0].id);
    return me_measure_sad_peak(E().dev[0].upload);
}

} /* extern "C" */

/* ---- more than one engine per process -------------------------------------------------------------------------- */
extern "C" int me_gpu_engine_create(MeGpuEngine **out) {
    if (!out)
        return fail(ME_GPU_ERR_BAD_ARG, "me_gpu_engine_create: null argument");
    *out = reinterpret_cast<MeGpuEngine *>(new (std::nothrow) Engine());
    return *out ? ME_GPU_OK : fail(ME_GPU_ERR_NO_MEMORY, "me_gpu_engine_create: out of host memory");
}

extern "C" void me_gpu_engine_bind(MeGpuEngine *e) { t_eng = e ? reinterpret_cast<Engine *>(e) : &g_default; }

extern "C" int me_gpu_engine_destroy(MeGpuEngine *e) {
    Engine *en = reinterpret_cast<Engine *>(e);
    if (!en)
        return ME_GPU_OK;
    if (en->initialised)
        return fail(ME_GPU_ERR_STATE, "me_gpu_engine_destroy: the engine is still initialised (bind it and call me_gpu_shutdown first)");
    if (t_eng == en)
        t_eng = &g_default;
    delete en;
    return ME_GPU_OK;
}
