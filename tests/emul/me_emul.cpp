/*
 * me_emul.cpp - TEST HARNESS (not a product path, never loaded by the engine).
 *
 * Runs the product's control logic (svt-av1-psyex_b200/csrc/me_ctrl.h, the exact code the device runs in
 * k_ctrl / k_epilogue) on the CPU, with trivial scalar loops standing in for the three CUDA search kernels.
 * `pytest -m "not gpu"` uses it to pin the control logic against the oracle without a GPU, so that a GPU
 * mismatch can only come from the search kernels themselves.
 */
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "me_ctrl.h"
#include "me_oracle.h" /* OraclePicture / OraclePlane layouts only */

namespace {

const OraclePlane *plane_of(const OraclePicture *pic, int level) {
    return level == 0 ? &pic->full : level == 1 ? &pic->quarter : &pic->sixteenth;
}
const uint8_t *at(const OraclePlane *pl, int x, int y) {
    return pl->buf + (ptrdiff_t)(pl->org_y + y) * pl->stride + pl->org_x + x;
}

void exec_search(const SearchTask &t, const OraclePicture *src, const OraclePicture *ref, SearchResult &r) {
    const OraclePlane *sp = plane_of(src, t.level), *rp = plane_of(ref, t.level);
    r.sad   = ME_SAD_LOOP_INIT;
    r.found = 0;
    r.x = r.y = 0;
    const bool skip = t.bw == 16 && t.rows <= 16 && t.skip_line;
    for (int y = 0; y < t.sa_h; y++) {
        if (skip && (y & 1) == 0)
            continue;
        for (int x = 0; x < t.sa_w; x++) {
            uint32_t sad = 0;
            for (int row = 0; row < t.rows; row++) {
                const uint8_t *a = at(sp, t.sx, t.sy + row * t.row_step);
                const uint8_t *b = at(rp, t.rx + x, t.ry + y + row * t.row_step);
                for (int i = 0; i < t.bw; i++) sad += (uint32_t)abs((int)a[i] - (int)b[i]);
            }
            if (sad < r.sad) {
                r.sad   = sad;
                r.x     = (int16_t)x;
                r.y     = (int16_t)y;
                r.found = 1;
            }
        }
    }
}

int zidx(int bx, int by, int levels) {
    int z = 0;
    for (int b = 0; b < levels; b++) z |= (((bx >> b) & 1) << (2 * b)) | (((by >> b) & 1) << (2 * b + 1));
    return z;
}

void exec_fs(const FsTask &t, const OraclePicture *src, const OraclePicture *ref, int org_x, int org_y, bool sub,
             unsigned long long *keys) {
    for (int y = 0; y < t.h; y++)
        for (int x = 0; x < t.w; x++) {
            uint32_t s[ME_NPU];
            memset(s, 0, sizeof s);
            for (int by = 0; by < 8; by++)
                for (int bx = 0; bx < 8; bx++) {
                    uint32_t v = 0;
                    for (int row = 0; row < 8; row += sub ? 2 : 1) {
                        const uint8_t *a = at(&src->full, org_x + 8 * bx, org_y + 8 * by + row);
                        const uint8_t *b = at(&ref->full, org_x + t.x0 + x + 8 * bx, org_y + t.y0 + y + 8 * by + row);
                        for (int i = 0; i < 8; i++) v += (uint32_t)abs((int)a[i] - (int)b[i]);
                    }
                    if (sub)
                        v <<= 1;
                    s[21 + zidx(bx, by, 3)] = v;
                    s[5 + zidx(bx >> 1, by >> 1, 2)] += v;
                    s[1 + zidx(bx >> 2, by >> 2, 1)] += v;
                    s[0] += v;
                }
            const uint32_t ord = t.probe ? 0u : 1u + (uint32_t)(y * t.w + x);
            for (int n = 0; n < ME_NPU; n++) {
                unsigned long long k = ((unsigned long long)s[n] << 32) | ord;
                if (k < keys[n])
                    keys[n] = k;
            }
        }
}

} // namespace

extern "C" int emul_me_picture(const OraclePicture *src, const OraclePicture *const refs[ME_NL][ME_NR],
                               const MeGpuParams *params, MeGpuOutputs *out, unsigned long long *ops_out) {
    MeJob j;
    memset(&j, 0, sizeof j);
    j.p = *params;
    me_make_geom(&j.g, (src->full.width + 7) & ~7, (src->full.height + 7) & ~7);
    const int nb = j.g.b64_count;
    if ((int)out->b64_count < nb)
        return ME_GPU_ERR_BAD_ARG;
    const OraclePicture *slot_pic[1 + ME_NREF] = {src};
    j.src_slot                                 = 0;
    for (int li = 0; li < ME_NL; li++)
        for (int ri = 0; ri < ME_NR; ri++) {
            j.ref_slot[li][ri]           = (uint16_t)(1 + li * ME_NR + ri);
            slot_pic[1 + li * ME_NR + ri] = refs[li][ri];
            j.ref_picture_number[li][ri] = refs[li][ri] ? refs[li][ri]->picture_number : 0;
        }
    j.pu_count = (uint32_t)me_exported_pu_count(params);
    std::vector<B64State>           state(nb);
    std::vector<SearchTask>         tasks((size_t)nb * ME_SLOTS);
    std::vector<SearchResult>       results((size_t)nb * ME_SLOTS);
    std::vector<FsTask>             fs((size_t)nb * ME_NREF);
    std::vector<unsigned long long> keys((size_t)nb * ME_NREF * ME_NPU);
    std::vector<uint32_t>           dist((size_t)6 * nb);
    std::vector<uint8_t>            flags((size_t)2 * nb);
    memset(tasks.data(), 0, tasks.size() * sizeof(SearchTask));
    j.state = state.data(); j.tasks = tasks.data(); j.results = results.data(); j.fs_tasks = fs.data();
    j.keys = keys.data(); j.dist = dist.data(); j.flags = flags.data();
    j.total_me_candidate_index = out->total_me_candidate_index;
    j.me_candidate_array       = out->me_candidate_array;
    j.me_mv_array              = out->me_mv_array;
    j.dbg_best_sad = out->p_sb_best_sad; j.dbg_best_mv = out->p_sb_best_mv; j.dbg_centre = out->hme_search_centre;
    j.dbg_hme_sad = (uint64_t *)out->hme_sad; j.dbg_do_ref = out->do_ref; j.dbg_area = out->search_area;
    unsigned long long totals[5] = {0, 0, 0, 0, 0};
    j.ops_total = totals;
    j.tf_best_sad = out->tf_best_sad; j.tf_best_mv = out->tf_best_mv; j.tf_hme_centre = out->tf_hme_centre;
    j.tf_hme_sad = out->tf_hme_sad; j.tf_early_exit = out->tf_early_exit;

    const bool     sub = params->me_search_method == ME_GPU_SUB_SAD_SEARCH;
    unsigned long long ops = 0;
    auto ctx_of = [&](int b) {
        CtrlCtx c;
        c.j = &j; c.p = &j.p; c.s = &state[b]; c.t = &tasks[(size_t)b * ME_SLOTS]; c.r = &results[(size_t)b * ME_SLOTS];
        c.f = &fs[(size_t)b * ME_NREF]; c.keys = &keys[(size_t)b * ME_NREF * ME_NPU];
        c.org_x = (b % j.g.b64_cols) * 64; c.org_y = (b / j.g.b64_cols) * 64;
        c.b64_w = (uint32_t)MC_MIN(64, j.g.width - c.org_x); c.b64_h = (uint32_t)MC_MIN(64, j.g.height - c.org_y);
        c.pic_w = j.g.width; c.pic_h = j.g.height;
        c.only = -1; c.lane = 0; c.nlanes = 1;
        c.fs_which = 0;
        return c;
    };
    /* the same stage sequence the engine runs (me_engine.cu, k_hme): with enable_me_sr_adjustment == 2 the integer-search
       stages are taken twice, reference (0, 0) first and the others after its search has finished */
    const bool two_pass = mc_fs_two_pass(params) != 0;
    const int  nseq = ST_COUNT + (two_pass ? 2 : 0);
    for (int seq = 0; seq < nseq; seq++) {
        const int stage = seq < ST_COUNT ? seq : ST_PROBE + (seq - ST_COUNT);
        const int fs_which = two_pass ? (seq < ST_COUNT ? 1 : 2) : 0;
        for (int b = 0; b < nb; b++) {
            CtrlCtx c = ctx_of(b);
            c.fs_which = fs_which;
            if (stage == ST_INIT)
                mc_init_keys(c.keys, 0, 1);
            mc_run_stage(&c, stage);
        }
        if (getenv("ME_EMUL_DUMP")) { /* task-shape histogram of this stage (kernel tuning aid) */
            std::map<std::string, int> hist;
            char                       key[128];
            if (stage <= ST_CENTRE) {
                for (size_t i = 0; i < tasks.size(); i++)
                    if (tasks[i].valid) {
                        const SearchTask &t = tasks[i];
                        snprintf(key, sizeof key, "L%d bw%-2d rows%-2d step%d skip%d sa %3dx%-3d", t.level, t.bw, t.rows, t.row_step,
                                 t.skip_line, t.sa_w, t.sa_h);
                        hist[key]++;
                    }
            } else {
                for (size_t i = 0; i < fs.size(); i++)
                    if (fs[i].valid) {
                        snprintf(key, sizeof key, "FS probe%d %3dx%-3d", fs[i].probe, fs[i].w, fs[i].h);
                        hist[key]++;
                    }
            }
            for (auto &kv : hist) fprintf(stderr, "stage %d  %-48s x %d\n", stage, kv.first.c_str(), kv.second);
        }
        if (stage <= ST_CENTRE) {
            for (size_t i = 0; i < tasks.size(); i++)
                if (tasks[i].valid)
                    exec_search(tasks[i], src, slot_pic[tasks[i].ref_slot], results[i]);
        } else {
            for (int b = 0; b < nb; b++)
                for (int k = 0; k < ME_NREF; k++) {
                    const FsTask &t = fs[(size_t)b * ME_NREF + k];
                    if (t.valid)
                        exec_fs(t, src, slot_pic[t.ref_slot], (b % j.g.b64_cols) * 64, (b / j.g.b64_cols) * 64, sub,
                                &keys[((size_t)b * ME_NREF + k) * ME_NPU]);
                }
        }
    }
    /* epilogue, in the order k_epilogue runs it */
    std::vector<uint32_t> bs((size_t)ME_NREF * ME_NPU), bm((size_t)ME_NREF * ME_NPU), md(ME_NPU);
    for (int b = 0; b < nb; b++) {
        CtrlCtx c = ctx_of(b);
        EpiCtx  e;
        e.j = &j; e.p = &j.p; e.s = &state[b]; e.keys = c.keys; e.b = (uint32_t)b; e.b64_w = c.b64_w; e.b64_h = c.b64_h;
        e.best_sad = (uint32_t(*)[ME_NPU])bs.data(); e.best_mv = (uint32_t(*)[ME_NPU])bm.data(); e.me_distortion = md.data();
        for (int k = 0; k < ME_NREF; k++)
            for (int n = 0; n < ME_NPU; n++) me_epi_decode(&e, k, n);
        if (j.p.me_type == ME_GPU_ME_MCTF) {
            for (int n = 0; n < ME_NPU; n++) me_epi_tf_pu(&e, n);
            me_epi_tf_block(&e);
            me_epi_taps(&e);
            if (j.dbg_best_sad) memcpy(j.dbg_best_sad + (size_t)b * ME_NREF * ME_NPU, bs.data(), bs.size() * 4);
            if (j.dbg_best_mv) memcpy(j.dbg_best_mv + (size_t)b * ME_NREF * ME_NPU, bm.data(), bm.size() * 4);
            ops += state[b].ops;
            continue;
        }
        if (j.p.enable_hme_flag && j.p.enable_me_hme_ref_pruning)
            me_epi_prune(&e);
        for (int n = 0; n < j.p.max_number_of_pus_per_sb; n++) me_epi_candidates(&e, n);
        me_epi_distortion(&e);
        me_epi_gm(&e);
        me_epi_taps(&e);
        if (j.dbg_best_sad) memcpy(j.dbg_best_sad + (size_t)b * ME_NREF * ME_NPU, bs.data(), bs.size() * 4);
        if (j.dbg_best_mv) memcpy(j.dbg_best_mv + (size_t)b * ME_NREF * ME_NPU, bm.data(), bm.size() * 4);
        ops += state[b].ops;
        out->me_64x64_distortion[b] = dist[0 * nb + b]; out->me_32x32_distortion[b] = dist[1 * nb + b];
        out->me_16x16_distortion[b] = dist[2 * nb + b]; out->me_8x8_distortion[b] = dist[3 * nb + b];
        out->me_8x8_cost_variance[b] = dist[4 * nb + b]; out->rc_me_distortion[b] = dist[5 * nb + b];
        out->stationary_block_present_sb[b] = flags[0 * nb + b]; out->rc_me_allow_gm[b] = flags[1 * nb + b];
    }
    out->b64_count = (uint32_t)nb;
    out->pu_count  = j.pu_count;
    out->tf_tot_horz_blks = (uint32_t)totals[2];
    out->tf_tot_vert_blks = (uint32_t)totals[3];
    if (ops_out)
        *ops_out = ops;
    return ME_GPU_OK;
}
