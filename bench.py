#!/usr/bin/env python
"""bench.py - open-loop ME frames/s of the B200 engine (and of the reference's CPU path beside it).

Workload (BASELINE.json configs[1]): 1920x1080 8-bit, preset 4, mini-GOP 16 (5-layer random access),
2 references per list, default HME + integer search areas, synthetic "motion" sequence.
One step = open-loop ME of the 16 pictures of one mini-GOP (POC 17..32 of a 33-picture sequence), i.e. the
whole hot path: pyramid, pre-HME/HME, integer full search, candidate construction, results in host memory.

  value  frames/s with the pictures (and pyramids) already resident in HBM; results still return to the host
  e2e    frames/s through the C-ABI with host buffers: every step uploads its 16 source planes from pinned
         host memory (pyramids built on the device), runs ME, downloads the MeSbResults images, releases
  roofline    dominant kernel vs the measured VABSDIFF4 issue rate (integer-SAD roofline)
  cpu_baseline  the reference's own svt_aom_motion_estimation_b64 + AVX2 kernels (oracle/_ref) on the host cores

Usage: python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--kind motion|iid]
                       [--workload 1080p-m4|1080p-m8|4k-m2|4k-m0]   (the other BASELINE.json configs)
Under torchrun every rank runs the same workload on its own GPU (weak scaling, no collective on the path).
"""
import argparse
import collections
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(REPO, "svt-av1-psyex_b200", "python"))
sys.path.insert(0, os.path.join(REPO, "tests"))

from me_b200 import abi, planes, synth  # noqa: E402

import workloads  # noqa: E402  (tests/workloads.py: the BASELINE.json configs as concrete workloads + golden digests)

WORKLOADS = workloads.WORKLOADS
WORKLOAD = "1080p-m4"
W, H, PRESET, NPIC, FORCE_REFS = 1920, 1080, 4, 33, (2, 2)
NSETS, SET_STRIDE = 3, 1000  # copies of the sequence in the pool (distinct picture numbers: poc + 1000 * set)
CPU_EVERY = 1


def set_workload(name):
    global WORKLOAD, W, H, PRESET, FORCE_REFS, NSETS, CPU_EVERY
    wl = WORKLOADS[name]
    WORKLOAD, W, H, PRESET, FORCE_REFS, NSETS, CPU_EVERY = name, wl["w"], wl["h"], wl["preset"], wl["refs"], wl["sets"], wl["cpu_every"]


def b64_count():
    return ((W + 63) // 64) * ((H + 63) // 64)


STEP_POCS = workloads.STEP_POCS  # decode order of mini-GOP 2 (POC 17..32)


def derive(poc):
    if WORKLOAD in WORKLOADS:  # the encoder's own reference lists at the workload's preset (tests/golden/encoder_lists.json)
        return workloads.derive_for(WORKLOAD, poc)
    from helpers import derive as _d  # experiments on other sizes (tools/profile_step.py w:h:preset:r0:r1): synth's table
    return _d(PRESET, W, H, poc, NPIC, force_refs=FORCE_REFS)


def bind_to_gpu_numa(index):
    """Pin this process (and the pinned buffers it is about to allocate: first touch) to the CPUs of the NUMA node the
    GPU hangs off, as read from sysfs. One process per GPU: without this the 8 ranks' uploads cross the socket link."""
    try:
        import torch
        bdf = torch.cuda.get_device_properties(index).pci_bus_id
        dom = torch.cuda.get_device_properties(index).pci_domain_id
        dev = torch.cuda.get_device_properties(index).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bdf:02x}:{dev:02x}.0/local_cpulist"
        cpus = set()
        for part in open(path).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return sorted(cpus)
    except Exception:
        pass
    return None


def rank_info():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


class ClockSampler(threading.Thread):
    """nvidia-smi style clock / throttle-reason samples during the timed region (pynvml)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.stop_flag, self.max_mhz = [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, n in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.25)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def make_pictures(kind, pinned, padded=True):
    """full-resolution luma planes of the 33 pictures (with or without the 68-pixel border: the engine generates it on the
    device for unpadded input), in pinned host memory when torch+CUDA are there"""
    seq = synth.Sequence(W, H, NPIC, kind=kind)
    rows, stride = (H + 136, W + 136) if padded else (H, W)
    out = {}
    for poc in range(NPIC):
        if pinned:
            import torch
            t = torch.empty((rows, stride), dtype=torch.uint8).pin_memory()
            a = t.numpy()
            out[poc] = (a, t)
        else:
            out[poc] = (np.empty((rows, stride), dtype=np.uint8), None)
        out[poc][0][...] = np.pad(seq.luma(poc), 68, mode="edge") if padded else seq.luma(poc)
    return out


def pinned_outputs(params, eng):
    """MeGpuOutputs (product fields only) carved out of ONE pinned buffer at the engine's slab offsets, so that
    the results of a picture come back with a single device->host copy"""
    import torch
    o = planes.Outputs(W, H, params, debug=False)
    offs, total = eng.output_layout(W, H, params)
    slab = torch.zeros(total, dtype=torch.uint8).pin_memory()
    base = slab.numpy()
    for i, name in enumerate(planes.PRODUCT_FIELDS):
        arr = o.arrays[name]
        a = base[offs[i]:offs[i] + arr.nbytes].view(arr.dtype).reshape(arr.shape)
        o.arrays[name] = a
        setattr(o.c, name, C.cast(a.ctypes.data, dict(abi.MeGpuOutputs._fields_)[name]))
    o._keep = slab
    return o


# --------------------------------------------------------------------------------------------------
# reference CPU arm (oracle/_ref: the reference's own sources compiled in place)
# --------------------------------------------------------------------------------------------------
class CpuArm:
    """The reference's own open-loop ME (svt_aom_motion_estimation_b64 + its SIMD kernels, oracle/_ref = the reference's
    sources compiled in place) over the step's pictures, on host threads, in C: every picture cut into the encoder's ME
    segments, all items in one queue, a MeContext per thread (oracle/ref_harness.c: ref_me_step_mt). Controls come from
    the reference's own svt_aom_sig_deriv_me (workloads.reference_params): no product library is loaded on this arm."""

    SIMD = {"avx2": 1, "avx512": 2, "c": 0}

    def __init__(self, kind, pocs):
        import checkers
        self.ref = checkers.Reference()
        self.pocs = pocs
        seq = synth.Sequence(W, H, NPIC, kind=kind)
        pics = self.pics = {}

        def get(k):
            if k not in pics:
                p = planes.Picture(W, H, k)
                p.set_luma(seq.luma(k))
                self.ref.build_pyramid(p)
                pics[k] = p
            return pics[k]

        self.jobs = []
        for poc in pocs:
            p, l0, l1 = workloads.reference_params(WORKLOAD, poc, self.ref)
            self.jobs.append((get(poc), [[get(k) for k in l0], [get(k) for k in l1]], p, planes.Outputs(W, H, p, debug=False)))
        self.b64_done = 0

    def step(self, threads, simd="avx2", every=None):
        """ME of (a sample of) every picture of the step once; returns seconds"""
        t0 = time.perf_counter()
        self.b64_done = self.ref.me_step_mt(self.jobs, threads, self.SIMD[simd], CPU_EVERY if every is None else every)
        return time.perf_counter() - t0

    def pictures(self):
        """whole pictures' worth of b64s the last step() covered"""
        return self.b64_done / b64_count()

    def measure(self, threads, simd, budget_s, max_reps=20):
        self.step(threads, simd)
        reps, dt, pics = 0, 0.0, 0.0
        while dt < budget_s and reps < max_reps:
            dt += self.step(threads, simd)
            pics += self.pictures()
            reps += 1
        return {"value": pics / dt, "unit": "frames/s", "cores": threads, "kind": "reference", "simd": simd,
                "sample": f"{self.sample_text()} x {reps} repetitions, {dt:.1f} s wall on {threads} threads"}

    def parity(self, kind):
        """the arm's own outputs against the frozen digests (only when whole pictures were searched)"""
        gold = workloads.load_hashes()
        keys = [workloads.hash_key(WORKLOAD, kind, poc) for poc in self.pocs]
        if CPU_EVERY != 1 or not all(k in gold for k in keys):
            return None
        return all(workloads.picture_hashes(j[3]) == gold[k] for j, k in zip(self.jobs, keys))

    def sample_text(self):
        return (f"all {b64_count()} b64 of each of the step's {len(self.pocs)} pictures" if CPU_EVERY == 1 else
                f"every {CPU_EVERY}th ME segment (of 48) of each of the step's {len(self.pocs)} pictures "
                f"({self.b64_done} of {len(self.pocs) * b64_count()} b64, frames/s scaled by the b64 count)")


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def physical_cores():
    """distinct (physical id, core id) pairs of /proc/cpuinfo; None when the VM does not tell"""
    try:
        seen, phys, core = set(), None, None
        for line in open("/proc/cpuinfo"):
            if line.startswith("physical id"):
                phys = line.split(":")[1].strip()
            elif line.startswith("core id"):
                core = line.split(":")[1].strip()
            elif not line.strip() and phys is not None and core is not None:
                seen.add((phys, core))
                phys = core = None
        return len(seen) or None
    except Exception:
        return None


def host_info():
    return {"cpu": cpu_model(), "logical_cpus": os.cpu_count(), "physical_cores": physical_cores(),
            "note": "a small host for a baseline: the GPU box's VM exposes 16 vCPUs of an unnamed Xeon; every ratio against "
                    "this arm is a ratio against THESE cores"}


def reference_main(args):
    rank, _, world = rank_info()
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    arm = CpuArm(args.kind, STEP_POCS)
    for _ in range(args.warmup):
        arm.step(threads)
    t_total, pics = 0.0, 0.0
    for _ in range(args.steps):
        t_total += arm.step(threads)
        pics += arm.pictures()
    value = pics / t_total
    sample = f"{arm.sample_text()}, per step; {threads} threads"
    others = [arm.measure(threads, "avx512", 3.0), arm.measure(1, "avx2", 3.0, max_reps=3)]
    line = {
        "impl": "reference", "metric": "open-loop ME frames/sec", "value": value, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args.kind),
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": threads, "kind": "reference", "sample": sample,
                         "simd": "avx2 (svt_*_avx2 kernels behind the reference's RTCD pointers)", "other_arms": others},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "parity": {"equal": arm.parity(args.kind), "against": "tests/golden/picture_hashes.json"}, "host": host_info(),
    }
    print(json.dumps(line))
    return 0


def workload_config(kind):
    wl = WORKLOADS[WORKLOAD]
    set_mb = NPIC * (W + 136 + 120) * (H + 136) * 21 // 16 / 1e6  # three planes of a picture ~ 21/16 of the padded full plane
    return {"workload": f"{wl['text']} (BASELINE.json configs[{wl['baseline_config']}])", "name": WORKLOAD, "width": W,
            "height": H, "preset": PRESET, "refs": list(FORCE_REFS) if FORCE_REFS else "preset default (4 + 3)",
            "pictures_per_step": 16, "sequence": f"synthetic {kind}, seed 1234, 33 pictures", "qp": 35,
            "l2": f"inputs larger than L2: {NSETS} resident copies of the sequence ({NSETS} x {set_mb:.0f} MB of planes) used "
                  f"in rotation, step k reads copy k % {NSETS}"}


# --------------------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------------------
def init_dist(torch, dist, local_rank, world):
    """NCCL for the timer's all_reduce(MAX) (the contract's launch), gloo for the sharded mode's control plane (which
    picture sits in which pool slot: a few bytes per picture). Neither carries a pixel."""
    if world <= 1:
        return None
    # NCCL prints its version banner on stdout when the box sets NCCL_DEBUG=VERSION: stdout carries ONE JSON line,
    # so fd 1 points at stderr while the communicator comes up (first collective included)
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    try:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
        torch.cuda.synchronize()
        ctl = dist.new_group(backend="gloo")
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)
    return ctl


def measure(args, kind, mode, steps, warmup, torch, dist, ctl, with_cpu):
    """One measurement of the current workload (set_workload) on this rank's GPU.
    mode "replicas": every rank runs the whole step on its own copy of the stream (weak scaling; configs[3]).
    mode "sharded":  ONE stream; the pictures of the step are dealt over the ranks by measured cost (shard.owners); a picture
                     is uploaded by its owner only and reaches the ranks that reference it GPU to GPU (CUDA IPC mapping of
                     the owner's pool, device-to-device copy over NVLink); strong scaling; no collective on the data path."""
    from me_b200 import shard
    from me_b200.engine import Engine
    global NSETS
    rank, local_rank, world = rank_info()
    sharded = mode == "sharded" and world > 1
    if sharded:
        NSETS = max(NSETS, 3)  # the pipelined e2e step below needs a set to rest for two steps before its slots are reused

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    pics = make_pictures(kind, pinned=True, padded=False)
    all_jobs = {poc: derive(poc) for poc in STEP_POCS}

    # Sharded: pictures are dealt by COST, the device time each one took the last time its position of the mini-GOP was
    # searched (what an encoder-side scheduler knows from the previous mini-GOP; here: one untimed pass on rank 0, every
    # rank gets the same table). A base-layer picture with 3 references and small scaled areas and a layer-3 picture with
    # 7 differ several times in work; dealing the decode order round-robin leaves GPUs idle.
    cost_ms = None
    if sharded:
        box = [None]
        if rank == 0:
            cal = Engine(W, H, 4, 3, pictures_in_flight=NPIC + 4, devices=[local_rank])
            for poc in range(NPIC):
                a, _ = pics[poc]
                cal.register_picture_unpadded_ptr(poc, a.ctypes.data, a.strides[0], W, H)
            cal.set_profiling(True)
            box[0] = {}
            for poc in STEP_POCS:
                _, p, l0, l1 = all_jobs[poc]
                out = planes.Outputs(W, H, p, debug=False)
                cal.submit(poc, p, [l0, l1])
                cal.wait(poc, out)
                box[0][poc] = cal.timings().total_ms
            cal.close()
        dist.broadcast_object_list(box, src=0, group=ctl)
        cost_ms = box[0]
    # the decode order starts with the key frame: mini-GOPs are [1:17], [17:33] of it; mini-GOP 1 and the key frame are only read
    order = synth.decode_order(NPIC)
    owner = ({**shard.owners(order[:17], world), **shard.owners(order[17:], world, cost=lambda poc: cost_ms[poc])} if sharded
             else {poc: rank for poc in range(NPIC)})
    my_pocs = [poc for poc in STEP_POCS if owner[poc] == rank]
    jobs = {poc: all_jobs[poc] for poc in my_pocs}
    needed = shard.pyramids_needed(my_pocs, lambda poc: jobs[poc][2:])  # what this rank's pictures read
    eng = Engine(W, H, 4, 3, pictures_in_flight=NSETS * NPIC + 4, devices=[local_rank])
    # pinned result buffers: one per picture of a step; a rank that has few pictures per step (sharded, N = 8: two) gets one
    # per (picture, resident set) so that as many pictures can be in flight as with a whole step per rank
    osets = NSETS if sharded and len(my_pocs) < 8 else 1
    outs_all = {(poc, o): pinned_outputs(jobs[poc][1], eng) for poc in my_pocs for o in range(osets)}
    outs = {poc: outs_all[(poc, 0)] for poc in my_pocs}

    def out_of(poc, st):
        return outs_all[(poc, st % osets)]
    state = {"k": 0}
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # profiling pass only
    num = lambda poc, st: poc + SET_STRIDE * st  # noqa: E731  picture numbers of a set are offset by a constant: distances are unchanged

    if sharded:
        handles = [None] * world
        dist.all_gather_object(handles, eng.ipc_export_pool(), group=ctl)
        for r, (h, sb) in enumerate(handles):
            if r != rank:
                eng.ipc_open_peer(r, h, sb)

    def reg(poc, st):
        a, _ = pics[poc]
        eng.register_picture_unpadded_ptr(num(poc, st), a.ctypes.data, a.strides[0], W, H)

    def exchange(pocs, st):
        """owners publish (slot, event) of the pictures they have just registered; everybody pulls the ones it reads"""
        mine = {poc: eng.ipc_picture(num(poc, st)) for poc in pocs if owner[poc] == rank}
        infos = [None] * world
        dist.all_gather_object(infos, mine, group=ctl)
        for poc in pocs:
            if owner[poc] != rank and poc in needed:
                slot, evh = infos[owner[poc]][poc]
                eng.register_picture_from_peer(num(poc, st), owner[poc], slot, evh, W, H)

    def held(poc):
        return owner[poc] == rank or poc in needed

    def submit(poc, st):
        _, p, l0, l1 = jobs[poc]
        q = abi.MeGpuParams.from_buffer_copy(p)
        q.picture_number = num(poc, st)
        eng.submit_out(num(poc, st), q, [[num(k, st) for k in l0], [num(k, st) for k in l1]], out_of(poc, st))

    # Pictures stream through the engine the way they stream through the encoder's ME stage: up to WINDOW pictures
    # are in flight, the oldest one is waited for (its results copied back) before the next is submitted, and a step
    # does not drain the pipeline; drain() (inside the timed region) waits for whatever is left.
    # < 16: the engine's job contexts, and the pinned result buffer of a POC is reused every 16 pictures
    # ... and a (picture, set) has one result buffer and one identity: never two of the same in flight
    WINDOW = min(int(os.environ.get("BENCH_WINDOW", 12)), len(my_pocs) * osets)
    in_flight = collections.deque()

    def retire_oldest():
        poc, st, _ = in_flight.popleft()
        eng.wait(num(poc, st), out_of(poc, st))

    def enqueue(poc, st):
        while len(in_flight) >= WINDOW:
            retire_oldest()
        submit(poc, st)
        in_flight.append((poc, st, state["k"]))

    def retire_before(k):
        while in_flight and in_flight[0][2] < k:
            retire_oldest()

    def drain():
        while in_flight:
            retire_oldest()

    def step_resident():
        st = state["k"] % NSETS
        state["k"] += 1
        for poc in my_pocs:
            enqueue(poc, st)

    uploaded = set()

    def step_e2e():
        st = state["k"] % NSETS
        state["k"] += 1
        if st in uploaded:  # the set was used NSETS steps ago: its jobs are long done, here and (a rendezvous per step) on the peers
            for poc in STEP_POCS:
                if held(poc):
                    eng.release_picture(num(poc, st))
        uploaded.add(st)
        if sharded:
            # Three resident sets: the slots released above were filled three steps ago. Every rank retires its jobs of two
            # steps ago before this step's rendezvous (the ticket exchange), so when a rank passes the NEXT rendezvous nobody
            # reads or pulls those slots any more - while the jobs of the previous step still run under this step's uploads
            # and pulls.
            retire_before(state["k"] - 1)  # tags are the counter after its increment: keeps only the previous step in flight
            for poc in my_pocs:
                reg(poc, st)
            exchange(STEP_POCS, st)
            for poc in my_pocs:
                enqueue(poc, st)
        else:
            for poc in my_pocs:
                while len(in_flight) >= WINDOW:
                    retire_oldest()
                reg(poc, st)
                enqueue(poc, st)

    def timed(step_fn, steps, warmup):
        for _ in range(warmup):
            step_fn()
        drain()
        barrier()
        sampler = ClockSampler(local_rank)
        if not os.environ.get("BENCH_NO_SAMPLER"):
            sampler.start()
        l0 = eng.timings().kernel_launches
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            step_fn()
        drain()
        barrier()
        dt = time.perf_counter() - t0
        sampler.stop_flag = True
        if sampler.is_alive():
            sampler.join()
        launches = eng.timings().kernel_launches - l0
        if world > 1:
            t = torch.tensor([dt, float(launches)], dtype=torch.float64, device="cuda")
            dist.all_reduce(t[:1], op=dist.ReduceOp.MAX)
            dist.all_reduce(t[1:], op=dist.ReduceOp.SUM)
            dt, launches = float(t[0].item()), int(t[1].item())
        return dt, launches, sampler.summary()

    # resident arm: every pyramid where it is read before the clock starts (sharded: uploaded by its owner, pulled by the rest)
    for st in range(NSETS):
        for poc in range(NPIC):
            if owner[poc] == rank:
                reg(poc, st)
        if sharded:
            exchange(list(range(NPIC)), st)
    peer_verified = None
    if sharded:
        pulled = [poc for poc in needed if owner[poc] != rank]
        if pulled:  # one replicated pyramid read back and compared with the one this device builds from the same luma
            poc = pulled[0]
            got = eng.read_pyramid(num(poc, 0), W, H)
            a, _ = pics[poc]
            eng.register_picture_unpadded_ptr(999999, a.ctypes.data, a.strides[0], W, H)
            own = eng.read_pyramid(999999, W, H)
            eng.release_picture(999999)
            peer_verified = all(np.array_equal(x, y) for x, y in zip(got, own))
    bytes0 = eng.ipc_bytes_pulled() if sharded else 0
    dt_res, launches, clocks = timed(step_resident, steps, warmup)

    # roofline pass: one picture at a time, every launch bracketed by events on its stream
    eng.set_profiling(True)
    prof = {"k_search_ms": 0.0, "k_fullsearch_ms": 0.0, "k_fullsearch_main_ms": 0.0, "k_ctrl_ms": 0.0, "k_epilogue_ms": 0.0,
            "total_ms": 0.0, "ops": 0, "ops_fs": 0, "ops_fs_inline": 0}
    for rep in range(2):
        for k in prof:
            prof[k] = 0 if k.startswith("ops") else 0.0
        for poc in my_pocs:
            flush_buf.zero_()
            torch.cuda.synchronize()
            submit(poc, rep)
            eng.wait(num(poc, rep), out_of(poc, rep))
            t = eng.timings()
            for k in ("k_search_ms", "k_fullsearch_ms", "k_fullsearch_main_ms", "k_ctrl_ms", "k_epilogue_ms", "total_ms"):
                prof[k] += getattr(t, k)
            prof["ops"] += t.byte_sad_ops
            prof["ops_fs"] += t.byte_sad_ops_fullsearch
            prof["ops_fs_inline"] += t.byte_sad_ops_fullsearch_inline
    eng.set_profiling(False)
    sad_peak = 4.0 * eng.sad_peak()  # byte-SADs / s

    # e2e arm: sources of the step uploaded every step (by their owner), pulled by the ranks that read them
    for st in range(NSETS):
        for poc in STEP_POCS:
            if held(poc):
                eng.release_picture(num(poc, st))
    barrier()
    bytes1 = eng.ipc_bytes_pulled() if sharded else 0
    dt_e2e, _, _ = timed(step_e2e, steps, warmup)
    bytes2 = eng.ipc_bytes_pulled() if sharded else 0

    # parity gate: the product arrays of this rank's pictures, as the e2e arm left them in host memory, against the
    # digests frozen from the reference's own AVX2 path on the same synthetic inputs (tests/golden/picture_hashes.json)
    gold = workloads.load_hashes()
    keys = {poc: workloads.hash_key(WORKLOAD, kind, poc) for poc in my_pocs}
    have = all(k in gold for k in keys.values())
    wrong = [poc for poc in my_pocs if have and workloads.picture_hashes(outs[poc]) != gold[keys[poc]]]
    checked = len(my_pocs) if have else 0
    nvlink = [bytes1 - bytes0, bytes2 - bytes1]
    d2h = sum(sum(a.nbytes for a in outs[p].arrays.values()) for p in my_pocs)  # this rank's result bytes per step
    d2h_all = d2h
    if world > 1:
        t = torch.tensor([len(wrong), checked, nvlink[0], nvlink[1], d2h], dtype=torch.int64, device="cuda")
        dist.all_reduce(t)
        n_wrong, checked, nvlink, d2h_all = int(t[0]), int(t[1]), [int(t[2]), int(t[3])], int(t[4])
    else:
        n_wrong = len(wrong)
    parity = {"pictures": checked, "equal": (n_wrong == 0) if checked else None, "mismatched_pocs_rank0": wrong,
              "against": "tests/golden/picture_hashes.json (reference svt_aom_motion_estimation_b64 + AVX2 kernels, all b64s, "
                         "every product array)" if checked else "no golden digests for this workload"}
    eng.close()

    n_frames = 16 * steps * (1 if sharded else world)
    value, e2e = n_frames / dt_res, n_frames / dt_e2e
    h2d = len(my_pocs) * W * H  # luma planes without their border: me_gpu_register_picture_unpadded pads on the device

    # byte-SADs where they are executed: the 8x8-variance probes and the integer searches of <= 256 positions run inside k_hme,
    # only the tiled areas in k_fullsearch
    ops_tiles = prof["ops_fs"] - prof["ops_fs_inline"]
    ops_hme = prof["ops"] - ops_tiles
    kernels = {
        "k_hme": {"ms_per_step": prof["k_search_ms"], "byte_sads_per_step": ops_hme},
        "k_fullsearch": {"ms_per_step": prof["k_fullsearch_ms"], "byte_sads_per_step": ops_tiles},
        "k_epilogue": {"ms_per_step": prof["k_epilogue_ms"]},
    }
    dom = "k_hme" if prof["k_search_ms"] >= prof["k_fullsearch_ms"] else "k_fullsearch"
    dom_ms, dom_ops = kernels[dom]["ms_per_step"], kernels[dom]["byte_sads_per_step"]
    achieved = dom_ops / (dom_ms * 1e-3) / 1e12 if dom_ms > 0 else 0.0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    # dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel: NOT measured by this run (it needs ncu);
    # read from profiles/ncu_traffic.json, which names the capture and the commit each figure comes from
    traffic, traffic_source = None, None
    try:
        e = json.load(open(os.path.join(REPO, "profiles", "ncu_traffic.json")))["entries"].get(f"{WORKLOAD}/{kind}/{dom}")
        if e:
            traffic, traffic_source = e["dram_bytes_per_launch"], f"profiles/{e['capture']} (commit {e['commit']})"
    except Exception:
        pass
    roofline = {
        "bound": "int_sad", "bound_note": "integer ALU (packed byte-SAD issue rate, SURVEY 8(d)): the path is neither HBM- nor "
        "tensor-bound; its HBM side is in traffic / hbm_gbs_achieved against hbm_peak_gbs", "kernel": dom, "achieved": achieved,
        "peak": sad_peak / 1e12, "unit": "Tbyte-SAD/s", "frac": achieved / (sad_peak / 1e12) if sad_peak else None,
        "traffic": traffic, "traffic_source": traffic_source,
        "traffic_unit": "DRAM bytes per launch (one picture); the three planes of a picture are 3.3 MB at 1080p, 12.4 MB at 4K",
        "hbm_gbs_achieved": (traffic * 16 / (dom_ms * 1e-3) / 1e9) if dom_ms > 0 and traffic else None,
        "peak_source": "VABSDIFF4.U8.ACC issue rate measured live by k_sad_peak on this GPU x 4 bytes",
        "launches_per_step": len(my_pocs),  # one k_hme, one k_fullsearch and one k_epilogue launch per picture (rank 0's pictures)
        "kernels": kernels, "serial_ms_per_step": prof["total_ms"],
        "whole_path_frac": (prof["ops"] / (prof["total_ms"] * 1e-3)) / sad_peak if sad_peak and prof["total_ms"] else None,
        "hbm_peak_gbs": peaks.get("hbm_gbs"),
    }
    cpu_baseline = None
    if rank == 0 and world == 1 and with_cpu:
        threads = os.cpu_count() or 1
        arm = CpuArm(kind, STEP_POCS)
        cpu_baseline = arm.measure(threads, "avx2", 4.0)
        cpu_baseline["other_arms"] = [arm.measure(threads, "avx512", 3.0), arm.measure(1, "avx2", 3.0, max_reps=3)]
        cpu_baseline["parity_equal"] = arm.parity(kind)
        cpu_baseline["host"] = host_info()
    cfg = workload_config(kind)
    cfg["mode"] = ("sharded: one stream, the step's pictures dealt over the ranks by measured cost (`dealing`); owners upload, "
                   "the others pull pyramids GPU to GPU (CUDA IPC + device-to-device copy over NVLink), no collective on the data path"
                   if sharded else "replicas: every rank runs the whole step on its own copy of the stream" if world > 1 else "one GPU")
    line = {
        "metric": "open-loop ME frames/sec", "value": value, "unit": "frames/s", "n_gpus": world, "steps": steps,
        "warmup": warmup, "ms_per_step": 1e3 * dt_res / steps, "higher_is_better": True,
        "scaling": "strong" if sharded else "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg,
        "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": 16 * W * H * (1 if sharded else world),
                "d2h_bytes_per_step": d2h_all, "ms_per_step": 1e3 * dt_e2e / steps,
                "h2d_gbs_per_rank": h2d / (dt_e2e / steps) / 1e9, "d2h_gbs_per_rank": d2h / (dt_e2e / steps) / 1e9,
                "note": "can exceed `value`: the resident arm rotates L2-busting copies of the sequence, this arm searches planes "
                        "it has just uploaded (warm in L2); uploads and result copies overlap the searches of other pictures"},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline, "parity": parity,
        "host": {"cpus": os.cpu_count()},
    }
    if sharded:
        line["nvlink_bytes_per_step"] = nvlink[1] / (steps + warmup)  # all ranks, e2e arm, warm-up steps counted on both sides
        line["peer_copy_verified"] = peer_verified
        line["pictures_per_rank_per_step"] = len(my_pocs)
        line["dealing"] = {"rule": "per mini-GOP, most expensive picture first to the least loaded rank (me_b200/shard.py owners)",
                           "cost": "device ms of the picture's ME, measured once on rank 0 before the clock starts",
                           "ms_per_rank": [round(sum(cost_ms[p] for p in STEP_POCS if owner[p] == r), 3) for r in range(world)]}
    return line


def gpu_main(args):
    import torch
    import torch.distributed as dist

    rank, local_rank, world = rank_info()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    numa_cpus = bind_to_gpu_numa(local_rank) if world > 1 else None
    ctl = init_dist(torch, dist, local_rank, world)
    mode = args.mode or ("sharded" if world > 1 and WORKLOAD.startswith("4k") else "replicas")
    line = measure(args, args.kind, mode, args.steps, args.warmup, torch, dist, ctl, with_cpu=not args.no_cpu)
    line["host"]["rank0_numa_cpus"] = len(numa_cpus) if numa_cpus else None
    ok = line["parity"]["equal"] is not False
    if world > 1 and mode == "replicas" and args.extra:
        # the multi-GPU path north_star describes, on BASELINE configs[2], in the same launch: one 4K stream sharded over the ranks
        main_wl = WORKLOAD
        set_workload("4k-m2")
        sh = measure(args, "motion", "sharded", max(3, args.steps), args.warmup, torch, dist, ctl, with_cpu=False)
        set_workload(main_wl)
        line["sharded_4k"] = {k: sh[k] for k in ("value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "scaling", "config", "e2e",
                                                  "parity", "nvlink_bytes_per_step", "peer_copy_verified", "pictures_per_rank_per_step", "dealing",
                                                  "clocks")}
        ok = ok and sh["parity"]["equal"] is not False
    extra = None
    if rank == 0 and world == 1 and args.extra and WORKLOAD == "1080p-m4" and args.kind == "motion":
        # the same measurement on the other configurations (worst-case "iid" input: no early exits, every search area at its
        # configured size - the cases where the integer full search dominates - and BASELINE configs[2]), each in a process of
        # its own after this one's engine is closed; the headline stays the workload above
        import subprocess
        extra = []
        for wl, kind in (("1080p-m4", "iid"), ("4k-m2", "motion"), ("4k-m0", "iid")):
            cmd = [sys.executable, os.path.abspath(__file__), "--workload", wl, "--kind", kind, "--steps", "10", "--warmup", "3",
                   "--no-cpu", "--no-extra"]
            try:
                r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
                d = json.loads(r.stdout.strip().splitlines()[-1])
                extra.append({"workload": wl, "sequence": kind, "value": d["value"], "e2e": d["e2e"]["value"], "unit": "frames/s",
                              "steps": d["steps"], "warmup": d["warmup"], "clocks": d["clocks"], "parity": d["parity"],
                              "roofline": {k: d["roofline"][k] for k in ("kernel", "achieved", "peak", "unit", "frac", "kernels")}})
            except Exception as e:  # reported, never silent
                extra.append({"workload": wl, "sequence": kind, "error": repr(e)[:200]})
    line["extra_workloads"] = extra
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0 if ok else 3  # results that differ from the reference's are not a benchmark


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--kind", default="motion", choices=["motion", "iid"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", dest="extra", action="store_false",
                    help="skip the runs of the other configurations (1080p-m4 iid, 4k-m2 motion, 4k-m0 iid) appended as extra_workloads at N=1")
    ap.add_argument("--mode", default=None, choices=["replicas", "sharded"],
                    help="N > 1: replicas = every rank its own copy of the stream (weak scaling; default at 1080p, configs[3]); sharded = "
                         "one stream, pictures dealt over the ranks, pyramids pulled GPU to GPU (strong scaling; default for the 4K workloads)")
    ap.add_argument("--workload", default="1080p-m4", choices=sorted(WORKLOADS),
                    help="BASELINE.json config to run (default: configs[1], the one the metric is quoted on at N=1)")
    args = ap.parse_args()
    set_workload(args.workload)
    if args.impl == "reference":
        return reference_main(args)
    return gpu_main(args)


if __name__ == "__main__":
    sys.exit(main())
