"""The BASELINE.json configurations as concrete workloads, shared by bench.py, the golden-hash generator
(tests/golden/make_picture_hashes.py) and the whole-picture GPU parity tests.

A workload is (size, preset, forced reference counts); one step = the 16 pictures of the second mini-GOP
(POC 17..32 of a 33-picture sequence) in decode order. `picture_hashes(out)` is the digest of a picture's product
outputs (planes.PRODUCT_FIELDS, the arrays the encoder consumes) that the goldens under tests/golden/ hold for the
reference's own AVX2 path on the same synthetic inputs.
"""
import hashlib
import json
import os

from me_b200 import planes, synth

# `refs` None = what the preset's mrp level gives (4 + 3 at presets <= 5).
# `sets` = resident copies of the sequence bench.py uses in rotation (inputs larger than the 126 MB L2 either way);
# `cpu_every` = the CPU arms time every n-th chunk of 6 b64s of each picture (a bounded sample, scaled back up).
WORKLOADS = {
    "1080p-m4": dict(w=1920, h=1080, preset=4, refs=(2, 2), sets=3, cpu_every=1, baseline_config=1,
                     text="1080p 8-bit preset 4, mini-GOP 16, 2 refs/list, default HME + integer search areas"),
    "1080p-m8": dict(w=1920, h=1080, preset=8, refs=(1, 1), sets=3, cpu_every=1, baseline_config=0,
                     text="1080p 8-bit preset 8, mini-GOP 16, 1 ref/list (configs[3] = this workload, one stream per GPU)"),
    "4k-m2": dict(w=3840, h=2160, preset=2, refs=None, sets=2, cpu_every=8, baseline_config=2,
                  text="4K preset 2 (8-bit ME planes: the MSB plane a 10-bit source is split into; synthetic 8-bit here), "
                       "mini-GOP 16, 4 + 3 refs, search areas up to 128x128 (x2 by MV)"),
    "4k-m0": dict(w=3840, h=2160, preset=0, refs=None, sets=2, cpu_every=48, baseline_config=4,
                  text="4K 8-bit preset 0, mini-GOP 16, 4 + 3 refs, HME-L0 up to 480x480, integer search up to 256x256 (x2 by MV)"),
}
NPIC = 33
# decode order of mini-GOP 2 (POC 17..32)
STEP_POCS = [32 if g == 0 else 16 + g for g in synth.RA5_DECODE_ORDER]
HASH_FILE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "picture_hashes.json")


LISTS_FILE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "encoder_lists.json")
_LISTS = None


def lists_for(name, poc):
    """(temporal layer, list0 pocs, list1 pocs) of picture `poc` of workload `name`: the lists the ENCODER's own prediction
    structure gives at the workload's preset (tests/golden/encoder_lists.json, frozen from a run of the encoder), cut to
    the workload's forced reference counts where BASELINE.json states them (configs[1]: "2 refs/list")."""
    global _LISTS
    if _LISTS is None:
        _LISTS = json.load(open(LISTS_FILE))["lists"]
    wl = WORKLOADS[name]
    tl, l0, l1 = _LISTS[str(wl["preset"])][str(poc)]
    if wl["refs"]:
        l0, l1 = l0[:wl["refs"][0]], l1[:wl["refs"][1]]
    return tl, list(l0), list(l1)


def derive_for(name, poc):
    """(cfg, MeGpuParams, list0 pocs, list1 pocs) of picture `poc` of workload `name`"""
    from helpers import derive
    wl = WORKLOADS[name]
    return derive(wl["preset"], wl["w"], wl["h"], poc, NPIC, force_refs=wl["refs"], lists=lists_for(name, poc))


def picture_hashes(out):
    """{field: first 16 hex digits of sha256 over the array bytes} for the product fields of an Outputs"""
    return {f: hashlib.sha256(out.arrays[f].tobytes()).hexdigest()[:16] for f in planes.PRODUCT_FIELDS}


def load_hashes():
    """{"<workload>/<kind>/<poc>": {field: digest}} frozen from the reference (AVX2 path), or {} when absent"""
    try:
        return json.load(open(HASH_FILE))["pictures"]
    except Exception:
        return {}


def hash_key(name, kind, poc):
    return f"{name}/{kind}/{poc}"


# What enc_handle.c's static set_mrp_ctrl() (Source/Lib/Globals/enc_handle.c:3397-3560, mrp levels 2 / 5 / 10) gives the
# workloads' presets - or what their forced reference counts stand for: (max refs list 0, max refs list 1,
# mrp_ctrls.safe_limit_nref, mrp_ctrls.safe_limit_zz_th, mrp_ctrls.only_l_bwd). The reference arm needs them to call the
# reference's own svt_aom_sig_deriv_me without going through the product's host library; tests/test_bench_contract.py
# checks that the two routes give the same MeGpuParams, byte for byte.
MRP = {"1080p-m4": (2, 2, 0, 0, 1), "1080p-m8": (1, 1, 2, 0, 1), "4k-m2": (4, 3, 0, 0, 1), "4k-m0": (4, 3, 0, 0, 1)}


def reference_params(name, poc, ref):
    """(MeGpuParams, list0 pocs, list1 pocs) of picture `poc` of workload `name`, controls by the REFERENCE's own
    svt_aom_sig_deriv_me (`ref`: checkers.Reference); no product library involved."""
    from me_b200 import abi
    wl = WORKLOADS[name]
    l0max, l1max, nref, zz, only_l_bwd = MRP[name]
    tl, l0, l1 = lists_for(name, poc)
    cfg = abi.MeGpuEncoderConfig(enc_mode=wl["preset"], qp=35, hierarchical_levels=4, width=wl["w"], height=wl["h"],
                                 picture_number=poc, temporal_layer_index=tl, is_ref=1 if tl < 4 else 0)
    p = ref.derive_params(cfg, nref, zz)
    p.only_l_bwd = only_l_bwd
    p.max_l0, p.max_refs = l0max, l0max + l1max
    p.max_cand = l0max + l1max + l0max * l1max + (l0max - 1) + (1 if l1max == 3 else 0)  # pcs.c:91-96
    p.num_of_list_to_search = 2 if l1 else 1
    p.num_of_ref_pic_to_search[0], p.num_of_ref_pic_to_search[1] = len(l0), len(l1)
    return p, l0, l1
