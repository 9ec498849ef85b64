"""Freezes whole-picture golden digests from the REFERENCE ITSELF for the BASELINE.json workloads.

    python tests/golden/make_picture_hashes.py [workload ...]

For every picture of the timed step (POC 17..32) of every workload in tests/workloads.py, on the "motion" and "iid"
sequences: the reference's own svt_aom_motion_estimation_b64 with its AVX2 kernels (oracle/_ref/libme_ref.so, built in
place from /root/reference) over ALL b64s of the picture -> sha256 per product array -> tests/golden/picture_hashes.json.
Before a picture's digests are trusted, a sample of 30 b64s spread over it is checked three ways:
reference AVX2 == reference C == the plain-C oracle (libme_oracle.so), every product array.
Runs in the build container only (needs /root/reference); the JSON travels to the GPU box.
"""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "svt-av1-psyex_b200", "python"))
sys.path.insert(0, os.path.join(HERE, ".."))

import workloads  # noqa: E402
from helpers import PictureSet  # noqa: E402
from me_b200 import planes  # noqa: E402
from checkers import Oracle, Reference  # noqa: E402

SAMPLE_B64 = 30


def main():
    names = sys.argv[1:] or list(workloads.WORKLOADS)
    ref, oracle = Reference(), Oracle()
    pool = ThreadPoolExecutor(os.cpu_count() or 1)
    try:
        doc = json.load(open(workloads.HASH_FILE))
    except Exception:
        doc = {"source": "reference svt_aom_motion_estimation_b64 + AVX2 kernels (oracle/_ref/libme_ref.so), whole pictures; "
                         "30-b64 samples cross-checked against the reference's C kernels and the plain-C oracle",
               "pictures": {}}
    for name in names:
        wl = workloads.WORKLOADS[name]
        w, h = wl["w"], wl["h"]
        nb = ((w + 63) // 64) * ((h + 63) // 64)
        for kind in ("motion", "iid"):
            ps = PictureSet(w, h, workloads.NPIC, kind, ref)
            for poc in workloads.STEP_POCS:
                t0 = time.time()
                cfg, p, l0, l1 = workloads.derive_for(name, poc)
                refs = [[ps.get(k) for k in l0], [ps.get(k) for k in l1]]
                full = ref.me_picture(ps.get(poc), refs, p, simd=1, debug=False)
                # three-way check on a sample of b64s spread over the picture (first, last, every (nb // 30)-th); at 4K
                # (seconds of scalar C per b64) on the base-layer, one layer-2 and one top-layer picture only
                step = max(1, nb // SAMPLE_B64)
                sample = list(range(0, nb, step))[:SAMPLE_B64 - 1] + [nb - 1]
                if w > 1920 and poc not in (32, 20, 17):
                    sample = []

                def check(first):
                    c_out = ref.me_picture(ps.get(poc), refs, p, first, 1, simd=0, debug=False)
                    o_out = oracle.me_picture(ps.get(poc), refs, p, first, 1, debug=False)
                    for f in planes.PRODUCT_FIELDS:
                        a = full.arrays[f][first]
                        assert np.array_equal(a, c_out.arrays[f][first]), (name, kind, poc, first, f, "avx2 != C")
                        assert np.array_equal(a, o_out.arrays[f][first]), (name, kind, poc, first, f, "reference != oracle")

                list(pool.map(check, sample))
                doc["pictures"][workloads.hash_key(name, kind, poc)] = workloads.picture_hashes(full)
                print(f"{name} {kind} poc {poc}: {nb} b64, refs {l0} {l1}, {len(sample)} b64 cross-checked, {time.time() - t0:.1f} s", flush=True)
            json.dump(doc, open(workloads.HASH_FILE, "w"), indent=0, sort_keys=True)


if __name__ == "__main__":
    main()
