"""Several engines in one process (me_gpu_engine_create / _bind / _destroy): the reference library serves several encoder
instances per process (EbComponentType handles); the calls of a thread go to the engine the thread has bound. Two engines
of different picture sizes on the same GPU, used alternately from one thread and at the same time from two threads, each
checked against the oracle; the default engine is untouched by what happens on the second one."""
import ctypes as C
import threading

import pytest

from helpers import PictureSet, derive, used_refs
from me_b200 import abi, planes

pytestmark = pytest.mark.gpu


def _run(eng, ps, w, h, preset, poc, n):
    cfg, p, l0, l1 = derive(preset, w, h, poc, n, force_refs=(2, 2))
    pocs = sorted(set([poc] + l0 + l1))
    for k in pocs:
        eng.register_picture(k, ps.get(k).planes[0], w, h)
    got = eng.me_picture(poc, w, h, p, [l0, l1], debug=False)
    for k in pocs:
        eng.release_picture(k)
    return poc, p, l0, l1, got


def test_two_engines_in_one_process():
    from checkers import Oracle
    from me_b200.engine import Engine
    oracle = Oracle()
    lib = abi.load_gpu_lib()
    second = C.c_void_p()
    assert lib.me_gpu_engine_create(C.byref(second)) == 0 and second.value
    sets = {"a": PictureSet(320, 192, 17, "motion", oracle), "b": PictureSet(200, 136, 17, "motion", oracle)}
    lib.me_gpu_engine_bind(None)
    a = Engine(320, 192, 4, 3, pictures_in_flight=8)  # the default engine
    lib.me_gpu_engine_bind(second)
    b = Engine(200, 136, 4, 3, pictures_in_flight=8)  # me_gpu_init on the second one: no "already initialised"
    try:
        # a picture registered on one engine does not exist on the other
        b.register_picture(900, sets["b"].get(3).planes[0], 200, 136)
        lib.me_gpu_engine_bind(None)
        assert lib.me_gpu_release_picture(900) == -3  # ME_GPU_ERR_NOT_FOUND on the default engine
        lib.me_gpu_engine_bind(second)
        b.release_picture(900)
        # alternately from one thread
        results = {}
        for i, (name, eng, handle, w, h, preset) in enumerate((("a", a, None, 320, 192, 4), ("b", b, second, 200, 136, 8)) * 2):
            lib.me_gpu_engine_bind(handle)
            results[(name, i)] = _run(eng, sets[name], w, h, preset, 4 + (i // 2) * 4, 17)
        # at the same time from two threads, each bound to its own engine
        errors = []

        def worker(name, eng, handle, w, h, preset):
            try:
                lib.me_gpu_engine_bind(handle)
                for rep in range(3):
                    results[(name, "t", rep)] = _run(eng, sets[name], w, h, preset, 4 + 4 * (rep % 2), 17)
            except Exception as e:  # noqa: BLE001
                errors.append((name, repr(e)))

        ts = [threading.Thread(target=worker, args=("a", a, None, 320, 192, 4)),
              threading.Thread(target=worker, args=("b", b, second, 200, 136, 8))]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert not errors, errors
        for key, (poc, p, l0, l1, got) in results.items():
            ps = sets[key[0]]
            exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p, debug=False)
            bad = planes.compare_outputs(exp, got, fields=planes.PRODUCT_FIELDS, used_refs=used_refs(l0, l1))
            assert not bad, (key, bad)
        # destroying an engine that is still initialised is refused
        assert lib.me_gpu_engine_destroy(second) == -7
    finally:
        lib.me_gpu_engine_bind(second)
        b.close()
        lib.me_gpu_engine_bind(None)
        a.close()
    assert lib.me_gpu_engine_destroy(second) == 0
