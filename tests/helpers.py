"""Shared case construction for the tests: encoder config -> ME controls, synthetic pictures, reference lists."""
import ctypes as C
import functools

import numpy as np

from me_b200 import abi, planes, synth


@functools.lru_cache(maxsize=None)
def host_lib():
    return abi.load_host_lib()


def derive(preset, width, height, poc, n_pictures, force_refs=None, qp=35, lists=None, **kw):
    """MeGpuParams + reference lists of picture `poc` in a random-access 5-layer sequence. `lists` = (temporal layer, list 0,
    list 1) given by the caller (the bench workloads: the encoder's own lists, workloads.lists_for); else synth's table."""
    tl = lists[0] if lists else synth.RA5_RPS[poc % 16][0]
    cfg = abi.MeGpuEncoderConfig(enc_mode=preset, qp=qp, hierarchical_levels=4, width=width, height=height,
                                 picture_number=poc, temporal_layer_index=tl, is_ref=1 if tl < 4 else 0, **kw)
    if force_refs:
        cfg.force_ref_l0, cfg.force_ref_l1 = force_refs
    p = abi.MeGpuParams()
    rc = host_lib().me_gpu_derive_params(C.byref(cfg), C.byref(p))
    assert rc == 0
    if lists:
        l0, l1 = list(lists[1])[:p.num_of_ref_pic_to_search[0]], list(lists[2])[:p.num_of_ref_pic_to_search[1]]
    else:
        _, l0, l1 = synth.ref_lists(poc, n_pictures, p.num_of_ref_pic_to_search[0], p.num_of_ref_pic_to_search[1])
    p.num_of_ref_pic_to_search[0] = len(l0)
    p.num_of_ref_pic_to_search[1] = len(l1)
    if not l1:
        p.num_of_list_to_search = 1
    return cfg, p, l0, l1


def derive_tf(level, width, height, centre, me_exit_th=0, qp_opt=0, qp=35, is_ref=1):
    """MeGpuParams of a temporal-filter ME call (ME_MCTF) for centre picture `centre`."""
    cfg = abi.MeGpuTfConfig(hme_me_level=level, qp_opt=qp_opt, qp=qp, is_ref=is_ref, temporal_layer_index=0,
                            me_exit_th=me_exit_th, width=width, height=height, picture_number=centre)
    p = abi.MeGpuParams()
    rc = host_lib().me_gpu_derive_params_tf(C.byref(cfg), C.byref(p))
    assert rc == 0
    return cfg, p


class PictureSet:
    """Synthetic pictures with CPU-built pyramids (by `builder`: an Oracle or Reference checker)."""

    def __init__(self, width, height, n_pictures, kind, builder, seed=1234):
        self.w, self.h, self.n = width, height, n_pictures
        self.seq = synth.Sequence(width, height, n_pictures, kind=kind, seed=seed)
        self.builder = builder
        self.pics = {}

    def get(self, poc):
        if poc not in self.pics:
            p = planes.Picture(self.w, self.h, poc)
            p.set_luma(self.seq.luma(poc))
            self.builder.build_pyramid(p)
            self.pics[poc] = p
        return self.pics[poc]


class Emulator:
    """tests/emul/libme_emul.so: the product's control logic on the CPU (test harness)."""

    def __init__(self):
        import os
        self.lib = C.CDLL(os.path.join(os.path.dirname(__file__), "emul", "libme_emul.so"))
        self.lib.emul_me_picture.argtypes = [C.POINTER(abi.OraclePicture), C.POINTER(abi.RefArray),
                                             C.POINTER(abi.MeGpuParams), C.POINTER(abi.MeGpuOutputs),
                                             C.POINTER(abi.u64)]

    def me_picture(self, src, refs, params):
        import checkers
        out = planes.Outputs(src.width, src.height, params)
        arr, keep = checkers._refs_array(refs)
        s = src.c_struct()
        ops = abi.u64(0)
        rc = self.lib.emul_me_picture(C.byref(s), C.byref(arr), C.byref(params), C.byref(out.c), C.byref(ops))
        assert rc == 0, rc
        out.byte_sad_ops = ops.value
        return out


def used_refs(l0, l1):
    return [len(l0), len(l1)]


# dynamic-GOP detector cases: (width, height, input_resolution class, kind, source poc, reference poc).
# "blocks" pictures are 8x8-pixel constant blocks of uniform noise: unrelated pictures whose sixteenth planes keep
# their contrast, so that tot_cplx counts; "flat" pictures are constant (90 for even, 60 for odd pocs): every search
# position ties, and the SAD of an even/odd pair is exactly the tot_cplx threshold 16*16*30 (not counted).
DG_CASES = [(320, 192, 1, "motion", 8, 0), (320, 192, 2, "iid", 4, 3), (320, 192, 4, "motion", 4, 3), (320, 192, 4, "blocks", 2, 1),
            (264, 200, 4, "motion", 8, 0), (264, 200, 2, "blocks", 1, 2), (136, 72, 3, "iid", 1, 0), (136, 72, 0, "blocks", 1, 0),
            (640, 360, 2, "motion", 8, 0), (640, 360, 4, "blocks", 3, 5), (64, 64, 4, "iid", 1, 0), (320, 192, 4, "flat", 2, 2), (200, 136, 2, "flat", 2, 1),
            (1280, 720, 3, "motion", 16, 0)]


class DgPictures(PictureSet):
    def __init__(self, width, height, kind, builder, seed=77):
        self.kind = kind
        super().__init__(width, height, 33, kind if kind in ("motion", "iid") else "iid", builder, seed=seed)

    def get(self, poc):
        if poc not in self.pics and self.kind in ("blocks", "flat"):
            rng = np.random.default_rng([self.seq.seed, poc, 5])
            small = rng.integers(0, 256, size=((self.h + 7) // 8, (self.w + 7) // 8), dtype=np.uint8)
            luma = np.repeat(np.repeat(small, 8, axis=0), 8, axis=1)[:self.h, :self.w]
            if self.kind == "flat":
                luma = np.full((self.h, self.w), 60 if poc & 1 else 90, dtype=np.uint8)
            p = planes.Picture(self.w, self.h, poc)
            p.set_luma(np.ascontiguousarray(luma))
            self.builder.build_pyramid(p)
            self.pics[poc] = p
        return super().get(poc)


# temporal-filter ME (ME_MCTF) cases: (width, height, tf hme_me_level, kind, centre poc, neighbour poc, me_exit_th)
TF_CASES = [(320, 192, 0, "motion", 8, 7, 0), (320, 192, 1, "motion", 8, 11, 400), (320, 192, 2, "iid", 4, 5, 20000),
            (320, 192, 3, "motion", 8, 9, 6000), (320, 192, 4, "iid", 8, 6, 0), (264, 200, 0, "iid", 4, 5, 20000),
            (264, 200, 1, "motion", 8, 5, 3000), (264, 200, 2, "motion", 8, 11, 400), (200, 136, 3, "iid", 8, 11, 400),
            (136, 72, 1, "motion", 2, 3, 1000)]


# Cases whose controls make a reference's searches depend on reference (0, 0)'s results (screen content: enable_me_sr_adjustment
# == 2, motion_estimation.c:1349-1364; low delay: reduce_hme_l0_sr_th_*, :1810-1825): (preset, poc, kind, encoder-config extras)
_SC, _RTC, _BOTH = dict(sc_class1=1), dict(rtc_tune=1), dict(rtc_tune=1, sc_class1=1)
DEP_CASES = [(0, 1, "motion", _SC), (2, 4, "motion", _SC), (4, 7, "iid", _SC), (2, 16, "iid", _SC), (8, 4, "motion", _RTC),
             (9, 1, "iid", _RTC), (10, 8, "motion", _RTC), (6, 4, "motion", _BOTH), (8, 7, "iid", _BOTH), (4, 1, "motion", _RTC)]
