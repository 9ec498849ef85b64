"""numpy-side picture planes and result slabs in the layouts the C-ABI uses."""
import ctypes as C

import numpy as np

from . import abi

PADS = (68, 32, 16)  # full / quarter / sixteenth padding (reference_object.c:249-295, enc_handle.c:4110)


def plane_geometry(width, height, level):
    """(plane_w, plane_h, pad, stride, rows) of pyramid level 0/1/2 for a width x height picture."""
    w, h, pad = width >> level, height >> level, PADS[level]
    return w, h, pad, w + 2 * pad, h + 2 * pad


def b64_grid(width, height):
    return (width + 63) // 64, (height + 63) // 64


class Picture:
    """Host copy of one PA reference object: padded full plane + 1/4 + 1/16 planes."""

    def __init__(self, width, height, picture_number=0):
        assert width % 8 == 0 and height % 8 == 0, "luma size must be a multiple of 8 (aligned_width/height)"
        self.width, self.height, self.picture_number = width, height, picture_number
        self.planes = []
        for lvl in range(3):
            w, h, pad, stride, rows = plane_geometry(width, height, lvl)
            self.planes.append(np.zeros((rows, stride), dtype=np.uint8))

    def interior(self, level):
        w, h, pad, _, _ = plane_geometry(self.width, self.height, level)
        return self.planes[level][pad:pad + h, pad:pad + w]

    def set_luma(self, y):
        """Store an unpadded luma image and edge-replicate the 68-pixel border
        (svt_aom_pad_input_pictures, pic_analysis_process.c:2198)."""
        assert y.shape == (self.height, self.width) and y.dtype == np.uint8
        self.planes[0][...] = np.pad(y, PADS[0], mode="edge")

    def c_struct(self):
        pic = abi.OraclePicture()
        for lvl, name in enumerate(("full", "quarter", "sixteenth")):
            w, h, pad, stride, _ = plane_geometry(self.width, self.height, lvl)
            pl = getattr(pic, name)
            pl.buf = self.planes[lvl].ctypes.data
            pl.stride, pl.org_x, pl.org_y, pl.width, pl.height = stride, pad, pad, w, h
        pic.picture_number = self.picture_number
        return pic


_OUT_FIELDS = [
    # name, dtype, per-b64 shape factory (P, max_cand, max_refs)
    ("total_me_candidate_index", np.uint8, lambda P, mc, mr: (P,)),
    ("me_candidate_array", np.uint8, lambda P, mc, mr: (P * mc,)),
    ("me_mv_array", np.uint32, lambda P, mc, mr: (P * mr,)),
    ("me_64x64_distortion", np.uint32, lambda P, mc, mr: ()),
    ("me_32x32_distortion", np.uint32, lambda P, mc, mr: ()),
    ("me_16x16_distortion", np.uint32, lambda P, mc, mr: ()),
    ("me_8x8_distortion", np.uint32, lambda P, mc, mr: ()),
    ("me_8x8_cost_variance", np.uint32, lambda P, mc, mr: ()),
    ("rc_me_distortion", np.uint32, lambda P, mc, mr: ()),
    ("stationary_block_present_sb", np.uint8, lambda P, mc, mr: ()),
    ("rc_me_allow_gm", np.uint8, lambda P, mc, mr: ()),
    ("p_sb_best_sad", np.uint32, lambda P, mc, mr: (2, 4, 85)),
    ("p_sb_best_mv", np.uint32, lambda P, mc, mr: (2, 4, 85)),
    ("hme_search_centre", np.int16, lambda P, mc, mr: (2, 4, 2)),
    ("hme_sad", np.uint64, lambda P, mc, mr: (2, 4)),
    ("do_ref", np.uint8, lambda P, mc, mr: (2, 4)),
    ("search_area", np.int16, lambda P, mc, mr: (2, 4, 4)),
    ("tf_best_sad", np.uint32, lambda P, mc, mr: (85,)),
    ("tf_best_mv", np.uint32, lambda P, mc, mr: (85,)),
    ("tf_hme_centre", np.int16, lambda P, mc, mr: (2,)),
    ("tf_hme_sad", np.uint32, lambda P, mc, mr: ()),
    ("tf_early_exit", np.uint8, lambda P, mc, mr: ()),
]
PRODUCT_FIELDS = [f[0] for f in _OUT_FIELDS[:11]]  # what the encoder consumes (TASK_PAME)
DEBUG_FIELDS = [f[0] for f in _OUT_FIELDS[11:17]]  # parity taps
TF_FIELDS = [f[0] for f in _OUT_FIELDS[17:]]  # what the temporal filter consumes (ME_MCTF jobs)


def exported_pu_count(params):
    if not params.enable_me_16x16:
        return 5
    return params.max_number_of_pus_per_sb if params.enable_me_8x8 else 21


class Outputs:
    """Zero-initialised result slabs + the MeGpuOutputs struct pointing at them."""

    def __init__(self, width, height, params, debug=True):
        bw, bh = b64_grid(width, height)
        self.b64_count = bw * bh
        P = exported_pu_count(params)
        self.arrays = {}
        self.c = abi.MeGpuOutputs()
        self.c.b64_count = self.b64_count
        tf = params.me_type == abi.ME_MCTF
        for name, dt, shp in _OUT_FIELDS:
            if not debug and name in DEBUG_FIELDS:
                continue
            if (name in TF_FIELDS) != tf and name not in DEBUG_FIELDS:
                continue
            a = np.zeros((self.b64_count,) + shp(P, params.max_cand, params.max_refs), dtype=dt)
            self.arrays[name] = a
            ptr_t = dict(abi.MeGpuOutputs._fields_)[name]
            setattr(self.c, name, C.cast(a.ctypes.data, ptr_t))

    def __getitem__(self, k):
        return self.arrays[k]

    @property
    def tf_blocks(self):
        """(tf_tot_horz_blks, tf_tot_vert_blks) of an ME_MCTF job"""
        return (int(self.c.tf_tot_horz_blks), int(self.c.tf_tot_vert_blks))


def compare_outputs(a, b, fields=None, used_refs=None):
    """Return a list of human-readable mismatches between two Outputs (empty = bit-exact).
    used_refs: optional [n_l0, n_l1]; debug taps of unused (list, ref) slots are ignored (the reference
    never initialises some of them)."""
    bad = []
    for name in fields or a.arrays.keys():
        if name not in a.arrays or name not in b.arrays:
            continue
        x, y = a.arrays[name], b.arrays[name]
        if used_refs is not None and name in DEBUG_FIELDS:
            mask = np.zeros((2, 4), dtype=bool)
            mask[0, :used_refs[0]] = True
            mask[1, :used_refs[1]] = True
            x, y = x[:, mask], y[:, mask]
        if x.shape != y.shape:
            bad.append(f"{name}: shape {x.shape} vs {y.shape}")
            continue
        if not np.array_equal(x, y):
            idx = np.argwhere(x != y)
            bad.append(f"{name}: {len(idx)} mismatches, first at {tuple(idx[0])}: {x[tuple(idx[0])]} vs {y[tuple(idx[0])]}")
    return bad
