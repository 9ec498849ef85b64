"""CPU checkers: the plain-C restatement (oracle/libme_oracle.so) and the reference's own sources
(oracle/_ref/libme_ref.so). TEST INFRASTRUCTURE, kept outside the product package: only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module."""
import ctypes as C
import os

from me_b200 import abi
from me_b200.planes import Outputs, Picture

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
ORACLE_LIB = os.path.join(REPO, "oracle", "libme_oracle.so")
REF_LIB = os.path.join(REPO, "oracle", "_ref", "libme_ref.so")


def _refs_array(refs):
    """refs: [[Picture...], [Picture...]] -> (RefArray, keepalive list)"""
    arr, keep = abi.RefArray(), []
    for li, lst in enumerate(refs):
        for ri, pic in enumerate(lst):
            s = pic.c_struct()
            keep.append(s)
            arr[li][ri] = C.pointer(s)
    return arr, keep


class Oracle:
    def __init__(self):
        if not os.path.exists(ORACLE_LIB):
            raise RuntimeError(f"{ORACLE_LIB} missing: run make -C oracle")
        self.lib = C.CDLL(ORACLE_LIB)
        self.lib.oracle_me_b64_range.argtypes = [C.POINTER(abi.OraclePicture), C.POINTER(abi.RefArray),
                                                 C.POINTER(abi.MeGpuParams), C.POINTER(abi.MeGpuOutputs),
                                                 abi.u32, abi.u32, C.POINTER(abi.u64)]
        self.lib.oracle_build_pyramid.argtypes = [C.POINTER(abi.OraclePicture)]
        self.lib.oracle_sad_loop.argtypes = [C.c_void_p, abi.u32, C.c_void_p, abi.u32, abi.u32, abi.u32,
                                             C.POINTER(abi.u64), C.POINTER(abi.i16), C.POINTER(abi.i16), abi.u32,
                                             abi.u8, abi.i16, abi.i16]
        self.lib.oracle_downsample_2d.argtypes = [C.c_void_p, abi.u32, abi.u32, abi.u32, C.c_void_p, abi.u32, abi.u32]
        self.lib.oracle_dg_detector.argtypes = [C.POINTER(abi.OraclePicture), C.POINTER(abi.OraclePicture), abi.u8,
                                                C.POINTER(abi.MeGpuDgMetrics)]

    def build_pyramid(self, pic: Picture):
        s = pic.c_struct()
        self.lib.oracle_build_pyramid(C.byref(s))

    def dg_detector(self, src: Picture, ref: Picture, input_resolution):
        a, b, m = src.c_struct(), ref.c_struct(), abi.MeGpuDgMetrics()
        rc = self.lib.oracle_dg_detector(C.byref(a), C.byref(b), input_resolution, C.byref(m))
        if rc != 0:
            raise RuntimeError(f"oracle_dg_detector failed: {rc}")
        return m.as_tuple()

    def me_picture(self, src: Picture, refs, params, first=0, count=1 << 30, debug=True):
        out = Outputs(src.width, src.height, params, debug=debug)
        arr, keep = _refs_array(refs)
        s = src.c_struct()
        ops = abi.u64(0)
        rc = self.lib.oracle_me_b64_range(C.byref(s), C.byref(arr), C.byref(params), C.byref(out.c), first, count,
                                          C.byref(ops))
        if rc != 0:
            raise RuntimeError(f"oracle_me_b64_range failed: {rc}")
        out.byte_sad_ops = ops.value
        return out


class Reference:
    """The reference's own C (simd=0) or AVX2 / AVX-512 (simd=1/2) ME, via oracle/ref_harness.c."""

    def __init__(self):
        if not os.path.exists(REF_LIB):
            raise RuntimeError(f"{REF_LIB} missing: run make -C oracle (needs /root/reference)")
        self.lib = C.CDLL(REF_LIB)
        self.lib.ref_me_b64_range.argtypes = [C.POINTER(abi.OraclePicture), C.POINTER(abi.RefArray),
                                              C.POINTER(abi.MeGpuParams), C.POINTER(abi.MeGpuOutputs), abi.u32,
                                              abi.u32, C.c_int]
        self.lib.ref_build_pyramid.argtypes = [C.POINTER(abi.OraclePicture)]
        self.lib.ref_derive_params.argtypes = [C.POINTER(abi.MeGpuEncoderConfig), C.POINTER(abi.RefMrpInputs),
                                               C.POINTER(abi.MeGpuParams)]
        self.lib.ref_sad_loop.argtypes = [C.c_void_p, abi.u32, C.c_void_p, abi.u32, abi.u32, abi.u32,
                                          C.POINTER(abi.u64), C.POINTER(abi.i16), C.POINTER(abi.i16), abi.u32,
                                          abi.u8, abi.i16, abi.i16]
        self.lib.ref_downsample_2d.argtypes = [C.c_void_p, abi.u32, abi.u32, abi.u32, C.c_void_p, abi.u32, abi.u32]
        self.lib.ref_dg_detector.argtypes = [C.POINTER(abi.OraclePicture), C.POINTER(abi.OraclePicture), abi.u8,
                                             C.POINTER(abi.MeGpuDgMetrics), C.c_int]

    def dg_detector(self, src: Picture, ref: Picture, input_resolution, simd=0):
        a, b, m = src.c_struct(), ref.c_struct(), abi.MeGpuDgMetrics()
        rc = self.lib.ref_dg_detector(C.byref(a), C.byref(b), input_resolution, C.byref(m), simd)
        if rc != 0:
            raise RuntimeError(f"ref_dg_detector failed: {rc}")
        return m.as_tuple()

    @staticmethod
    def available():
        return os.path.exists(REF_LIB)

    def build_pyramid(self, pic: Picture):
        s = pic.c_struct()
        self.lib.ref_build_pyramid(C.byref(s))

    def me_picture(self, src: Picture, refs, params, first=0, count=1 << 30, simd=0, debug=True):
        out = Outputs(src.width, src.height, params, debug=debug)
        arr, keep = _refs_array(refs)
        s = src.c_struct()
        rc = self.lib.ref_me_b64_range(C.byref(s), C.byref(arr), C.byref(params), C.byref(out.c), first, count, simd)
        if rc != 0:
            raise RuntimeError(f"ref_me_b64_range failed: {rc}")
        return out

    def me_step_mt(self, jobs, threads, simd=1, every=1):
        """jobs: [(src Picture, refs, params, Outputs)]: open-loop ME of all of them on `threads` threads (the timed CPU
        arm: segments of all pictures in one queue, a MeContext per thread). Returns the b64s searched."""
        class RefStepJob(C.Structure):
            _fields_ = [("src", C.POINTER(abi.OraclePicture)), ("refs", C.POINTER(abi.RefArray)),
                        ("p", C.POINTER(abi.MeGpuParams)), ("out", C.POINTER(abi.MeGpuOutputs))]
        arr = (RefStepJob * len(jobs))()
        keep = []
        for i, (src, refs, p, out) in enumerate(jobs):
            ra, k = _refs_array(refs)
            s = src.c_struct()
            keep.append((ra, k, s))
            arr[i].src, arr[i].refs, arr[i].p, arr[i].out = C.pointer(s), C.pointer(ra), C.pointer(p), C.pointer(out.c)
        done = C.c_long(0)
        self.lib.ref_me_step_mt.argtypes = [C.POINTER(RefStepJob), C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_long)]
        rc = self.lib.ref_me_step_mt(arr, len(jobs), threads, simd, every, C.byref(done))
        if rc != 0:
            raise RuntimeError(f"ref_me_step_mt failed: {rc}")
        return done.value

    def has_hook(self):
        return hasattr(self.lib, "ref_hook_me_picture")

    def hook_me_picture(self, src: Picture, refs, params, gpu_library_path):
        """The reference-side binding (integration/svt_me_gpu_hook.c, compiled against the reference's headers) driven
        like the encoder drives its ME stage: a real pcs / MeContext in, pcs->pa_me_data->me_results and the per-b64
        pcs arrays out. Runs the CUDA engine through dlopen(gpu_library_path): GPU tests only."""
        self.lib.ref_hook_me_picture.argtypes = [C.POINTER(abi.OraclePicture), C.POINTER(abi.RefArray),
                                                 C.POINTER(abi.MeGpuParams), C.POINTER(abi.MeGpuOutputs), C.c_char_p]
        out = Outputs(src.width, src.height, params, debug=False)
        arr, keep = _refs_array(refs)
        s = src.c_struct()
        rc = self.lib.ref_hook_me_picture(C.byref(s), C.byref(arr), C.byref(params), C.byref(out.c),
                                          gpu_library_path.encode())
        if rc != 0:
            raise RuntimeError(f"ref_hook_me_picture failed: {rc}")
        return out

    def hook_tf_pair(self, src: Picture, ref: Picture, params, gpu_library_path):
        """svt_me_gpu_tf_pair + svt_me_gpu_tf_load_block of the binding for one (centre, neighbour) pair"""
        self.lib.ref_hook_tf_pair.argtypes = [C.POINTER(abi.OraclePicture), C.POINTER(abi.OraclePicture),
                                              C.POINTER(abi.MeGpuParams), C.POINTER(abi.MeGpuOutputs), C.c_char_p]
        out = Outputs(src.width, src.height, params, debug=False)
        a, b = src.c_struct(), ref.c_struct()
        rc = self.lib.ref_hook_tf_pair(C.byref(a), C.byref(b), C.byref(params), C.byref(out.c), gpu_library_path.encode())
        if rc != 0:
            raise RuntimeError(f"ref_hook_tf_pair failed: {rc}")
        return out

    def hook_dg_detector(self, src: Picture, ref: Picture, input_resolution, gpu_library_path):
        self.lib.ref_hook_dg_detector.argtypes = [C.POINTER(abi.OraclePicture), C.POINTER(abi.OraclePicture), abi.u8,
                                                  C.POINTER(abi.MeGpuDgMetrics), C.c_char_p]
        a, b, m = src.c_struct(), ref.c_struct(), abi.MeGpuDgMetrics()
        rc = self.lib.ref_hook_dg_detector(C.byref(a), C.byref(b), input_resolution, C.byref(m), gpu_library_path.encode())
        if rc != 0:
            raise RuntimeError(f"ref_hook_dg_detector failed: {rc}")
        return m.as_tuple()

    def hook_pyramid_for_analysis(self, pic: Picture, cpu: Picture, gpu_library_path):
        """svt_me_gpu_pyramid_for_analysis of the binding: fills pic's quarter / sixteenth planes from the device; returns the
        reference's calculate_histogram of the sixteenth plane of `cpu` (CPU-built) and of the fetched one"""
        import numpy as np
        self.lib.ref_hook_pyramid_for_analysis.argtypes = [C.POINTER(abi.OraclePicture), C.POINTER(abi.OraclePicture), C.c_void_p,
                                                           C.c_void_p, C.c_char_p]
        a, b = pic.c_struct(), cpu.c_struct()
        h0, h1 = np.zeros(256, dtype=np.uint32), np.zeros(256, dtype=np.uint32)
        rc = self.lib.ref_hook_pyramid_for_analysis(C.byref(a), C.byref(b), h0.ctypes.data, h1.ctypes.data, gpu_library_path.encode())
        if rc != 0:
            raise RuntimeError(f"ref_hook_pyramid_for_analysis failed: {rc}")
        return h0, h1

    def derive_params_tf(self, cfg):
        p = abi.MeGpuParams()
        self.lib.ref_derive_params_tf.argtypes = [C.POINTER(abi.MeGpuTfConfig), C.POINTER(abi.MeGpuParams)]
        rc = self.lib.ref_derive_params_tf(C.byref(cfg), C.byref(p))
        if rc != 0:
            raise RuntimeError("ref_derive_params_tf failed")
        return p

    def derive_params(self, cfg, safe_limit_nref, safe_limit_zz_th):
        p = abi.MeGpuParams()
        mrp = abi.RefMrpInputs(safe_limit_nref, safe_limit_zz_th)
        rc = self.lib.ref_derive_params(C.byref(cfg), C.byref(mrp), C.byref(p))
        if rc != 0:
            raise RuntimeError("ref_derive_params failed")
        return p
