"""The product from Python: a thin ctypes wrapper over libme_gpu.so (CUDA, sm_100a).

There is no fallback of any kind: if the library is missing, or no B200 is visible, construction fails.
Method names follow the C-ABI (include/me_gpu.h), which in turn follows the reference's ME surface.
"""
import ctypes as C

import numpy as np

from . import abi
from .planes import Outputs, PADS, plane_geometry


class MeGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"me_gpu error {code}: {msg}")
        self.code = code


class Engine:
    def __init__(self, max_width, max_height, max_refs_l0=4, max_refs_l1=3, pictures_in_flight=40, devices=None):
        self.lib = abi.load_gpu_lib()
        ids = None
        n = 0
        if devices:
            n = len(devices)
            ids = (C.c_int * n)(*devices)
        self._check(self.lib.me_gpu_init(n, ids, max_width, max_height, max_refs_l0, max_refs_l1, pictures_in_flight))
        self._open = True

    def _check(self, rc):
        if rc != 0:
            raise MeGpuError(rc, self.lib.me_gpu_last_error().decode())

    def close(self):
        if getattr(self, "_open", False):
            self.lib.me_gpu_shutdown()
            self._open = False

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # pictures -------------------------------------------------------------------------------------
    def register_picture(self, picture_number, padded_plane, width, height):
        """padded_plane: 2-D uint8 array, the padded full-resolution luma (68-pixel border already replicated)."""
        a = padded_plane
        assert a.dtype == np.uint8 and a.ndim == 2 and a.strides[1] == 1
        self._check(self.lib.me_gpu_register_picture(picture_number, a.ctypes.data, a.strides[0], PADS[0], PADS[0],
                                                     width, height))

    def register_picture_ptr(self, picture_number, ptr, stride, width, height):
        self._check(self.lib.me_gpu_register_picture(picture_number, ptr, stride, PADS[0], PADS[0], width, height))

    def register_picture_unpadded(self, picture_number, luma):
        a = luma
        assert a.dtype == np.uint8 and a.ndim == 2 and a.strides[1] == 1
        self._check(self.lib.me_gpu_register_picture_unpadded(picture_number, a.ctypes.data, a.strides[0], a.shape[1],
                                                              a.shape[0]))

    def register_picture_unpadded_ptr(self, picture_number, ptr, stride, width, height):
        self._check(self.lib.me_gpu_register_picture_unpadded(picture_number, ptr, stride, width, height))

    # one process per GPU: pyramids pulled from the owner's pool over NVLink (CUDA IPC) ---------------
    def ipc_export_pool(self):
        h, sb = C.create_string_buffer(64), abi.u64(0)
        self._check(self.lib.me_gpu_ipc_export_pool(h, C.byref(sb)))
        return h.raw, sb.value

    def ipc_open_peer(self, peer, handle, slot_bytes):
        self._check(self.lib.me_gpu_ipc_open_peer(peer, handle, slot_bytes))

    def ipc_picture(self, picture_number):
        slot, h = C.c_int(0), C.create_string_buffer(64)
        self._check(self.lib.me_gpu_ipc_picture(picture_number, C.byref(slot), h))
        return slot.value, h.raw

    def register_picture_from_peer(self, picture_number, peer, peer_slot, event_handle, width, height):
        self._check(self.lib.me_gpu_register_picture_from_peer(picture_number, peer, peer_slot, event_handle, width, height))

    def ipc_bytes_pulled(self):
        return int(self.lib.me_gpu_ipc_bytes_pulled())

    def release_picture(self, picture_number):
        self._check(self.lib.me_gpu_release_picture(picture_number))

    def read_pyramid(self, picture_number, width, height):
        planes = []
        for lvl in range(3):
            w, h, pad, stride, rows = plane_geometry(width, height, lvl)
            planes.append(np.zeros((rows, stride), dtype=np.uint8))
        self._check(self.lib.me_gpu_read_pyramid(picture_number, planes[0].ctypes.data, planes[1].ctypes.data,
                                                 planes[2].ctypes.data))
        return planes

    # motion estimation ----------------------------------------------------------------------------
    def submit(self, picture_number, params, ref_numbers):
        """ref_numbers: [[list0 picture numbers], [list1 picture numbers]]"""
        arr = abi.RefNumbers()
        for li, lst in enumerate(ref_numbers):
            for ri, n in enumerate(lst):
                arr[li][ri] = n
        self._check(self.lib.me_gpu_submit(picture_number, C.byref(params), C.byref(arr)))

    def submit_out(self, picture_number, params, ref_numbers, out: Outputs):
        """submit with the result arrays known up front: the D2H copies are queued behind the last kernel"""
        arr = abi.RefNumbers()
        for li, lst in enumerate(ref_numbers):
            for ri, n in enumerate(lst):
                arr[li][ri] = n
        self._check(self.lib.me_gpu_submit_out(picture_number, C.byref(params), C.byref(arr), C.byref(out.c)))

    def output_layout(self, width, height, params):
        offs, total = (C.c_size_t * 11)(), C.c_size_t(0)
        self._check(self.lib.me_gpu_output_layout(width, height, C.byref(params), C.byref(offs), C.byref(total)))
        return list(offs), total.value

    def wait(self, picture_number, out: Outputs):
        self._check(self.lib.me_gpu_wait(picture_number, C.byref(out.c)))
        return out

    def me_picture(self, picture_number, width, height, params, ref_numbers, debug=True):
        out = Outputs(width, height, params, debug=debug)
        self.submit(picture_number, params, ref_numbers)
        self.wait(picture_number, out)
        t = self.timings()
        out.byte_sad_ops = t.byte_sad_ops
        return out

    # dynamic-GOP detector -------------------------------------------------------------------------
    def dg_detector(self, src_picture_number, ref_picture_number, input_resolution):
        """(tot_dist, tot_cplx, tot_active, sum_in_vectors) of the pair, as early_hme leaves them in DGDetectorMetrics"""
        m = abi.MeGpuDgMetrics()
        self._check(self.lib.me_gpu_dg_detector(src_picture_number, ref_picture_number, input_resolution, C.byref(m)))
        return m.as_tuple()

    def timings(self):
        t = abi.MeGpuTimings()
        self._check(self.lib.me_gpu_get_timings(C.byref(t)))
        return t

    def set_profiling(self, on):
        self.lib.me_gpu_set_profiling(1 if on else 0)

    def sad_peak(self):
        """measured VABSDIFF4 lane-ops/s of the device (x4 = byte-SADs/s)"""
        return float(self.lib.me_gpu_measure_sad_peak())
