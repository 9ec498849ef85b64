"""ctypes mirror of include/me_gpu.h (and of the plain picture descriptor the CPU checkers under tests/ take).

Harness plumbing only: the product is libme_gpu.so (CUDA behind a C-ABI); Python is used by tests and
bench.py to feed it host buffers. Field order must match the C headers exactly (checked by
tests/test_abi.py against sizeof/offset probes compiled from the headers).
"""
import ctypes as C
import os

MAX_LISTS, MAX_REFS, NPU = 2, 4, 85
REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", "..", ".."))
PKG = os.path.join(REPO, "svt-av1-psyex_b200")

u8, u16, u32, u64, i8, i16, i32 = (C.c_uint8, C.c_uint16, C.c_uint32, C.c_uint64, C.c_int8, C.c_int16, C.c_int32)


class SearchArea(C.Structure):
    _fields_ = [("width", u16), ("height", u16)]


class SearchAreaMinMax(C.Structure):
    _fields_ = [("sa_min", SearchArea), ("sa_max", SearchArea)]


class MeGpuParams(C.Structure):
    _fields_ = [
        ("picture_number", u64), ("input_resolution", u8), ("enable_me_8x8", u8), ("enable_me_16x16", u8),
        ("max_number_of_pus_per_sb", u8), ("gm_enabled", u8), ("gm_use_distance_based_active_th", u8),
        ("hierarchical_levels", u8), ("pcs_temporal_layer_index", u8), ("similar_brightness_refs", u8),
        ("only_l_bwd", u8), ("max_cand", u8), ("max_refs", u8), ("max_l0", u8),
        ("num_of_list_to_search", u8), ("num_of_ref_pic_to_search", u8 * MAX_LISTS), ("temporal_layer_index", u8),
        ("is_ref", u8),
        ("enable_hme_flag", u8), ("enable_hme_level0_flag", u8), ("enable_hme_level1_flag", u8),
        ("enable_hme_level2_flag", u8), ("hme_search_method", u8), ("me_search_method", u8),
        ("use_best_unipred_cand_only", u8), ("reduce_hme_l0_sr_th_min", u8), ("reduce_hme_l0_sr_th_max", u8),
        ("num_hme_sa_w", u16), ("num_hme_sa_h", u16),
        ("me_sa", SearchAreaMinMax), ("hme_l0_sa", SearchAreaMinMax), ("hme_l1_sa", SearchArea),
        ("hme_l2_sa", SearchArea),
        ("enable_me_hme_ref_pruning", u8), ("prune_ref_if_hme_sad_dev_bigger_than_th", u16),
        ("prune_ref_if_me_sad_dev_bigger_than_th", u16), ("zz_sad_th", u32), ("zz_sad_pct", u16),
        ("phme_sad_th", u32), ("phme_sad_pct", u16),
        ("enable_me_sr_adjustment", u8), ("reduce_me_sr_based_on_mv_length_th", u16),
        ("stationary_hme_sad_abs_th", u16), ("stationary_me_sr_divisor", u16),
        ("reduce_me_sr_based_on_hme_sad_abs_th", u16), ("me_sr_divisor_for_low_hme_sad", u16),
        ("distance_based_hme_resizing", u8),
        ("me_8x8_var_enabled", u8), ("me_sr_div4_th", u32), ("me_sr_div2_th", u32), ("me_sr_mult2_th", u32),
        ("mv_sa_adj_enabled", u8), ("mv_sa_adj_nearest_ref_only", u8), ("mv_sa_adj_mv_size_th", u16),
        ("mv_sa_adj_sa_multiplier", u16),
        ("prehme_enable", u8), ("prehme_skip_search_line", u8), ("prehme_l1_early_exit", u8),
        ("prehme_sa_cfg", SearchAreaMinMax * 2),
        ("prune_me_candidates_th", i32), ("me_early_exit_th", u32), ("me_safe_limit_zz_th", u32),
        ("prev_me_stage_based_exit_th", u32),
        ("me_type", u8), ("tf_me_exit_th", u16),
    ]

    def as_dict(self):
        def conv(v):
            if isinstance(v, C.Structure):
                return {f[0]: conv(getattr(v, f[0])) for f in v._fields_}
            if isinstance(v, C.Array):
                return [conv(x) for x in v]
            return int(v)
        return {f[0]: conv(getattr(self, f[0])) for f in self._fields_}

    @classmethod
    def from_dict(cls, d):
        def fill(obj, dd):
            for k, v in dd.items():
                cur = getattr(obj, k)
                if isinstance(cur, C.Structure):
                    fill(cur, v)
                elif isinstance(cur, C.Array):
                    for i, x in enumerate(v):
                        if isinstance(cur[i], C.Structure):
                            fill(cur[i], x)
                        else:
                            cur[i] = x
                else:
                    setattr(obj, k, v)
        p = cls()
        fill(p, d)
        return p


class MeGpuEncoderConfig(C.Structure):
    _fields_ = [
        ("enc_mode", i8), ("rtc_tune", u8), ("sc_class1", u8), ("qp", u8), ("rate_control_vbr", u8),
        ("fast_decode", u8), ("hierarchical_levels", u8), ("temporal_layer_index", u8), ("is_ref", u8),
        ("slice_type", u8), ("similar_brightness_refs", u8), ("low_frame_rate", u8), ("force_ref_l0", i8),
        ("force_ref_l1", i8), ("width", u16), ("height", u16), ("picture_number", u64),
    ]


class MeGpuTfConfig(C.Structure):
    _fields_ = [("hme_me_level", u8), ("qp_opt", u8), ("qp", u8), ("is_ref", u8), ("temporal_layer_index", u8),
                ("me_exit_th", u16), ("width", u16), ("height", u16), ("picture_number", u64)]


ME_MCTF, ME_OPEN_LOOP = 1, 3  # MeGpuParams.me_type


class MeGpuOutputs(C.Structure):
    _fields_ = [
        ("b64_count", u32), ("pu_count", u32),
        ("total_me_candidate_index", C.POINTER(u8)), ("me_candidate_array", C.POINTER(u8)),
        ("me_mv_array", C.POINTER(u32)),
        ("me_64x64_distortion", C.POINTER(u32)), ("me_32x32_distortion", C.POINTER(u32)),
        ("me_16x16_distortion", C.POINTER(u32)), ("me_8x8_distortion", C.POINTER(u32)),
        ("me_8x8_cost_variance", C.POINTER(u32)), ("rc_me_distortion", C.POINTER(u32)),
        ("stationary_block_present_sb", C.POINTER(u8)), ("rc_me_allow_gm", C.POINTER(u8)),
        ("p_sb_best_sad", C.POINTER(u32)), ("p_sb_best_mv", C.POINTER(u32)),
        ("hme_search_centre", C.POINTER(i16)), ("hme_sad", C.POINTER(u64)), ("do_ref", C.POINTER(u8)),
        ("search_area", C.POINTER(i16)),
        ("tf_best_sad", C.POINTER(u32)), ("tf_best_mv", C.POINTER(u32)), ("tf_hme_centre", C.POINTER(i16)),
        ("tf_hme_sad", C.POINTER(u32)), ("tf_early_exit", C.POINTER(u8)),
        ("tf_tot_horz_blks", u32), ("tf_tot_vert_blks", u32),
    ]


class MeGpuTimings(C.Structure):
    _fields_ = [("pyramid_ms", C.c_float), ("hme_ms", C.c_float), ("search_ms", C.c_float),
                ("epilogue_ms", C.c_float), ("total_ms", C.c_float), ("kernel_launches", u64),
                ("byte_sad_ops", u64), ("byte_sad_ops_fullsearch", u64), ("k_search_ms", C.c_float),
                ("k_fullsearch_ms", C.c_float), ("k_fullsearch_main_ms", C.c_float), ("k_ctrl_ms", C.c_float),
                ("k_epilogue_ms", C.c_float), ("byte_sad_ops_fullsearch_inline", u64)]


class MeGpuDgMetrics(C.Structure):
    _fields_ = [("tot_dist", u64), ("tot_cplx", u32), ("tot_active", u32), ("sum_in_vectors", C.c_int32)]

    def as_tuple(self):
        return (self.tot_dist, self.tot_cplx, self.tot_active, self.sum_in_vectors)


class OraclePlane(C.Structure):
    _fields_ = [("buf", C.c_void_p), ("stride", i32), ("org_x", i32), ("org_y", i32), ("width", i32),
                ("height", i32)]


class OraclePicture(C.Structure):
    _fields_ = [("full", OraclePlane), ("quarter", OraclePlane), ("sixteenth", OraclePlane),
                ("picture_number", u64)]


class RefMrpInputs(C.Structure):
    _fields_ = [("safe_limit_nref", u8), ("safe_limit_zz_th", u32)]


RefArray = (C.POINTER(OraclePicture) * MAX_REFS) * MAX_LISTS
RefNumbers = (u64 * MAX_REFS) * MAX_LISTS

GPU_LIB = os.environ.get("ME_B200_TRACE_LIB") or os.path.join(PKG, "libme_gpu.so")  # tools/trace_hme.py points at the instrumented build
HOST_LIB = os.path.join(PKG, "libme_host.so")

# every symbol include/me_gpu.h declares (tests/test_abi.py checks the .so exports all of them)
GPU_SYMBOLS = [
    "me_gpu_init", "me_gpu_shutdown", "me_gpu_last_error", "me_gpu_device_count", "me_gpu_register_picture",
    "me_gpu_register_picture_unpadded", "me_gpu_release_picture", "me_gpu_read_pyramid", "me_gpu_plane_geometry",
    "me_gpu_submit", "me_gpu_wait", "me_gpu_derive_params", "me_gpu_get_timings", "me_gpu_measure_sad_peak", "me_gpu_set_profiling", "me_gpu_submit_out", "me_gpu_output_layout",
    "me_gpu_dg_detector", "me_gpu_derive_params_tf",
    "me_gpu_wait_pair", "me_gpu_cancel", "me_gpu_wait_done", "me_gpu_read_plane",
    "me_gpu_ipc_export_pool", "me_gpu_ipc_open_peer", "me_gpu_ipc_picture", "me_gpu_register_picture_from_peer",
    "me_gpu_ipc_bytes_pulled",
    "me_gpu_dg_detector_noalloc", "me_gpu_dg_detector_noalloc_setup", "me_gpu_dg_detector_noalloc_error",
    "me_gpu_engine_create", "me_gpu_engine_bind", "me_gpu_engine_destroy",
]


def load_gpu_lib():
    """Load the CUDA engine. Fails loudly if it is not built: there is no fallback."""
    if not os.path.exists(GPU_LIB):
        raise RuntimeError(f"{GPU_LIB} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = C.CDLL(GPU_LIB)
    lib.me_gpu_last_error.restype = C.c_char_p
    lib.me_gpu_init.argtypes = [C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.me_gpu_register_picture.argtypes = [u64, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.me_gpu_register_picture_unpadded.argtypes = [u64, C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib.me_gpu_release_picture.argtypes = [u64]
    lib.me_gpu_read_pyramid.argtypes = [u64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.me_gpu_plane_geometry.argtypes = [C.c_int, C.c_int, C.c_int] + [C.POINTER(C.c_int)] * 5
    lib.me_gpu_submit.argtypes = [u64, C.POINTER(MeGpuParams), C.POINTER(RefNumbers)]
    lib.me_gpu_wait.argtypes = [u64, C.POINTER(MeGpuOutputs)]
    lib.me_gpu_wait_pair.argtypes = [u64, u8, u64, C.POINTER(MeGpuOutputs)]
    lib.me_gpu_cancel.argtypes = [u64]
    lib.me_gpu_read_plane.argtypes = [u64, C.c_int, C.c_void_p, C.c_int]
    lib.me_gpu_ipc_export_pool.argtypes = [C.c_char_p, C.POINTER(u64)]
    lib.me_gpu_ipc_open_peer.argtypes = [C.c_int, C.c_char_p, u64]
    lib.me_gpu_ipc_picture.argtypes = [u64, C.POINTER(C.c_int), C.c_char_p]
    lib.me_gpu_register_picture_from_peer.argtypes = [u64, C.c_int, C.c_int, C.c_char_p, C.c_int, C.c_int]
    lib.me_gpu_ipc_bytes_pulled.restype = u64
    lib.me_gpu_wait_done.argtypes = [u64]
    lib.me_gpu_submit_out.argtypes = [u64, C.POINTER(MeGpuParams), C.POINTER(RefNumbers), C.POINTER(MeGpuOutputs)]
    lib.me_gpu_output_layout.argtypes = [C.c_int, C.c_int, C.POINTER(MeGpuParams), C.POINTER(C.c_size_t * 11),
                                         C.POINTER(C.c_size_t)]
    lib.me_gpu_derive_params.argtypes = [C.POINTER(MeGpuEncoderConfig), C.POINTER(MeGpuParams)]
    lib.me_gpu_derive_params_tf.argtypes = [C.POINTER(MeGpuTfConfig), C.POINTER(MeGpuParams)]
    lib.me_gpu_get_timings.argtypes = [C.POINTER(MeGpuTimings)]
    lib.me_gpu_dg_detector.argtypes = [u64, u64, u8, C.POINTER(MeGpuDgMetrics)]
    lib.me_gpu_measure_sad_peak.restype = C.c_double
    lib.me_gpu_dg_detector_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, u8,
                                               C.POINTER(MeGpuDgMetrics)]
    lib.me_gpu_dg_detector_noalloc_error.restype = C.c_char_p
    lib.me_gpu_engine_create.argtypes = [C.POINTER(C.c_void_p)]
    lib.me_gpu_engine_bind.argtypes = [C.c_void_p]
    lib.me_gpu_engine_destroy.argtypes = [C.c_void_p]
    return lib


def load_host_lib():
    """Host-only part of the product (me_gpu_derive_params); needs no GPU."""
    lib = C.CDLL(HOST_LIB)
    lib.me_gpu_derive_params.argtypes = [C.POINTER(MeGpuEncoderConfig), C.POINTER(MeGpuParams)]
    lib.me_gpu_derive_params_tf.argtypes = [C.POINTER(MeGpuTfConfig), C.POINTER(MeGpuParams)]
    lib.me_gpu_derive_mrp_inputs.argtypes = [C.POINTER(MeGpuEncoderConfig), C.POINTER(u8), C.POINTER(u32)]
    return lib
