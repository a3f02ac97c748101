"""ctypes binding of libsgpu.so — exactly the entry points include/sgpu.h declares."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SGPU_LIB") or os.path.join(HERE, "libsgpu.so")


SIMPLE_SUM, SIMPLE_MAX, SIMPLE_MIN = 0, 1, 2      # include/sgpu.h SGPU_SIMPLE_*
PLANE_MASK, PLANE_DRIZZLE = 0, 1                  # include/sgpu.h SGPU_PLANE_*


class SgpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"sgpu error {code}: {msg}")
        self.code = code


class BlockDesc(C.Structure):
    _fields_ = [("nb_frames", C.c_int), ("nb_layers", C.c_int), ("width", C.c_int), ("height", C.c_int),
                ("dtype", C.c_int)]


class StackArgs(C.Structure):
    _fields_ = [("is_mean", C.c_int), ("type_of_rejection", C.c_int), ("sig", C.c_float * 2),
                ("critical_value", C.c_void_p), ("normalize", C.c_int),
                ("scale", C.c_void_p), ("offset", C.c_void_p), ("mul", C.c_void_p),
                ("weights", C.c_void_p), ("shiftx", C.c_void_p),
                ("output_norm", C.c_int), ("use_32bit_output", C.c_int),
                ("create_rejmaps", C.c_int), ("merge_lowhigh_rejmaps", C.c_int), ("nb_critical", C.c_int),
                ("source_bitpix", C.c_int), ("out_plane_stride", C.c_size_t), ("engine", C.c_int)]


class ImStats(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("total", "ngoodpix", "mean", "median", "sigma", "avgDev", "mad",
                                          "sqrtbwmv", "location", "scale", "min", "max", "normValue", "bgnoise")]

    def asdict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


REDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int)   # sgpu_reduce_fn


class FitsInfo(C.Structure):
    _fields_ = [("bitpix", C.c_int), ("naxes", C.c_long * 3), ("bzero", C.c_double), ("bscale", C.c_double),
                ("datamax", C.c_double), ("has_datamax", C.c_int), ("data_offset", C.c_long),
                ("device_readable", C.c_int)]


class Rect(C.Structure):
    _fields_ = [("x", C.c_int), ("y", C.c_int), ("w", C.c_int), ("h", C.c_int)]


class OverlapStats(C.Structure):
    _fields_ = [("Nij", C.c_size_t)] + [(n, C.c_float) for n in ("medij", "medji", "madij", "madji", "locij", "locji",
                                                                   "scaij", "scaji")]


class SerInfo(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("color_id", "little_endian", "width", "height", "bit_depth", "frame_count", "byte_depth",
                                       "planes")] + [("data_offset", C.c_long), ("device_readable", C.c_int)]


class ImageBlock(C.Structure):
    _fields_ = [("channel", C.c_long), ("start_row", C.c_long), ("end_row", C.c_long), ("height", C.c_long)]


# every symbol include/sgpu.h declares: name -> (restype, argtypes)
P = C.c_void_p
SIGNATURES = {
    "sgpu_abi_version": (C.c_int, []),
    "sgpu_ctx_create": (C.c_int, [C.c_int, C.POINTER(P)]),
    "sgpu_ctx_destroy": (None, [P]),
    "sgpu_ctx_set_stream": (C.c_int, [P, P]),
    "sgpu_ctx_synchronize": (C.c_int, [P]),
    "sgpu_last_error": (C.c_char_p, [P]),
    "sgpu_launch_count": (C.c_uint64, [P]),
    "sgpu_handover_count": (C.c_uint64, [P]),
    "sgpu_last_engine": (C.c_int, [P]),
    "sgpu_simple_block": (C.c_int, [P, P, C.c_int, P, P, C.c_int, C.c_int, P, C.POINTER(C.c_double)]),
    "sgpu_sum_convert": (C.c_int, [P, P, C.c_int, C.c_size_t, C.c_double, C.c_int, P]),
    "sgpu_stats_window_misses": (C.c_uint64, [P]),
    "sgpu_ctx_set_cancel_callback": (C.c_int, [P, P, P]),
    "sgpu_mem_info": (C.c_int, [P, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "sgpu_device_alloc": (C.c_int, [P, C.c_size_t, C.POINTER(P)]),
    "sgpu_device_free": (None, [P, P]),
    "sgpu_download": (C.c_int, [P, P, P, C.c_size_t]),
    "sgpu_ipc_export": (C.c_int, [P, P, P]),
    "sgpu_ipc_open": (C.c_int, [P, P, C.POINTER(P)]),
    "sgpu_ipc_close": (C.c_int, [P, P]),
    "sgpu_host_alloc": (C.c_int, [P, C.c_size_t, C.POINTER(P)]),
    "sgpu_host_free": (None, [P, P]),
    "sgpu_block_create": (C.c_int, [P, C.POINTER(BlockDesc), C.POINTER(P)]),
    "sgpu_block_wrap": (C.c_int, [P, C.POINTER(BlockDesc), P, C.POINTER(P)]),
    "sgpu_block_destroy": (None, [P]),
    "sgpu_block_device_ptr": (P, [P]),
    "sgpu_block_upload_frame": (C.c_int, [P, C.c_int, P, C.c_int, C.c_int, C.c_int]),
    "sgpu_maximized_framing": (C.c_int, [C.c_int, P, P, P, P, C.c_int, C.POINTER(C.c_long), C.POINTER(C.c_long), C.POINTER(C.c_int)]),
    "sgpu_block_upload_frame_sized": (C.c_int, [P, C.c_int, P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "sgpu_block_upload_weight_plane": (C.c_int, [P, C.c_int, C.c_int, P, C.c_int, C.c_int, C.c_int]),
    "sgpu_block_set_weight_planes": (C.c_int, [P, P, P]),
    "sgpu_block_mask_ramp": (C.c_int, [P, C.c_int]),
    "sgpu_fits_probe": (C.c_int, [C.c_char_p, C.POINTER(FitsInfo)]),
    "sgpu_block_upload_fits": (C.c_int, [P, C.c_int, P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "sgpu_block_upload_fits_fd": (C.c_int, [P, C.c_int, C.c_int, C.c_long, C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.c_int]),
    "sgpu_stack_block": (C.c_int, [P, P, C.POINTER(StackArgs), P, P, P, P]),
    "sgpu_stack_host": (C.c_int, [P, C.POINTER(BlockDesc), C.POINTER(P), P, C.POINTER(StackArgs), P, P, P, P]),
    "sgpu_stack_fits": (C.c_int, [P, C.c_int, C.POINTER(C.c_char_p), P, C.POINTER(StackArgs), P, P, P, P]),
    "sgpu_fitseq_probe": (C.c_int, [C.c_char_p, P, C.c_int, C.POINTER(C.c_int)]),
    "sgpu_stack_fitseq": (C.c_int, [P, C.c_char_p, C.c_int, P, P, C.POINTER(StackArgs), P, P, P, P]),
    "sgpu_ser_probe": (C.c_int, [C.c_char_p, C.POINTER(SerInfo)]),
    "sgpu_block_upload_ser_fd": (C.c_int, [P, C.c_int, C.c_int, C.POINTER(SerInfo), C.c_int, C.c_int, C.c_int]),
    "sgpu_stack_ser": (C.c_int, [P, C.c_char_p, C.c_int, P, P, C.POINTER(StackArgs), P, P, P, P]),
    "sgpu_minmax_nonzero": (C.c_int, [P, P, C.c_size_t, C.c_int, C.POINTER(C.c_float)]),
    "sgpu_minmax_nonzero_static": (C.c_int, [P, P, C.c_size_t, C.c_int, C.POINTER(C.c_float)]),
    "sgpu_rescale_0_1": (C.c_int, [P, P, C.c_size_t, C.c_float, C.c_float]),
    "sgpu_image_stats": (C.c_int, [P, P, C.c_int, C.c_int, C.c_long, C.c_long, C.c_int, C.POINTER(ImStats)]),
    "sgpu_image_stats_sharded": (C.c_int, [P, P, C.c_int, C.c_int, C.c_long, C.c_long, C.c_long, C.c_long, C.c_int, P, P,
                                           C.POINTER(ImStats)]),
    "sgpu_imstats_seq_record": (C.c_int, [C.c_int, C.c_int, C.POINTER(ImStats), C.c_char_p, C.c_size_t]),
    "sgpu_imstats_from_seq_record": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(ImStats)]),
    "sgpu_estimators_from_stats": (C.c_int, [C.c_int, C.c_int, C.POINTER(ImStats), C.POINTER(C.c_double),
                                             C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "sgpu_compute_factors": (C.c_int, [C.c_int] * 6 + [P, P, P]),
    "sgpu_noise_weights": (C.c_int, [C.c_int, C.c_int, P, P, P]),
    "sgpu_compute_overlap": (C.c_size_t, [C.c_int] * 4 + [C.c_double] * 4 + [C.POINTER(Rect), C.POINTER(Rect)]),
    "sgpu_overlap_stats_pair": (C.c_int, [P, P, C.c_int, C.c_int, P, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(Rect),
                                          C.POINTER(Rect), C.c_int, C.POINTER(OverlapStats)]),
    "sgpu_overlap_solve": (C.c_int, [C.c_int, C.c_int, C.c_int, P, P, P, P, P, P]),
    "sgpu_wfwhm_weights": (C.c_int, [C.c_int, C.c_int, P, P]),
    "sgpu_nbstars_weights": (C.c_int, [C.c_int, C.c_int, P, P]),
    "sgpu_nbstack_weights": (C.c_int, [C.c_int, C.c_int, P, P]),
    "sgpu_compute_parallel_blocks": (C.c_int, [P, C.c_int, C.c_long, C.POINTER(C.c_long), C.c_int,
                                               C.POINTER(C.c_long), C.POINTER(C.c_int)]),
    "sgpu_synth_fill": (C.c_int, [P, P, C.c_uint64, C.c_int, C.c_int]),
}

_lib = None


def lib():
    """Load libsgpu.so; there is no fallback — a missing library is an error."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(nvcc, sm_100a). siril_b200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)   # AttributeError if the library does not export a declared symbol
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


class Context:
    """sgpu_ctx wrapper: one per GPU / rank."""

    def __init__(self, device=0, stream=None):
        self.L = lib()
        h = P()
        rv = self.L.sgpu_ctx_create(device, C.byref(h))
        if rv:
            raise SgpuError(rv, f"cannot create a context on CUDA device {device} (no GPU? there is no CPU fallback)")
        self.h = h
        self.device = device
        if stream is not None:
            self.check(self.L.sgpu_ctx_set_stream(self.h, P(stream)))

    def check(self, rv):
        if rv:
            raise SgpuError(rv, (self.L.sgpu_last_error(self.h) or b"").decode())

    def synchronize(self):
        self.check(self.L.sgpu_ctx_synchronize(self.h))

    @property
    def launches(self):
        return int(self.L.sgpu_launch_count(self.h))

    @property
    def handovers(self):
        """pixels the last stack call handed from the register / sorted engine to the exact engine"""
        return int(self.L.sgpu_handover_count(self.h))

    @property
    def stats_window_misses(self):
        """medians of image_stats that took the full-histogram path (window missed the median)"""
        return int(self.L.sgpu_stats_window_misses(self.h))

    @property
    def last_engine(self):
        return {1: "stack_exact_kernel", 2: "stack_sorted_kernel",
                5: "stack_lsort_kernel", 6: "stack_psel_kernel"}.get(int(self.L.sgpu_last_engine(self.h)), "none")

    def close(self):
        if getattr(self, "h", None):
            self.L.sgpu_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
