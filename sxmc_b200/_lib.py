"""ctypes binding of libsxmc_b200.so (the C ABI declared in include/sxmc_b200.h).

There is no fallback: if the shared library is missing, import of the compute API fails
loudly (build it with ``python -c "import __graft_entry__ as g; g.build()"`` or
``make -C sxmc_b200/csrc``)."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# SXMC_B200_LIB: tuning experiments load an alternative build of the same library
LIB_PATH = os.environ.get("SXMC_B200_LIB") or os.path.join(HERE, "libsxmc_b200.so")

c_dp = C.POINTER(C.c_double)
c_fp = C.POINTER(C.c_float)
c_ip = C.POINTER(C.c_int)
c_up = C.POINTER(C.c_uint)
c_sp = C.POINTER(C.c_short)
c_vp = C.c_void_p


class SxmcError(RuntimeError):
    """Any non-zero status from the C ABI.  ``code`` is the SXMC_E* value; for
    SXMC_EINVAL the message is the reference's pdfz::Error text."""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code
        self.msg = msg


class ChainConfig(C.Structure):
    _fields_ = [("nsources", C.c_int), ("nsignals", C.c_int), ("nparameters", C.c_int),
                ("nobservables", C.c_int), ("parameter_means", c_dp),
                ("parameter_sigma", c_dp), ("parameter_fixed", c_ip),
                ("systematics_fixed", C.c_int), ("nsystematics", C.c_int),
                ("pdfs", C.POINTER(c_vp)), ("nexpected", c_dp), ("n_mc", c_up),
                ("source_id", c_sp)]


# every symbol include/sxmc_b200.h declares: name -> (restype, argtypes)
_SIZE = C.c_size_t
SYMBOLS = {
    "sxmc_last_error": (C.c_char_p, []),
    "sxmc_version": (C.c_char_p, []),
    "sxmc_init": (C.c_int, [C.c_int]),
    "sxmc_device_synchronize": (C.c_int, []),
    "sxmc_buf_alloc": (C.c_int, [_SIZE, C.POINTER(c_vp)]),
    "sxmc_buf_free": (C.c_int, [c_vp]),
    "sxmc_buf_upload": (C.c_int, [c_vp, c_vp, _SIZE]),
    "sxmc_buf_download": (C.c_int, [c_vp, c_vp, _SIZE]),
    "sxmc_buf_zero": (C.c_int, [c_vp, _SIZE]),
    "sxmc_host_alloc": (C.c_int, [_SIZE, C.POINTER(c_vp)]),
    "sxmc_host_free": (C.c_int, [c_vp]),
    "sxmc_hist_create": (C.c_int, [c_fp, _SIZE, C.c_int, C.c_int, c_dp, C.c_int, c_dp, C.c_int,
                                   c_ip, C.c_int, C.c_uint, C.POINTER(c_vp)]),
    "sxmc_hist_destroy": (C.c_int, [c_vp]),
    "sxmc_hist_add_systematic": (C.c_int, [c_vp, C.c_int, C.c_int, C.c_int, C.c_int, c_sp]),
    "sxmc_hist_set_eval_points": (C.c_int, [c_vp, c_fp, _SIZE]),
    "sxmc_hist_set_pdf_buffer": (C.c_int, [c_vp, c_vp, C.c_long, C.c_long]),
    "sxmc_hist_set_norm_buffer": (C.c_int, [c_vp, c_vp, C.c_long]),
    "sxmc_hist_set_param_buffer": (C.c_int, [c_vp, c_vp, C.c_long, C.c_long]),
    "sxmc_hist_eval_async": (C.c_int, [c_vp, C.c_int]),
    "sxmc_hist_eval_finished": (C.c_int, [c_vp]),
    "sxmc_hists_eval_async": (C.c_int, [C.POINTER(c_vp), C.c_int, C.c_int]),
    "sxmc_hists_eval_host": (C.c_int, [C.POINTER(c_vp), C.c_int, c_dp, C.c_int, C.c_long, c_up, c_fp]),
    "sxmc_hist_total_nbins": (C.c_int, [c_vp]),
    "sxmc_hist_bin_volume": (C.c_double, [c_vp]),
    "sxmc_hist_nsamples": (_SIZE, [c_vp]),
    "sxmc_hist_neval_points": (_SIZE, [c_vp]),
    "sxmc_hist_get_bins": (C.c_int, [c_vp, c_up]),
    "sxmc_hist_get_read_bins": (C.c_int, [c_vp, c_ip]),
    "sxmc_hist_get_samples": (C.c_int, [c_vp, c_fp]),
    "sxmc_hist_set_layout": (C.c_int, [c_vp, C.c_int]),
    "sxmc_hist_prepare": (C.c_int, [c_vp]),
    "sxmc_set_bin_tuning": (C.c_int, [C.c_int, C.c_int, C.c_int]),
    "sxmc_enable_kernel_timing": (C.c_int, [C.c_int]),
    "sxmc_last_kernel_ms": (C.c_int, [c_fp, c_fp]),
    "sxmc_bin_debug_stamps": (C.c_int, [C.c_int, C.POINTER(C.c_uint64)]),
    "sxmc_nll_debug_stamps": (C.c_int, [C.c_int, C.POINTER(C.c_uint64)]),
    "sxmc_jit_compile_check": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint, C.POINTER(_SIZE)]),
    "sxmc_nll_eval": (C.c_int, [c_vp, _SIZE, _SIZE, c_vp, _SIZE, _SIZE, c_vp, c_vp, c_vp, c_vp,
                                c_vp, c_vp, c_dp, c_dp]),
    "sxmc_nll_eval_noalloc": (C.c_int, [c_vp, _SIZE, _SIZE, c_vp, _SIZE, _SIZE, c_vp, c_vp, c_vp, c_vp,
                                        c_vp, c_vp, c_vp, c_vp]),
    "sxmc_chain_create": (C.c_int, [C.POINTER(ChainConfig), C.POINTER(c_vp)]),
    "sxmc_chain_destroy": (C.c_int, [c_vp]),
    "sxmc_chain_set_seed": (C.c_int, [c_vp, C.c_uint64, C.c_uint64]),
    "sxmc_chain_set_tapes": (C.c_int, [c_vp, c_dp, _SIZE, c_dp, _SIZE]),
    "sxmc_chain_set_comm": (C.c_int, [c_vp, c_vp]),
    "sxmc_chain_run": (C.c_int, [c_vp, c_fp, _SIZE, C.c_uint, C.c_float, C.c_int, C.c_uint,
                                 C.POINTER(C.c_long)]),
    "sxmc_chain_begin": (C.c_int, [c_vp, c_fp, _SIZE, C.c_uint, C.c_float, C.c_int]),
    "sxmc_chain_advance": (C.c_int, [c_vp, C.c_uint]),
    "sxmc_chain_end": (C.c_int, [c_vp, C.POINTER(C.c_long)]),
    "sxmc_chain_advance_timed": (C.c_int, [c_vp, C.c_uint, _SIZE, c_fp, c_fp, c_fp]),
    "sxmc_chain_read": (C.c_int, [c_vp, c_fp, C.c_long]),
    "sxmc_chain_accepted": (C.c_int, [c_vp, C.POINTER(C.c_long)]),
    "sxmc_chain_loop_ms": (C.c_int, [c_vp, c_fp]),
    "sxmc_chain_launches": (C.c_int, [c_vp, C.POINTER(C.c_long)]),
    "sxmc_chain_enable_trace": (C.c_int, [c_vp, C.c_int]),
    "sxmc_chain_read_trace": (C.c_int, [c_vp, c_dp, c_dp, C.c_uint]),
    "sxmc_chain_set_data": (C.c_int, [c_vp, c_fp, _SIZE]),
    "sxmc_chain_nll_at": (C.c_int, [c_vp, c_dp, C.c_int, c_dp, c_dp]),
    "sxmc_chain_get_norms": (C.c_int, [c_vp, c_up]),
    "sxmc_chain_get_lut": (C.c_int, [c_vp, c_fp]),
    "sxmc_chain_get_bins": (C.c_int, [c_vp, c_up]),
    "sxmc_chain_set_mode": (C.c_int, [c_vp, C.c_int]),
    "sxmc_chain_last_comm_ms": (C.c_int, [c_vp, c_fp]),
    "sxmc_chain_bin_kernel_kind": (C.c_int, [c_vp, C.POINTER(C.c_int)]),
    "sxmc_chain_bin_plan": (C.c_int, [c_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "sxmc_chain_last_kernel_ms": (C.c_int, [c_vp, c_fp, c_fp]),
    "sxmc_chain_last_path": (C.c_int, [c_vp, C.POINTER(C.c_int)]),
    "sxmc_chain_phase_stamps": (C.c_int, [c_vp, C.c_int, C.POINTER(C.c_uint64)]),
    "sxmc_generator_create": (C.c_int, [C.POINTER(c_vp), C.c_int, c_dp, c_up, c_dp, C.c_int, c_fp, c_fp,
                                        C.c_int, C.POINTER(c_vp)]),
    "sxmc_generator_destroy": (C.c_int, [c_vp]),
    "sxmc_generator_efficiency": (C.c_int, [c_vp, c_dp]),
    "sxmc_generator_sample": (C.c_int, [c_vp, C.c_int, C.c_uint64, C.c_uint64, c_fp, _SIZE, c_up,
                                        C.POINTER(_SIZE)]),
    "sxmc_rows_contour": (C.c_int, [c_vp, C.c_long, C.c_int, C.c_float, c_fp, c_fp, c_fp, c_fp,
                                    C.POINTER(C.c_long)]),
    "sxmc_chain_contour": (C.c_int, [c_vp, C.c_float, c_fp, c_fp, c_fp, c_fp, C.POINTER(C.c_long)]),
    "sxmc_batch_create": (C.c_int, [C.POINTER(ChainConfig), C.c_int, C.POINTER(c_vp)]),
    "sxmc_batch_destroy": (C.c_int, [c_vp]),
    "sxmc_batch_set_seed": (C.c_int, [c_vp, C.c_int, C.c_uint64, C.c_uint64]),
    "sxmc_batch_set_tapes": (C.c_int, [c_vp, C.c_int, c_dp, _SIZE, c_dp, _SIZE]),
    "sxmc_batch_set_data": (C.c_int, [c_vp, C.c_int, c_fp, _SIZE]),
    "sxmc_batch_generate_data": (C.c_int, [c_vp, C.c_int, c_vp, C.c_int, C.c_uint64, C.c_uint64, c_up]),
    "sxmc_batch_enable_trace": (C.c_int, [c_vp, C.c_int]),
    "sxmc_batch_run": (C.c_int, [c_vp, C.c_int, C.c_uint, C.c_float, C.c_int]),
    "sxmc_batch_nrows": (C.c_int, [c_vp, C.c_int, C.POINTER(C.c_long), C.POINTER(C.c_long)]),
    "sxmc_batch_read": (C.c_int, [c_vp, C.c_int, c_fp, C.c_long]),
    "sxmc_batch_read_trace": (C.c_int, [c_vp, C.c_int, c_dp, c_dp, C.c_uint]),
    "sxmc_batch_contour": (C.c_int, [c_vp, C.c_int, C.c_float, c_fp, c_fp, c_fp, c_fp, C.POINTER(C.c_long)]),
    "sxmc_batch_loop_ms": (C.c_int, [c_vp, c_fp]),
    "sxmc_batch_last_path": (C.c_int, [c_vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "sxmc_batch_phase_stamps": (C.c_int, [c_vp, C.c_int, C.POINTER(C.c_uint64)]),
    "sxmc_chain_p2p_export": (C.c_int, [c_vp, c_vp]),
    "sxmc_chain_p2p_connect": (C.c_int, [c_vp, c_vp, C.c_int, C.c_int]),
    "sxmc_chain_p2p_status": (C.c_int, [c_vp, C.POINTER(C.c_uint), C.POINTER(C.c_int)]),
    "sxmc_chain_p2p_clocks": (C.c_int, [c_vp, C.POINTER(C.c_double)]),
    "sxmc_comm_unique_id": (C.c_int, [c_vp]),
    "sxmc_comm_create": (C.c_int, [c_vp, C.c_int, C.c_int, C.POINTER(c_vp)]),
    "sxmc_comm_destroy": (C.c_int, [c_vp]),
    "sxmc_comm_allreduce_u32": (C.c_int, [c_vp, c_vp, _SIZE]),
}

_lib = None


def lib():
    """The loaded library with prototypes set.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "sxmc_b200: %s is missing -- build the CUDA library first "
                "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU "
                "fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise SxmcError(rc, lib().sxmc_last_error().decode())


def ptr(a, t):
    return a.ctypes.data_as(t)
