"""ctypes binding of libmerl_b200.so (include/merl_b200.h).

This is the only door into the compute path: there is no Python/torch/CPU fallback.  If the
shared library is missing or a call fails, a MerlError is raised with the library's message.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MERL_B200_LIB") or os.path.join(_PKG, "libmerl_b200.so")  # env: tuning builds only

# every symbol include/merl_b200.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "merl_table_params_default", "merl_table_create_from_file", "merl_table_create_from_host",
    "merl_table_destroy", "merl_table_info", "merl_table_set_cdf", "merl_table_download",
    "merl_table_refresh_sampler",
    "merl_table_update_cells",
    "merl_grad_clear_counts",
    "merl_file_report", "merl_table_save_packed", "merl_table_create_from_packed",
    "merl_bins", "merl_eval", "merl_pdf", "merl_eval_pdf", "merl_sample", "merl_backward",
    "merl_eval_pdf_aos", "merl_sample_aos", "merl_eval_pdf_host", "merl_sample_host",
    "merl_eval_pdf_host_static",
    "merl_render", "merl_set_render_graphs",
    "merl_comm_unique_id", "merl_comm_create", "merl_comm_destroy", "merl_allreduce_sum_f32",
    "merl_grad_apply_mask",
    "merl_peerbuf_create", "merl_peerbuf_destroy", "merl_peerbuf_info", "merl_peerbuf_allreduce",
    "merl_last_error", "merl_abi_version", "merl_launch_count", "merl_set_sample_search",
]

BWD_WARP_AGGREGATE = 1
BWD_DEFER_MASK = 2
BWD_CTA_AGGREGATE = 4
XCHG_PEER, XCHG_NCCL, XCHG_NVLS = 0, 1, 2
ERR_UNSUPPORTED = -5
STATIC_STAGE_MAX_QUERIES = 65536


class MerlError(RuntimeError):
    pass


class TableParams(C.Structure):
    _fields_ = [
        ("nh", C.c_int32), ("nd", C.c_int32), ("np", C.c_int32), ("theta_h_sqrt", C.c_int32),
        ("ni", C.c_int32), ("clamp_negative", C.c_int32), ("cdf_nm", C.c_int32), ("cdf_nq", C.c_int32),
        ("l2_persist", C.c_int32), ("reserved", C.c_int32),
        ("cdf_floor", C.c_float), ("spec_weight", C.c_float), ("scale", C.c_double * 3),
    ]


class TableInfo(C.Structure):
    _fields_ = [
        ("nh", C.c_int32), ("nd", C.c_int32), ("np", C.c_int32), ("ni", C.c_int32),
        ("theta_h_sqrt", C.c_int32), ("device", C.c_int32),
        ("ncell", C.c_int64), ("table_bytes", C.c_int64),
        ("d_cells", C.c_void_p), ("d_cdf", C.c_void_p), ("d_dens", C.c_void_p),
        ("spec_weight", C.c_float), ("smem_eval_bytes", C.c_int32), ("smem_sample_bytes", C.c_int32),
    ]


class PeerbufInfo(C.Structure):
    _fields_ = [("data", C.c_void_p), ("count", C.c_int64), ("transport", C.c_int32), ("nranks", C.c_int32),
                ("rank", C.c_int32), ("status", C.c_int32), ("registration", C.c_char * 64)]


class FileReport(C.Structure):
    _fields_ = [
        ("dims", C.c_int32 * 3), ("is_merl_layout", C.c_int32),
        ("ncell", C.c_int64), ("file_bytes", C.c_int64), ("expected_bytes", C.c_int64),
        ("negative", C.c_int64 * 3), ("nonfinite", C.c_int64 * 3),
        ("min_valid", C.c_double * 3), ("max_valid", C.c_double * 3),
    ]


_lib = None


def lib():
    """Load libmerl_b200.so; fail loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MerlError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C mitsuba_customization_b200/csrc` (there is no fallback path)")
        L = C.CDLL(LIB_PATH)
        L.merl_last_error.restype = C.c_char_p
        L.merl_launch_count.restype = C.c_int64
        vp, i64, i32, u32 = C.c_void_p, C.c_int64, C.c_int, C.c_uint32
        L.merl_table_params_default.argtypes = [C.POINTER(TableParams)]
        L.merl_table_params_default.restype = None
        L.merl_table_create_from_file.argtypes = [C.c_char_p, C.POINTER(TableParams), i32, C.POINTER(vp)]
        L.merl_table_create_from_host.argtypes = [vp, C.POINTER(TableParams), i32, C.POINTER(vp)]
        L.merl_table_destroy.argtypes = [vp]
        L.merl_table_info.argtypes = [vp, C.POINTER(TableInfo)]
        L.merl_table_set_cdf.argtypes = [vp, vp]
        L.merl_table_download.argtypes = [vp, vp, vp, vp]
        L.merl_table_refresh_sampler.argtypes = [vp, vp]
        L.merl_table_update_cells.argtypes = [vp, vp, vp]
        L.merl_grad_clear_counts.argtypes = [vp, vp, vp]
        L.merl_file_report.argtypes = [C.c_char_p, C.POINTER(FileReport)]
        L.merl_table_save_packed.argtypes = [vp, C.c_char_p]
        L.merl_table_create_from_packed.argtypes = [C.c_char_p, i32, C.POINTER(vp)]
        L.merl_bins.argtypes = [vp] + [vp] * 6 + [vp, i64, vp, vp]
        L.merl_eval.argtypes = [vp] + [vp] * 6 + [vp, i64, vp, vp, vp, vp]
        L.merl_pdf.argtypes = [vp] + [vp] * 6 + [vp, i64, vp, vp]
        L.merl_eval_pdf.argtypes = [vp] + [vp] * 6 + [vp, i64, vp, vp, vp, vp, vp]
        L.merl_sample.argtypes = [vp] + [vp] * 6 + [vp, i64] + [vp] * 8 + [vp]
        L.merl_backward.argtypes = [vp] + [vp] * 9 + [vp, i64, vp, u32, vp]
        L.merl_eval_pdf_aos.argtypes = [vp, vp, vp, i64, vp, vp]
        L.merl_sample_aos.argtypes = [vp, vp, vp, i64, vp, vp, vp]
        L.merl_eval_pdf_host.argtypes = [vp] + [vp] * 6 + [i64] + [vp] * 4
        L.merl_sample_host.argtypes = [vp] + [vp] * 6 + [i64] + [vp] * 8
        L.merl_eval_pdf_host_static.argtypes = [vp] + [vp] * 6 + [i64] + [vp] * 4 + [vp]
        L.merl_set_sample_search.argtypes = [i32]
        L.merl_set_render_graphs.argtypes = [i32]
        L.merl_render.argtypes = [vp, i32, vp, i32, vp, vp, vp, vp]
        L.merl_comm_unique_id.argtypes = [vp]
        L.merl_comm_create.argtypes = [vp, i32, i32, i32, C.POINTER(vp)]
        L.merl_comm_destroy.argtypes = [vp]
        L.merl_allreduce_sum_f32.argtypes = [vp, vp, i64, vp]
        L.merl_grad_apply_mask.argtypes = [vp, vp, vp]
        L.merl_peerbuf_create.argtypes = [vp, i64, i32, C.POINTER(vp)]
        L.merl_peerbuf_destroy.argtypes = [vp]
        L.merl_peerbuf_info.argtypes = [vp, C.POINTER(PeerbufInfo)]
        L.merl_peerbuf_allreduce.argtypes = [vp, i64, vp, vp]
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise MerlError(f"libmerl_b200 error {rc}: {lib().merl_last_error().decode(errors='replace')}")


def default_params() -> TableParams:
    p = TableParams()
    lib().merl_table_params_default(C.byref(p))
    return p


def launch_count() -> int:
    return int(lib().merl_launch_count())
