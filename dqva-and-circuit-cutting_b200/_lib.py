"""ctypes binding of libqdc_b200.so (the C-ABI declared in include/qdc_b200.h).

There is deliberately no fallback: if the shared library is missing or a call fails, a
`QdcError` is raised.  Build it with `python __graft_entry__.py build` (or `make -C csrc`).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libqdc_b200.so")

c_i64 = C.c_int64
c_u64p = C.POINTER(C.c_uint64)
vp = C.c_void_p


class QdcError(RuntimeError):
    pass


# name -> (restype, argtypes); must list every symbol include/qdc_b200.h declares
SIGNATURES = {
    "qdc_version": (C.c_int, []),
    "qdc_last_error": (C.c_char_p, []),
    "qdc_device_info": (C.c_int, [C.POINTER(C.c_int)] * 3),
    "qdc_mis_build_tables": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_int), c_u64p, C.c_int,
                                       C.c_uint64, C.c_uint64, vp, vp, vp, vp, vp, vp]),
    "qdc_sv_compile": (C.c_int, [C.c_int, vp, C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_int,
                                 C.POINTER(vp)]),
    "qdc_sv_free": (None, [vp]),
    "qdc_sv_info": (C.c_int, [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                              C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_uint64)]),
    "qdc_sv_dump": (C.c_int, [vp] * 9),
    "qdc_sv_run": (C.c_int, [vp, vp, c_i64, c_i64, c_i64, vp, C.c_int, C.POINTER(C.c_int32),
                             C.c_int, C.POINTER(C.c_double), C.c_int, vp, vp]),
    "qdc_sv_score_workspace_bytes": (c_i64, [vp, c_i64]),
    "qdc_sv_run_score": (C.c_int, [vp, vp, c_i64, vp, C.c_int, vp, vp, C.c_uint64, C.c_int, vp, vp, vp]),
    "qdc_sv_probs": (C.c_int, [vp, C.c_int, c_i64, C.c_uint64, vp, vp]),
    "qdc_sv_sample_workspace_doubles": (c_i64, [C.c_int, c_i64]),
    "qdc_sv_sample_counts": (C.c_int, [vp, C.c_int, c_i64, c_i64, C.c_uint64, c_i64, C.c_uint64, C.c_uint64,
                                       C.c_uint64, vp, vp, c_i64, vp, vp]),
    "qdc_sv_pauli_tensor": (C.c_int, [vp, C.c_int, c_i64, C.c_uint64, C.c_int, C.c_int,
                                      C.POINTER(C.c_double), C.POINTER(C.c_int32), C.c_int,
                                      C.POINTER(c_i64), vp, c_i64, vp]),
    "qdc_ml_workspace_bytes": (c_i64, [c_i64, C.c_int]),
    "qdc_ml_project": (C.c_int, [vp, c_i64, C.c_int, c_i64, vp, vp, vp, vp, vp, c_i64, vp, C.c_int, vp]),
    "qdc_reduce_partials": (C.c_int, []),
    "qdc_mis_reduce": (C.c_int, [vp, C.c_int, c_i64, c_i64, c_i64, vp, C.c_double,
                                 vp, vp, vp, vp, vp, vp, vp, vp, vp, c_i64, vp, vp]),
    "qdc_topk_select": (C.c_int, [vp, vp, vp, c_i64, C.c_int, vp, vp, vp]),
    "qdc_contract_packed_doubles": (c_i64, [c_i64, C.c_int]),
    "qdc_contract_pack": (C.c_int, [vp, c_i64, C.c_int, vp, vp]),
    "qdc_contract_f64": (C.c_int, [vp, c_i64, vp, c_i64, c_i64, c_i64, C.c_int, C.c_double, vp,
                                   C.c_double, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int,
                                   vp, vp, vp, vp, c_i64, vp, vp]),
    "qdc_penalty_workspace_bytes": (C.c_int, []),
    "qdc_penalty_reduce": (C.c_int, [vp, C.c_int, C.c_int, C.c_uint64, vp, C.c_int, C.c_uint64, vp, vp, vp, vp]),
    "qdc_i8_operand_bytes": (c_i64, [c_i64, C.c_int, C.c_int]),
    "qdc_i8_prepare": (C.c_int, [vp, c_i64, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp]),
    "qdc_colsum_partials": (C.c_int, []),
    "qdc_colsum": (C.c_int, [vp, c_i64, c_i64, c_i64, c_i64, C.c_int, vp, vp]),
    "qdc_contract_i8_prepared": (C.c_int, [vp, c_i64, vp, c_i64, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int,
                                           C.c_double, C.c_int, vp, vp, C.c_int, C.c_double,
                                           vp, vp, c_i64, vp, vp, vp, vp, C.c_int,
                                           vp, vp, vp, c_i64, C.c_int, vp, vp, vp, vp]),
    "qdc_cutqc_measure": (C.c_int, [vp, C.c_int, C.c_int, c_i64, C.c_uint64, vp, vp, vp]),
    "qdc_cutqc_collapse": (C.c_int, [vp, c_i64, C.c_int, vp, vp, vp, vp, vp, vp]),
    "qdc_cutqc_merge": (C.c_int, [vp, c_i64, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, vp, vp]),
    "qdc_cutqc_gather": (C.c_int, [vp, c_i64, vp, C.c_int, C.c_int, vp, vp]),
    "qdc_khatri_rao": (C.c_int, [vp, c_i64, vp, c_i64, C.c_int, vp, vp]),
    "qdc_norm_workspace_doubles": (C.c_int, [C.c_int]),
    "qdc_contract_norm": (C.c_int, [vp, c_i64, vp, c_i64, C.c_int, C.c_double, vp, vp, vp]),
    "qdc_contract_norm_static_max_kdim": (C.c_int, []),
    "qdc_contract_norm_static": (C.c_int, [vp, c_i64, vp, c_i64, C.c_int, C.c_double, vp, vp]),
}

_lib = None


def load():
    """Load the library once; raise QdcError (never fall back) if that is impossible."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise QdcError(f"{LIB_PATH} not found: the CUDA extension is not built "
                       "(run `python __graft_entry__.py build`); there is no CPU fallback")
    try:
        lib = C.CDLL(LIB_PATH)
    except OSError as exc:
        raise QdcError(f"cannot load {LIB_PATH}: {exc}") from exc
    for name, (restype, argtypes) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as exc:
            raise QdcError(f"{LIB_PATH} does not export {name}") from exc
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().qdc_last_error()
        raise QdcError(f"{what} failed with code {rc}: {msg.decode() if msg else ''}")


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(stream=None):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)
