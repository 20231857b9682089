"""ctypes binding of ``libqsv.so`` (C ABI declared in ``include/qsv.h``).

There is no fallback: if the shared library is missing or the CUDA device is
absent, every compute entry point raises ``QsvError``.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libqsv.so")

QSV_C128, QSV_C64 = 0, 1
DENSE1, PERMX, DIAG1, DENSEK, ROWS2 = 0, 1, 2, 3, 4
MAX_TARGETS = 5

# numpy mirror of qsv_gate_t (64 bytes)
GATE_DTYPE = np.dtype([
    ("kind", np.uint32), ("n_targets", np.uint32), ("target_bits", np.uint8, (8,)),
    ("ctrl_mask", np.uint64), ("ctrl_value", np.uint64), ("row0", np.uint64), ("row1", np.uint64),
    ("matrix_offset", np.uint64), ("reserved", np.uint64)], align=True)
assert GATE_DTYPE.itemsize == 64


class QsvStats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("sweeps", C.c_uint64), ("amp_bytes", C.c_uint64),
                ("gates_applied", C.c_uint64), ("fused_sweeps", C.c_uint64), ("reserved", C.c_uint64 * 3)]


class QsvError(RuntimeError):
    pass


_SIGNATURES = {
    "qsv_version": (C.c_char_p, []),
    "qsv_last_error": (C.c_char_p, [C.c_void_p]),
    "qsv_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int]),
    "qsv_destroy": (None, [C.c_void_p]),
    "qsv_get_stats": (C.c_int, [C.c_void_p, C.POINTER(QsvStats)]),
    "qsv_reset_stats": (C.c_int, [C.c_void_p]),
    "qsv_apply_gate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "qsv_apply_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "qsv_upload_program": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]),
    "qsv_fuse": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]),
    "qsv_run": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.c_void_p]),
    "qsv_fuse_to_layout": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]),
    "qsv_run_exchange": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "qsv_plan_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "qsv_set_profiling": (C.c_int, [C.c_void_p, C.c_int]),
    "qsv_get_sweep_profile": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_int)]),
    "qsv_init_basis_columns": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.c_void_p]),
    "qsv_project_ancilla": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_void_p]),
    "qsv_init_basis_columns_psi": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.POINTER(C.c_double), C.c_void_p]),
    "qsv_project_ancilla_psi": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.POINTER(C.c_double), C.c_void_p]),
    "qsv_permute_bits": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "qsv_permute_exchange": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "qsv_compare": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_double, C.c_double,
                              C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.c_void_p]),
    "qsv_allprop": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_double, C.c_double, C.POINTER(C.c_int),
                              C.POINTER(C.c_double), C.c_void_p]),
    "qsv_compare_noalloc": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_double, C.c_double,
                                      C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.c_void_p]),
}
NOALLOC_MAX_AMPS = 1 << 28

_lib = None


def exported_symbols():
    return sorted(_SIGNATURES)


def load():
    """Load libqsv.so once; raise QsvError (never fall back) when it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise QsvError(f"{LIB_PATH} is not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                       f"(or make -C quantum-simulations_b200/csrc). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header and library out of sync
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, ctx=None):
    if rc != 0:
        msg = load().qsv_last_error(ctx)
        raise QsvError(f"libqsv error {rc}: {msg.decode() if msg else '?'}")
