"""ctypes binding of the TEST-ONLY entry point ``qsv_hostcheck_fused`` (include/qsv_hostcheck.h): the fused plan of a gate
program interpreted on the host, so the fusion pass can be checked against the oracle without a GPU.  Lives under tests/
on purpose -- the package never calls it."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from quantum_simulations_b200 import _lib  # noqa: E402

HOSTCHECK_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libqsv_hostcheck.so")   # built next to libqsv.so, never loaded by the package
_fn = None


def _load():
    global _fn
    if _fn is None:
        if not os.path.exists(HOSTCHECK_PATH):
            import __graft_entry__
            __graft_entry__.build()
        lib = C.CDLL(HOSTCHECK_PATH)
        _fn = lib.qsv_hostcheck_fused
        _fn.restype = C.c_int
        _fn.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p,
                        C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_char_p, C.c_uint64]
    return _fn


def run_fused_on_host(recs, pool, state, n_bits, tile_bits=0, dtype=np.complex128):
    """Apply flattened gate records to ``state`` (flat, 2^n_bits) through the fused plan, on the host.
    Returns (result, n_sweeps, n_fft_sweeps)."""
    fn = _load()
    recs = np.ascontiguousarray(recs, dtype=_lib.GATE_DTYPE)
    pool = np.ascontiguousarray(pool, dtype=np.complex128)
    out = np.ascontiguousarray(state, dtype=dtype).reshape(-1).copy()
    assert out.size == 1 << n_bits
    sweeps, fft = C.c_uint64(0), C.c_uint64(0)
    err = C.create_string_buffer(512)
    rc = fn(_lib.QSV_C128 if dtype == np.complex128 else _lib.QSV_C64, n_bits, recs.ctypes.data, len(recs),
            pool.ctypes.data if pool.size else None, pool.size, tile_bits, out.ctypes.data, C.byref(sweeps), C.byref(fft), err, 512)
    if rc:
        raise RuntimeError(f"qsv_hostcheck_fused rc={rc}: {err.value.decode()}")
    return out, int(sweeps.value), int(fft.value)


def run_fused_to_layout_on_host(recs, pool, state, n_bits, final_pos, carry_exchange=False, tile_bits=0):
    """Like run_fused_on_host for a plan with a requested end layout (logical bit b -> physical position final_pos[b])."""
    _load()
    lib = C.CDLL(HOSTCHECK_PATH)
    fn = lib.qsv_hostcheck_fused_layout
    fn.restype = C.c_int
    fn.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                   C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_char_p, C.c_uint64]
    recs = np.ascontiguousarray(recs, dtype=_lib.GATE_DTYPE)
    pool = np.ascontiguousarray(pool, dtype=np.complex128)
    out = np.ascontiguousarray(state, dtype=np.complex128).reshape(-1).copy()
    fp = np.ascontiguousarray(final_pos, dtype=np.int32)
    sweeps, fft = C.c_uint64(0), C.c_uint64(0)
    err = C.create_string_buffer(512)
    rc = fn(_lib.QSV_C128, n_bits, recs.ctypes.data, len(recs), pool.ctypes.data if pool.size else None, pool.size, tile_bits,
            fp.ctypes.data, int(carry_exchange), out.ctypes.data, C.byref(sweeps), C.byref(fft), err, 512)
    if rc:
        raise RuntimeError(f"qsv_hostcheck_fused_layout rc={rc}: {err.value.decode()}")
    return out, int(sweeps.value), int(fft.value)


def plan_shape(recs, pool, n_bits, tile_bits=0):
    """(n_sweeps, n_fft_sweeps) of the fused plan for a register of any size -- nothing is executed."""
    fn = _load()
    recs = np.ascontiguousarray(recs, dtype=_lib.GATE_DTYPE)
    pool = np.ascontiguousarray(pool, dtype=np.complex128)
    sweeps, fft = C.c_uint64(0), C.c_uint64(0)
    err = C.create_string_buffer(512)
    rc = fn(_lib.QSV_C128, n_bits, recs.ctypes.data, len(recs), pool.ctypes.data if pool.size else None, pool.size, tile_bits,
            None, C.byref(sweeps), C.byref(fft), err, 512)
    if rc:
        raise RuntimeError(f"qsv_hostcheck_fused rc={rc}: {err.value.decode()}")
    return int(sweeps.value), int(fft.value)
