"""ctypes wrapper of oracle/liborc.so (C + OpenMP restatement).  TEST INFRASTRUCTURE / CPU BASELINE ONLY."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "liborc.so")
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            import subprocess
            subprocess.check_call(["make", "-C", _HERE])
        _lib = C.CDLL(_PATH)
        _lib.orc_run.restype = C.c_int
        _lib.orc_run.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p]
        _lib.orc_set_threads.argtypes = [C.c_int]
        _lib.orc_get_threads.restype = C.c_int
    return _lib


def run_records(recs, pool, state, n_bits, batch_bits=0):
    """In place on a contiguous complex128 array of 2^n_bits amplitudes."""
    assert state.dtype == np.complex128 and state.flags.c_contiguous and state.size == 1 << n_bits
    recs = np.ascontiguousarray(recs)
    pool = np.ascontiguousarray(pool, dtype=np.complex128)
    rc = load().orc_run(state.ctypes.data_as(C.c_void_p), n_bits, batch_bits, recs.ctypes.data_as(C.c_void_p), len(recs),
                        pool.ctypes.data_as(C.c_void_p))
    if rc:
        raise RuntimeError("orc_run failed")
    return state


def set_threads(n: int = 0) -> int:
    """Use ``n`` OpenMP threads (0 = every core this process may run on), whatever OMP_NUM_THREADS says --
    torch.distributed.run exports OMP_NUM_THREADS=1 to its workers.  Returns the count in effect."""
    if n <= 0:
        try:
            n = len(os.sched_getaffinity(0))
        except AttributeError:
            n = os.cpu_count() or 1
    load().orc_set_threads(int(n))
    return threads()


def threads() -> int:
    """The thread count the OpenMP loops will really use."""
    return int(load().orc_get_threads())
