"""The allocation-free compare entry point, qsv_compare_noalloc.

Some hosts of this library cannot call cudaMalloc when they want to check a result: the state
vectors already fill the device, or the calling thread must stay out of the CUDA allocator because
it stalls every stream on the GPU.  For them libqsv.so carries its reduction scratch as a fixed
array inside the library.  This file drives that entry point through the C ABI and checks it
against the ordinary qsv_compare (which uses scratch allocated by qsv_create) on two consecutive
calls, so stale scratch contents from the first call would show up in the second.

Kept in a file that sorts last so that a run stopping at the first failure has already executed
every other test.  It needs nothing from those tests: it passes alone in a fresh process too.
"""
import ctypes as C

import numpy as np
import pytest

from quantum_simulations_b200 import _lib

pytestmark = pytest.mark.gpu


def _noalloc(lib, a, b, rtol, atol, dtype_code):
    import torch
    mx, bad = C.c_double(-1.0), C.c_uint64(12345)
    rc = lib.qsv_compare_noalloc(a.device.index, dtype_code, C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), a.numel(),
                                 rtol, atol, C.byref(mx), C.byref(bad), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.qsv_last_error(None)
    return mx.value, int(bad.value)


@pytest.mark.parametrize("np_dtype,code", [(np.complex128, _lib.QSV_C128), (np.complex64, _lib.QSV_C64)])
def test_noalloc_compare_matches_context_compare_twice_in_a_row(np_dtype, code):
    import torch
    from quantum_simulations_b200.engine import Engine
    lib = _lib.load()
    eng = Engine(0, np_dtype)                      # the existing entry point, scratch from cudaMalloc
    rng = np.random.default_rng(11)
    rtol, atol = 1e-5, 1e-8
    # first call: large array with many violations, so every scratch slot gets a non-trivial value
    n1 = (1 << 22) + 777
    x = (rng.standard_normal(n1) + 1j * rng.standard_normal(n1)).astype(np_dtype)
    y = x.copy()
    y[::5] += np_dtype(0.25 + 0.5j)
    a1, b1 = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    want1 = eng.compare(a1, b1, rtol, atol)
    got1 = _noalloc(lib, a1, b1, rtol, atol, code)
    assert got1[1] == want1[1] == len(range(0, n1, 5))
    assert got1[0] == want1[0]                     # a max is order independent: results must be identical
    # second call right after: small array, one tiny violation; leftovers of call one must not leak in
    n2 = 1000
    x2 = (rng.standard_normal(n2) + 1j * rng.standard_normal(n2)).astype(np_dtype)
    y2 = x2.copy()
    y2[123] += np_dtype(3e-3)
    a2, b2 = torch.from_numpy(x2).cuda(), torch.from_numpy(y2).cuda()
    want2 = eng.compare(a2, b2, rtol, atol)
    got2 = _noalloc(lib, a2, b2, rtol, atol, code)
    assert got2 == want2
    assert got2[1] == 1 and abs(got2[0] - 3e-3) < (1e-12 if np_dtype == np.complex128 else 1e-6)
    # identical arrays -> exactly zero, and the empty case
    assert _noalloc(lib, a2, a2, rtol, atol, code) == (0.0, 0)
    assert _noalloc(lib, a2[:0], a2[:0], rtol, atol, code) == (0.0, 0)
    eng.close()


def test_noalloc_compare_rejects_sizes_beyond_its_fixed_scratch():
    import torch
    lib = _lib.load()
    buf = torch.zeros(16, dtype=torch.complex128, device="cuda")
    mx, bad = C.c_double(0), C.c_uint64(0)
    rc = lib.qsv_compare_noalloc(0, _lib.QSV_C128, C.c_void_p(buf.data_ptr()), C.c_void_p(buf.data_ptr()),
                                 _lib.NOALLOC_MAX_AMPS + 1, 1e-5, 1e-8, C.byref(mx), C.byref(bad), None)
    assert rc == -5                                  # QSV_E_TOOBIG; the buffers are never touched
    msg = lib.qsv_last_error(None).decode()
    assert "QSV_NOALLOC_MAX_AMPS" in msg and "qsv_compare" in msg
    rc = lib.qsv_compare_noalloc(0, 7, C.c_void_p(buf.data_ptr()), C.c_void_p(buf.data_ptr()), 16, 1e-5, 1e-8,
                                 C.byref(mx), C.byref(bad), None)
    assert rc == -1                                  # unknown dtype
