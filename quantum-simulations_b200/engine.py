"""Host driver of the CUDA statevector engine.

PyTorch is used for plumbing only -- device buffers, the current stream and (in
``sharded.py``) ``torch.distributed``; every amplitude is touched by the kernels
of ``libqsv.so`` through the C ABI of ``include/qsv.h``.  Nothing here evaluates
gates on the CPU: without the library or without a GPU the calls raise.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import QSV_C128, QSV_C64, QsvError, QsvStats, check
from .program import Program, flatten, flatten_two_level


def _torch():
    import torch
    return torch


def _require_cuda():
    torch = _torch()
    if not torch.cuda.is_available():
        raise QsvError("no CUDA device: the B200 statevector engine has no CPU path")
    return torch


_NP2QSV = {np.dtype(np.complex128): QSV_C128, np.dtype(np.complex64): QSV_C64}


def _torch_dtype(np_dtype):
    torch = _torch()
    return torch.complex128 if np.dtype(np_dtype) == np.complex128 else torch.complex64


class Engine:
    """One libqsv context on one device for one amplitude type."""

    def __init__(self, device=None, dtype=np.complex128):
        torch = _require_cuda()
        self.lib = _lib.load()
        self.np_dtype = np.dtype(dtype)
        if self.np_dtype not in _NP2QSV:
            raise ValueError("dtype must be complex128 or complex64")
        self.device = torch.cuda.current_device() if device is None else int(device)
        self._ctx = C.c_void_p()
        check(self.lib.qsv_create(C.byref(self._ctx), self.device, _NP2QSV[self.np_dtype]), None)
        self._program = None
        self._keep = None

    def close(self):
        if self._ctx:
            self.lib.qsv_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- helpers ---------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(_torch().cuda.current_stream(self.device).cuda_stream)

    def empty(self, n_amps: int):
        torch = _torch()
        return torch.empty(int(n_amps), dtype=_torch_dtype(self.np_dtype), device=f"cuda:{self.device}")

    def from_host(self, array):
        """Copy a host array to this engine's device as a flat tensor of the engine's amplitude type."""
        flat = np.ascontiguousarray(array, dtype=self.np_dtype).reshape(-1)
        return _torch().from_numpy(flat).to(f"cuda:{self.device}")

    def synchronize(self):
        _torch().cuda.synchronize(self.device)

    def stats(self) -> dict:
        s = QsvStats()
        check(self.lib.qsv_get_stats(self._ctx, C.byref(s)), self._ctx)
        return {k: int(getattr(s, k)) for k, _ in QsvStats._fields_ if k != "reserved"}

    def reset_stats(self):
        check(self.lib.qsv_reset_stats(self._ctx), self._ctx)

    # -- programs --------------------------------------------------------------------
    def upload(self, program: Program, batch_bits: int, n_bits: int = None):
        recs = program.records(batch_bits)
        pool = program.pool()
        if n_bits is None:
            n_bits = program.n_qubits + batch_bits
        self._keep = (recs, pool)
        check(self.lib.qsv_upload_program(self._ctx, n_bits, batch_bits, recs.ctypes.data_as(C.c_void_p), len(recs),
                                          pool.ctypes.data_as(C.c_void_p), pool.size), self._ctx)
        self._program = program
        self._n_bits = n_bits
        self._batch_bits = batch_bits

    def upload_records(self, recs: np.ndarray, pool: np.ndarray, n_bits: int, batch_bits: int = 0):
        recs = np.ascontiguousarray(recs, dtype=_lib.GATE_DTYPE)
        pool = np.ascontiguousarray(pool, dtype=np.complex128)
        self._keep = (recs, pool)
        check(self.lib.qsv_upload_program(self._ctx, n_bits, batch_bits, recs.ctypes.data_as(C.c_void_p), len(recs),
                                          pool.ctypes.data_as(C.c_void_p), pool.size), self._ctx)
        self._n_bits = n_bits
        self._batch_bits = batch_bits

    def fuse(self, max_tile_bits: int = 0) -> int:
        n = C.c_uint64(0)
        check(self.lib.qsv_fuse(self._ctx, max_tile_bits, C.byref(n)), self._ctx)
        return int(n.value)

    def fuse_to_layout(self, final_pos, carry_exchange=False, max_tile_bits: int = 0) -> int:
        """:meth:`fuse` with a requested end layout (logical bit b -> physical position ``final_pos[b]``)."""
        n = C.c_uint64(0)
        p = np.ascontiguousarray(final_pos, dtype=np.int32)
        check(self.lib.qsv_fuse_to_layout(self._ctx, max_tile_bits, p.ctypes.data_as(C.c_void_p), int(bool(carry_exchange)), C.byref(n)),
              self._ctx)
        return int(n.value)

    def run_exchange(self, state, spare, peer_ptrs, g_bits: int, rank: int):
        """Run a plan built with ``carry_exchange``: its last sweep stores into the peers' buffers."""
        arr = (C.c_void_p * len(peer_ptrs))(*peer_ptrs)
        sp = C.c_void_p(spare.data_ptr()) if spare is not None else C.c_void_p()
        check(self.lib.qsv_run_exchange(self._ctx, C.c_void_p(state.data_ptr()), sp, arr, g_bits, rank, self._stream()), self._ctx)

    def plan_info(self):
        """(fused sweeps, out-of-place sweeps) of the plan built by :meth:`fuse`."""
        a, b = C.c_uint64(0), C.c_uint64(0)
        check(self.lib.qsv_plan_info(self._ctx, C.byref(a), C.byref(b)), self._ctx)
        return int(a.value), int(b.value)

    def run(self, state, spare=None):
        """Apply the uploaded program; returns the tensor that holds the result."""
        flag = C.c_int(0)
        sp = C.c_void_p(spare.data_ptr()) if spare is not None else C.c_void_p()
        check(self.lib.qsv_run(self._ctx, C.c_void_p(state.data_ptr()), sp, C.byref(flag), self._stream()), self._ctx)
        return spare if flag.value else state

    def set_profiling(self, on: bool):
        check(self.lib.qsv_set_profiling(self._ctx, int(bool(on))), self._ctx)

    def sweep_profile(self, cap: int = 4096):
        """CUDA-event durations (ms) of the fused sweeps launched since the last call."""
        buf = (C.c_float * cap)()
        n = C.c_int(0)
        check(self.lib.qsv_get_sweep_profile(self._ctx, buf, cap, C.byref(n)), self._ctx)
        return [float(buf[i]) for i in range(min(cap, n.value))]

    # -- single operations -----------------------------------------------------------
    def apply_gate_record(self, state, n_bits: int, rec, matrix):
        rec = np.ascontiguousarray(rec, dtype=_lib.GATE_DTYPE)
        m = None if matrix is None else np.ascontiguousarray(matrix, dtype=np.complex128)
        check(self.lib.qsv_apply_gate(self._ctx, C.c_void_p(state.data_ptr()), n_bits, rec.ctypes.data_as(C.c_void_p),
                                      m.ctypes.data_as(C.c_void_p) if m is not None else C.c_void_p(), self._stream()), self._ctx)

    def apply_rows(self, state, dim: int, batch: int, rows, matrix):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        m = np.ascontiguousarray(matrix, dtype=np.complex128)
        check(self.lib.qsv_apply_rows(self._ctx, C.c_void_p(state.data_ptr()), dim, batch, rows.ctypes.data_as(C.c_void_p),
                                      len(rows), m.ctypes.data_as(C.c_void_p), self._stream()), self._ctx)

    @staticmethod
    def _psi(ancilla_state):
        """None (= |0>) or the four doubles (re a, im a, re b, im b) of a|0> + b|1>."""
        if ancilla_state is None:
            return None
        st = np.asarray(ancilla_state, dtype=np.complex128)
        assert st.shape == (2,), f'state vector must be a 1D array of length 2, but got {st.shape}'
        if st[0] == 1 and st[1] == 0:
            return None
        return (C.c_double * 4)(st[0].real, st[0].imag, st[1].real, st[1].imag)

    def init_basis_columns(self, state, n_bits: int, batch_bits: int, col0: int = 0, ancilla_mask: int = 0, ancilla_state=None):
        psi = self._psi(ancilla_state)
        if psi is None or not ancilla_mask:
            check(self.lib.qsv_init_basis_columns(self._ctx, C.c_void_p(state.data_ptr()), n_bits, batch_bits, col0, ancilla_mask,
                                                  self._stream()), self._ctx)
        else:
            check(self.lib.qsv_init_basis_columns_psi(self._ctx, C.c_void_p(state.data_ptr()), n_bits, batch_bits, col0,
                                                      ancilla_mask, psi, self._stream()), self._ctx)

    def project_ancilla(self, src, n_bits: int, batch_bits: int, ancilla_mask: int, ancilla_state=None):
        n_out = (1 << n_bits) >> bin(ancilla_mask).count("1")
        out = self.empty(n_out)
        psi = self._psi(ancilla_state)
        if psi is None:
            check(self.lib.qsv_project_ancilla(self._ctx, C.c_void_p(src.data_ptr()), C.c_void_p(out.data_ptr()), n_bits,
                                               batch_bits, ancilla_mask, self._stream()), self._ctx)
        else:
            check(self.lib.qsv_project_ancilla_psi(self._ctx, C.c_void_p(src.data_ptr()), C.c_void_p(out.data_ptr()), n_bits,
                                                   batch_bits, ancilla_mask, psi, self._stream()), self._ctx)
        return out

    def permute_bits(self, src, n_bits: int, perm):
        perm = np.ascontiguousarray(perm, dtype=np.int32)
        out = self.empty(1 << n_bits)
        check(self.lib.qsv_permute_bits(self._ctx, C.c_void_p(src.data_ptr()), C.c_void_p(out.data_ptr()), n_bits,
                                        perm.ctypes.data_as(C.c_void_p), self._stream()), self._ctx)
        return out

    def compare(self, a, b, rtol=1e-5, atol=1e-8):
        """(max |a-b|, number of entries violating |a-b| <= atol + rtol*|b|), computed on the device."""
        assert a.numel() == b.numel()
        mx = C.c_double(0)
        bad = C.c_uint64(0)
        check(self.lib.qsv_compare(self._ctx, C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), a.numel(), rtol, atol,
                                   C.byref(mx), C.byref(bad), self._stream()), self._ctx)
        return mx.value, int(bad.value)

    def allprop(self, a, b, rtol=1e-5, atol=1e-8):
        """(ok, ratio): is ``a == ratio * b`` for one complex scalar?  Device version of the reference's allprop."""
        assert a.numel() == b.numel()
        ok = C.c_int(0)
        ratio = (C.c_double * 2)()
        check(self.lib.qsv_allprop(self._ctx, C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), a.numel(), rtol, atol,
                                   C.byref(ok), ratio, self._stream()), self._ctx)
        return bool(ok.value), complex(ratio[0], ratio[1])

    def allclose(self, a, b, rtol=1e-5, atol=1e-8) -> bool:
        return self.compare(a, b, rtol, atol)[1] == 0


_ENGINES = {}
_WARNED = set()


def _try_fuse(eng) -> bool:
    """Build the fused plan; when the fusion pass refuses the program ("unsupported": e.g. fewer index bits than a
    register stage) the program still runs on the GPU, one kernel per gate -- many times slower on long streams, so
    say so once per reason instead of downgrading silently.  Any other error propagates."""
    import warnings
    try:
        eng.fuse()
        return True
    except QsvError as e:
        msg = str(e)
        if "unsupported" not in msg.lower():
            raise
        if msg not in _WARNED:
            _WARNED.add(msg)
            warnings.warn(f"fused sweeps unavailable for this program, running one kernel per gate: {msg}", RuntimeWarning, stacklevel=3)
        return False


def get_engine(dtype=np.complex128, device=None) -> Engine:
    torch = _require_cuda()
    dev = torch.cuda.current_device() if device is None else int(device)
    key = (dev, np.dtype(dtype))
    if key not in _ENGINES:
        _ENGINES[key] = Engine(dev, dtype)
    return _ENGINES[key]


def _pow2_ceil_bits(n: int) -> int:
    return max(0, (int(n) - 1).bit_length())


def simulate_columns(program: Program, col0: int, n_cols_bits: int, dtype=np.complex128, fuse=True, engine=None,
                     drop_ancilla=True, return_device=False, ancilla_state=None):
    """Run ``program`` on the basis columns ``col0 .. col0 + 2^n_cols_bits`` (ancillas in |0>, or in the
    single-qubit pure state ``ancilla_state``, which is then also the state they are projected on).

    Returns the ``[rows][columns]`` block, rows restricted to ancilla bits = 0 when
    ``drop_ancilla``.  This is the column-batched unitary reconstruction that replaces
    ``reduce(@)`` + ``dela`` + ``inflate``.
    """
    eng = engine or get_engine(dtype)
    torch = _torch()
    n = program.n_qubits
    b = n_cols_bits
    n_bits = n + b
    anc = program.ancilla_mask
    state = eng.empty(1 << n_bits)
    eng.init_basis_columns(state, n_bits, b, col0, anc, ancilla_state)
    eng.upload(program, b)
    spare = None
    if fuse and _try_fuse(eng):
        spare = eng.empty(1 << n_bits)
    res = eng.run(state, spare)
    rows_bits = n
    if drop_ancilla and anc:
        res = eng.project_ancilla(res, n_bits, b, anc, ancilla_state)
        rows_bits = n - bin(anc).count("1")
    if return_device:
        return res.view(1 << rows_bits, 1 << b)
    torch.cuda.synchronize(eng.device)
    return res.view(1 << rows_bits, 1 << b).cpu().numpy()


def reconstruct_unitary(gates_time_order, register=None, drop_ancilla=False, dtype=np.complex128, fuse=True, ancilla_state=None):
    """Dense operator of a CtrlGate stream over ``register`` (default: sorted union)."""
    program = flatten(gates_time_order, register)
    anc = program.ancilla_mask
    if drop_ancilla or not anc:
        n_free = program.n_qubits - bin(anc).count("1")
        return simulate_columns(program, 0, n_free, dtype=dtype, fuse=fuse, drop_ancilla=True, ancilla_state=ancilla_state)
    # full operator including ancilla=1 columns: treat ancillas as ordinary qubits
    eng = get_engine(dtype)
    n = program.n_qubits
    state = eng.empty(1 << (2 * n))
    eng.init_basis_columns(state, 2 * n, n, 0, 0)
    eng.upload(program, n)
    res = eng.run(state, None)
    _torch().cuda.synchronize(eng.device)
    return res.view(1 << n, 1 << n).cpu().numpy()


def reconstruct_unitary_rows(unitaries_time_order, dimension: int, dtype=np.complex128):
    """Dense operator of a UnitaryM stream of arbitrary (not necessarily 2^n) dimension."""
    eng = get_engine(dtype)
    torch = _torch()
    d = int(dimension)
    state = torch.eye(d, dtype=_torch_dtype(dtype), device=f"cuda:{eng.device}").reshape(-1).contiguous()
    for u in unitaries_time_order:
        if u.dimension != d:
            raise ValueError('matmul: Input operands have dimension mismatch.')
        eng.apply_rows(state, d, d, u.core, np.asarray(u.matrix, dtype=np.complex128) * complex(u.phase))
    torch.cuda.synchronize(eng.device)
    return state.view(d, d).cpu().numpy()


def apply_circuit(program: Program, states: np.ndarray, dtype=np.complex128, fuse=True, engine=None):
    """Apply ``program`` to host state vectors ``[2^n]`` or ``[2^n, B]``; returns the same shape."""
    eng = engine or get_engine(dtype)
    torch = _torch()
    states = np.asarray(states)
    single = states.ndim == 1
    x = states.reshape(states.shape[0], -1)
    dim, batch = x.shape
    n = program.n_qubits
    if dim != 1 << n:
        raise ValueError(f"state has {dim} rows, register needs {1 << n}")
    b = _pow2_ceil_bits(batch)
    padded = np.zeros((dim, 1 << b), dtype=eng.np_dtype)
    padded[:, :batch] = x
    dev = torch.from_numpy(padded.reshape(-1)).to(f"cuda:{eng.device}")
    eng.upload(program, b)
    spare = None
    if fuse and _try_fuse(eng):
        spare = torch.empty_like(dev)
    res = eng.run(dev, spare)
    torch.cuda.synchronize(eng.device)
    out = res.view(dim, 1 << b).cpu().numpy()[:, :batch]
    return out[:, 0].copy() if single else np.ascontiguousarray(out)
