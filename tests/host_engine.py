"""Oracle-backed stand-ins for the device entry points the Python layer calls (TEST ONLY).  They let the logic that
sits ABOVE the engine -- argument plumbing, buffer bookkeeping, result handling of ``QasmCircuit.unitary``,
``optimize.verify`` and friends -- run on a CPU box; the device work itself is covered by the ``-m gpu`` tests."""
import numpy as np
import torch

from oracle import sv_oracle as orc


def simulate_columns(program, col0, n_cols_bits, dtype=np.complex128, fuse=True, engine=None, drop_ancilla=True,
                     return_device=False):
    n, b, anc = program.n_qubits, n_cols_bits, program.ancilla_mask
    keep = [i for i in range(1 << n) if not (i & anc)]
    st = np.zeros((1 << n, 1 << b), dtype=np.complex128)
    for c in range(1 << b):
        st[keep[col0 + c], c] = 1
    flat = st.reshape(-1)
    orc.run_records(program.records(b), program.pool(), flat, n + b)
    out = flat.reshape(1 << n, 1 << b)
    if drop_ancilla and anc:
        out = out[keep, :]
    out = np.ascontiguousarray(out.astype(dtype))
    return torch.from_numpy(out) if return_device else out


class HostCompareEngine:
    def compare(self, a, b, rtol=1e-5, atol=1e-8):
        a, b = a.numpy().reshape(-1), b.numpy().reshape(-1)
        d = np.abs(a - b)
        return float(d.max()), int((d > atol + rtol * np.abs(b)).sum())

    def allprop(self, a, b, rtol=1e-5, atol=1e-8):
        from quantum_simulations_b200.construct.types import proportional
        ok, lam = proportional(a.numpy().reshape(-1), b.numpy().reshape(-1), rtol, atol)
        return bool(ok), complex(lam)


def install(monkeypatch):
    import quantum_simulations_b200.engine as E
    monkeypatch.setattr(E, "simulate_columns", simulate_columns)
    monkeypatch.setattr(E, "get_engine", lambda *a, **k: HostCompareEngine())
