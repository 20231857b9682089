"""Equivalence of two gate streams, decided on the GPU.

The reference's optimiser test multiplies every leaf into a dense matrix on the host before and after and compares
(``tests/qompiler/optimize/test_optimizer.py:165-169``).  Here both streams are run on all basis columns by the CUDA
engine and compared where they lie (``qsv_compare`` / ``qsv_allprop``): nothing but one flag and one number comes back.
"""
import numpy as np

from ..construct.qspace import qubit_key
from ..program import flatten


def _leaves(x):
    """Time-ordered gate list of a Bytecode tree (or a list already in time order)."""
    if hasattr(x, "children"):
        from ..stream import iter_leaves
        return [g for g in iter_leaves(x) if hasattr(g, "controls")]
    return list(x)


def equal_on_device(before, after, register=None, up_to_phase=False, rtol=1e-5, atol=1e-8, dtype=np.complex128, info=None):
    """True iff the two streams (trees or time-ordered gate lists) are the same operator over ``register`` (default: the
    sorted union of their qubits; ancillas count as ordinary qubits).  Same tolerances as ``np.allclose`` /
    ``allprop`` (``quompiler/utils/mfun.py:34-52``)."""
    from ..engine import get_engine, simulate_columns
    a, b = _leaves(before), _leaves(after)
    if register is None:
        register = sorted({q for g in a + b for q in g.qspace}, key=qubit_key)
    plain = {q: type(q)(i) for i, q in enumerate(register)}   # ancillas as ordinary qubits: compare the full operator
    eng = get_engine(dtype)
    outs = []
    for gates in (a, b):
        prog = flatten([], [plain[q] for q in register])
        for g in gates:
            prog.add_ctrl_gate(_renamed(g, plain))
        n = prog.n_qubits
        outs.append(simulate_columns(prog, 0, n, dtype=dtype, engine=eng, drop_ancilla=False, return_device=True).reshape(-1))
    if up_to_phase:
        ok, lam = eng.allprop(outs[0], outs[1], rtol, atol)
        if info is not None:
            info["ratio"] = lam
        return bool(ok)
    max_abs, bad = eng.compare(outs[0], outs[1], rtol, atol)
    if info is not None:
        info["max_abs_diff"] = max_abs
        info["violations"] = bad
    return bad == 0


def _renamed(g, plain):
    from ..construct.cgate import CtrlGate
    return CtrlGate(g.matrix(), g.controls(), [plain[q] for q in g.qspace], g.phase())


def optimize_checked(root, optimizer, **kw):
    """Run ``optimizer.optimize(root)`` and verify on the GPU that the operator did not change; raises otherwise."""
    before = _leaves(root)
    out = optimizer.optimize(root)
    info = {}
    if not equal_on_device(before, out, info=info, **kw):
        raise AssertionError(f"optimiser changed the operator: {info}")
    return out
