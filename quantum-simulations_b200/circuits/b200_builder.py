"""``B200Builder``: the CUDA statevector engine behind the ``CircuitBuilder`` plug-in API.

Drop-in sibling of the reference's ``CirqBuilder``
(``quompiler/circuits/cirq_factory/cirq_builder.py:15-79``): it is selected by
``QompilePlatform.B200``, fed by ``QRenderer.render`` in circuit time order
(``quompiler/circuits/render.py:21-32``), and ``finish()`` returns a circuit object
offering ``unitary(qubit_order)`` like ``cirq.Circuit.unitary`` does at
``tests/qompiler/circuits/test_cirq_circuit.py:39`` -- plus ``apply(states)`` for
state-vector batches.  Unlike the Cirq builder it accepts ``UnitaryM`` gates too
(``cirq_builder.py:57`` ignores them with a warning).
"""
import numpy as np

from ..construct.qspace import qubit_key
from ..program import Program, flatten
from .qbuilder import CircuitBuilder


class B200Circuit:
    """Immutable snapshot of the gates built so far (circuit time order)."""

    def __init__(self, gates, qubits, dtype=np.complex128):
        self.gates = list(gates)
        self.qubits = list(qubits)
        self.dtype = np.dtype(dtype)

    def __len__(self):
        return len(self.gates)

    def all_qubits(self):
        return list(self.qubits)

    def program(self, register=None) -> Program:
        reg = list(register) if register is not None else self.qubits
        missing = {q for g in self.gates for q in g.qspace} - set(reg)
        if missing:
            raise ValueError(f"qubit_order lacks {sorted(missing, key=qubit_key)}")
        return flatten(self.gates, reg)

    def unitary(self, qubit_order=None, drop_ancilla=False, fuse=True) -> np.ndarray:
        """Dense operator over ``qubit_order`` (default: all registered qubits, sorted).

        Ancillas are kept as ordinary qubits unless ``drop_ancilla``; with it the result
        is <0|U|0> on the ancillas, i.e. the reference's ``.dela().inflate()``.
        """
        from ..engine import _try_fuse, get_engine, simulate_columns
        import torch
        prog = self.program(qubit_order)
        if drop_ancilla:
            n_free = prog.n_qubits - bin(prog.ancilla_mask).count("1")
            return simulate_columns(prog, 0, n_free, dtype=self.dtype, fuse=fuse, drop_ancilla=True)
        eng = get_engine(self.dtype)
        n = prog.n_qubits
        state = eng.empty(1 << (2 * n))
        eng.init_basis_columns(state, 2 * n, n, 0, 0)
        eng.upload(prog, n)
        spare = None
        self.fused = bool(fuse and _try_fuse(eng))     # False = ran one kernel per gate (a RuntimeWarning said why)
        if self.fused:
            spare = eng.empty(1 << (2 * n))
        res = eng.run(state, spare)
        torch.cuda.synchronize(eng.device)
        return res.view(1 << n, 1 << n).cpu().numpy()

    def apply(self, states, qubit_order=None, fuse=True) -> np.ndarray:
        """Apply the circuit to ``[2^n]`` or ``[2^n, B]`` host state vectors."""
        from ..engine import apply_circuit
        return apply_circuit(self.program(qubit_order), states, dtype=self.dtype, fuse=fuse)

    def simulate_columns(self, col0=0, n_cols_bits=None, fuse=True) -> np.ndarray:
        """Columns ``col0..`` of <0|U|0>: the column-sharded form of ``unitary(drop_ancilla=True)``."""
        from ..engine import simulate_columns
        prog = self.program()
        n_free = prog.n_qubits - bin(prog.ancilla_mask).count("1")
        b = n_free if n_cols_bits is None else n_cols_bits
        return simulate_columns(prog, col0, b, dtype=self.dtype, fuse=fuse, drop_ancilla=True)


class B200Builder(CircuitBuilder):
    def __init__(self, dtype=np.complex128):
        self._qubits = {}
        self._gates = []
        self._dtype = dtype

    def register(self, qspace) -> None:
        for q in qspace:
            self._qubits.setdefault(q, q)

    def build_gate(self, m) -> None:
        if hasattr(m, 'qspace') and hasattr(m, 'controls'):
            unknown = [q for q in m.qspace if q not in self._qubits]
            if unknown:
                raise KeyError(f"gate uses unregistered qubits {unknown}")
            self._gates.append(m)
            return
        if hasattr(m, 'core') and hasattr(m, 'dimension'):
            # a bare UnitaryM: place it on the registered qubits (all of them, sorted)
            from ..construct.cgate import CtrlGate
            reg = self.all_qubits()
            if m.dimension != 1 << len(reg):
                raise ValueError(f"UnitaryM of order {m.dimension} does not fit {len(reg)} registered qubits")
            folded = type(m)(m.dimension, m.core, np.asarray(m.matrix) * m.phase)  # phase lives in the core block
            g = CtrlGate.convert(folded, reg)
            self._gates.append(g)
            return
        raise TypeError(f"cannot build a gate from {type(m)}")

    def finish(self, optimized=False) -> B200Circuit:
        return B200Circuit(self._gates, self.all_qubits(), self._dtype)

    def all_qubits(self) -> list:
        return sorted(self._qubits, key=qubit_key)
