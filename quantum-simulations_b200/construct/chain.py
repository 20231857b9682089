"""Lazy product of gates, evaluated on the GPU.

The reference's verification idiom is
``reduce(lambda a, b: a @ b, leaves).sorted().dela().inflate()``
(``tests/qompiler/circuits/test_qompiler.py:79-82``): every ``@`` multiplies two
dense operators lifted to the union of their qubits
(``quompiler/construct/cgate.py:104-121``).  Here ``@`` only concatenates;
``inflate()`` sends the whole chain to the CUDA statevector engine once: the
identity's columns are the batch, the chain is applied right-to-left (the last
factor acts first, ``bytecode.py:47-66``), ancillas start in |0> and are projected
on <0| at the end (``cgate.py:289-305``).  There is no CPU evaluation path.
"""
import numpy as np

from .qspace import qubit_key


class GateChain:
    def __init__(self, gates, register=None, drop_ancilla=False, ancilla_state=None):
        self.gates = list(gates)          # matrix-product order: gates[0] is the left-most factor
        self._register = list(register) if register is not None else None
        self._drop = drop_ancilla
        self._anc_state = ancilla_state   # None = |0>
        kinds = {hasattr(g, 'qspace') for g in self.gates}
        assert len(kinds) == 1, 'cannot mix UnitaryM and CtrlGate in one product'
        self._qubit_gates = kinds.pop()
        if not self._qubit_gates:
            dims = {g.dimension for g in self.gates}
            if len(dims) != 1:
                raise ValueError('matmul: Input operands have dimension mismatch.')
            self.dimension = dims.pop()

    # -- structure ---------------------------------------------------------------------
    @property
    def qspace(self):
        if not self._qubit_gates:
            raise AttributeError('a UnitaryM chain has no qspace')
        if self._register is not None:
            reg = self._register
        else:
            seen = set()
            for g in self.gates:
                seen.update(g.qspace)
            reg = sorted(seen, key=qubit_key)
        return [q for q in reg if not (self._drop and q.ancilla)]

    def _full_register(self):
        if self._register is not None:
            return list(self._register)
        seen = set()
        for g in self.gates:
            seen.update(g.qspace)
        return sorted(seen, key=qubit_key)

    def order(self) -> int:
        return (1 << len(self.qspace)) if self._qubit_gates else self.dimension

    def __matmul__(self, other):
        rhs = other.gates if isinstance(other, GateChain) else [other]
        if self._drop or (isinstance(other, GateChain) and other._drop):
            raise NotImplementedError('multiply before projecting the ancillas')
        return GateChain(self.gates + list(rhs))

    def __rmatmul__(self, other):
        return GateChain([other] + self.gates)

    def sorted(self, sorting=None):
        if sorting is not None:
            raise NotImplementedError('a lazy product is always reported over the sorted register')
        return GateChain(self.gates, None, self._drop, self._anc_state)

    def dela(self, state=None):
        """Ancillas fixed in ``state`` (default |0>) on both sides: the sandwich ``<state| U |state>`` of
        ``cgate.py:289-305`` taken on the product's dense operator, evaluated on the device (ancilla bits of the
        basis columns start in ``state``, the read-out contracts them with ``<state|``).  Unlike the reference, a
        non-unitary outcome is returned rather than rejected: the lazy product never builds a ``CtrlGate``."""
        if state is not None:
            assert isinstance(state, np.ndarray), 'state vector must be a np.ndarray.'
            assert state.shape == (2,), f'state vector must be a 1D array of length 2, but got {state.shape}'
        return GateChain(self.gates, self._register, True, state)

    # -- evaluation ----------------------------------------------------------------------
    def inflate(self, dtype=np.complex128) -> np.ndarray:
        from ..engine import reconstruct_unitary, reconstruct_unitary_rows
        if self._qubit_gates:
            return reconstruct_unitary(self.gates[::-1], self._full_register(), drop_ancilla=self._drop, dtype=dtype,
                                       ancilla_state=self._anc_state if self._drop else None)
        return reconstruct_unitary_rows(self.gates[::-1], self.dimension, dtype=dtype)

    def __array__(self, dtype=None, copy=None):
        m = self.inflate()
        return m if dtype is None else m.astype(dtype)

    def __repr__(self):
        return f'GateChain{{{len(self.gates)} gates}}'
