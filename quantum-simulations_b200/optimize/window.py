"""Window operators of the optimiser (mirror of ``quompiler/optimize/window.py:13-104``).

Both operators walk the leaves in matrix-product order and keep, per qubit, the stack of leaves seen so far; a leaf
combines with its predecessors only while they act on the same qubits with the same roles (``qontrol`` equal).
The products taken here are between gates of identical ``qontrol``, i.e. products of their small core matrices
(``cgate.py:104-111``): they stay on the host.  What the reference checks with dense ``reduce(@)`` products
(``tests/qompiler/optimize/test_optimizer.py:137-173``) -- that the optimised tree is the same operator -- is what
runs on the GPU here (``verify.py``).
"""
from collections import defaultdict

import numpy as np

from ..construct.cgate import CtrlGate
from ..construct.types import GateGrain
from ..utils.granularity import granularity


def same_space_product(a, b) -> CtrlGate:
    """``a @ b`` for two gates with equal ``qontrol`` (cgate.py:107-111): direct core product, after sorting both
    when their qubit lists are ordered differently."""
    if a.qspace == b.qspace and [int(c) for c in a.controls()] == [int(c) for c in b.controls()]:
        return CtrlGate(a.matrix() @ b.matrix(), a.controls(), a.qspace, a.phase() * b.phase())
    assert a.qontrol() == b.qontrol()
    sa, sb = a.sorted(), b.sorted()
    return CtrlGate(sa.matrix() @ sb.matrix(), sa.controls(), sa.qspace, sa.phase() * sb.phase())


def is_identity_operator(g) -> bool:
    """``np.allclose(np.array(g), np.eye(g.order()))`` without inflating: every entry outside the controlled block is
    exactly 0 or 1, so the dense comparison reduces to the block ``phase * matrix``."""
    m = np.asarray(g.matrix()) * g.phase()
    return bool(np.allclose(m, np.eye(m.shape[0])))


class WindowOperator:
    def __init__(self, window: int = 0, emit=GateGrain.CLIFFORD_T):
        assert window >= 0
        self.window = window
        self.emit = emit
        self.prefix = defaultdict(list)

    def run(self, node):
        raise NotImplementedError

    def _stacks(self, gate):
        stacks = [self.prefix[q] for q in gate.qspace]
        max_length = min(len(s) for s in stacks)
        return stacks, (max_length if self.window == 0 else min(max_length, self.window))


class AnnihilateOperator(WindowOperator):
    """Cancels a leaf against the run of predecessors whose product with it is the identity (window.py:25-63)."""

    def run(self, node):
        g = node.data
        assert hasattr(g, "controls")
        if np.allclose(g.matrix(), np.eye(1 << len(g.target_qids()))):
            return
        stacks, window_size = self._stacks(g)
        k = self.compatible_length(stacks, window_size, g.qontrol())
        for stack in stacks:
            stack.append(node)
        product = g
        for kk in range(2, k + 2):
            product = same_space_product(stacks[0][-kk].data, product)
            if is_identity_operator(product):
                for i in range(1, 1 + kk):
                    stacks[0][-i].skip = True
                for stack in stacks:
                    while stack and stack[-1].skip:
                        stack.pop()
                return

    @staticmethod
    def compatible_length(stacks, window_size, qontrol):
        for k in range(1, window_size + 1):
            predecessor = {id(s[-k]) if len(s) >= k else None for s in stacks}
            if len(predecessor) > 1 or (None in predecessor):
                return k - 1
            if stacks[0][-k].data.qontrol() != qontrol:
                return k - 1
        return window_size


class ConsolidateOperator(WindowOperator):
    """Folds a leaf into its predecessors while the product stays at least as fine as ``emit`` (window.py:66-104)."""

    def run(self, node):
        g = node.data
        assert hasattr(g, "controls")
        stacks, window_size = self._stacks(g)
        k, product = self.compatible(stacks, window_size, node)
        for i in range(1, k):
            stacks[0][-i].skip = True
        for stack in stacks:
            while stack and stack[-1].skip:
                stack.pop()
        if node.data is not product:
            node.metadata['optimized'] = 'ConsolidateOperator'
        node.data = product
        for stack in stacks:
            stack.append(node)

    def compatible(self, stacks, window_size, node):
        qontrol = node.data.qontrol()
        product = node.data
        for k in range(1, window_size + 1):
            predecessor = {id(s[-k]) if len(s) >= k else None for s in stacks}
            if len(predecessor) > 1 or (None in predecessor):
                return k, product
            prior = stacks[0][-k]
            if prior.data.qontrol() != qontrol:
                return k, product
            tentative = same_space_product(prior.data, product)
            if granularity(tentative) < self.emit:
                return k, product
            product = tentative
        return window_size + 1, product
