"""Peephole pass over flattened gate records: the compile-time twin of the fusion pass.

What the reference's ``SlidingWindowCombiner`` decides on a tree of gate objects with per-qubit stacks of tree nodes
(``quompiler/optimize/basic_optimizer.py:11-49``, ``window.py:25-104``) is decided here on the SAME POD records the
statevector engine executes (``qsv_gate_t``: control mask / value, target bits, a ``2^t x 2^t`` matrix with the phase
folded in), with one table: for every index bit, the last surviving record that touches it, and from every record a
link per bit to the one before.  Two records can meet only if all bits of the later one name the same predecessor and
both have the same key ``(ctrl_mask, ctrl_value, target bits)``:

* pass 1 (cancel): a run of such neighbours whose product is the identity disappears
  (X X, CX CX, T TD, ...; ``AnnihilateOperator``);
* pass 2 (merge): a run is replaced by its product as long as the product is no coarser than the emit level
  (T T -> S, S S -> Z, H X H -> Z at CTRL_PRUNED; ``ConsolidateOperator`` and ``utils/granularity.py:30-44``).

``window`` bounds the run length exactly as in the reference (0 = unbounded).  Records are taken in MATRIX-PRODUCT order
(the order of ``BytecodeIter``): record i stands left of record i+1.  The output is pinned to the reference's own on
thirteen compiled / hand-made streams (``tests/golden/optimizer.npz``) and the operator is compared before and after on
the GPU (``optimize.verify``).
"""
import numpy as np

from .._lib import DENSE1, DENSEK, DIAG1, PERMX
from ..construct.types import GateGrain, UnivGate

_X = np.array([[0, 1], [1, 0]], dtype=np.complex128)


def _matrix_of(rec, pool):
    if int(rec["kind"]) == PERMX:
        return _X
    t = int(rec["n_targets"])
    off = int(rec["matrix_offset"])
    return np.asarray(pool[off:off + (1 << (2 * t))], dtype=np.complex128).reshape(1 << t, 1 << t)


def _sorted_targets(tbits, m):
    """Key form of a multi-target record: target bits in descending order (most significant first), matrix permuted to
    match -- two records over the same bits in a different order get the same key."""
    t = len(tbits)
    order = sorted(range(t), key=lambda j: -tbits[j])
    if order == list(range(t)):
        return tuple(tbits), m
    idx = np.zeros(1 << t, dtype=np.int64)
    for v in range(1 << t):            # v indexes the sorted order; find the index in the record's own order
        o = 0
        for jn, jo in enumerate(order):
            if (v >> (t - 1 - jn)) & 1:
                o |= 1 << (t - 1 - jo)
        idx[v] = o
    return tuple(tbits[j] for j in order), m[np.ix_(idx, idx)]


def grain_of(m, n_controls):
    """Granularity of a product as the reference's ``granularity_gate`` sees it (a product is never an RGate)."""
    t = m.shape[0].bit_length() - 1
    named = UnivGate.get_prop(m) if t == 1 else None
    if named is not None:
        if n_controls > 1:
            return GateGrain.SINGLET
        return GateGrain.CLIFFORD_T if named in UnivGate.cliffordt() else GateGrain.CTRL_PRUNED
    if t > 1:
        return GateGrain.MULTI_TARGET
    return GateGrain.CTRL_PRUNED if n_controls < 2 else GateGrain.SINGLET


class _Table:
    """Per index bit: the last surviving record; per record and bit: the surviving record before it."""

    def __init__(self):
        self.last = {}
        self.prev = {}

    def chain(self, bits, key, keys, limit):
        """Records that precede a new record over ``bits`` on ALL of them, newest first, while they carry ``key``."""
        out = []
        cursor = {b: self.last.get(b, -1) for b in bits}
        while limit == 0 or len(out) < limit:
            heads = set(cursor.values())
            if len(heads) != 1:
                break
            p = heads.pop()
            if p < 0 or keys[p] != key:
                break
            out.append(p)
            cursor = {b: self.prev[p][b] for b in bits}
        return out

    def push(self, i, bits):
        self.prev[i] = {b: self.last.get(b, -1) for b in bits}
        for b in bits:
            self.last[b] = i

    def drop_tail(self, bits, n):
        """Unlink the n newest records over ``bits`` (they are the tail on every one of these bits)."""
        for _ in range(n):
            p = self.last[next(iter(bits))]
            for b in bits:
                self.last[b] = self.prev[p][b]


def peephole_records(recs, pool, window=0, emit=GateGrain.CLIFFORD_T, core_is_identity=None):
    """Returns ``(keep, matrices, absorbed)``: indexes of the surviving records (ascending); for each its matrix in the
    key form (target bits descending); and the indexes of the records it swallowed in the second pass.  ``core_is_identity[i]`` (optional) marks records whose matrix is the identity
    before its phase is folded in -- the reference leaves those alone in the first pass; without it the folded matrix
    decides."""
    n = len(recs)
    keys, mats, bits = [], [], []
    for i in range(n):
        r = recs[i]
        if int(r["kind"]) not in (DENSE1, PERMX, DIAG1, DENSEK):
            raise ValueError("the peephole pass works on qubit gates (row-pair records have no neighbours to merge with)")
        t = int(r["n_targets"])
        tb, m = _sorted_targets([int(x) for x in r["target_bits"][:t]], _matrix_of(r, pool))
        cm, cv = int(r["ctrl_mask"]), int(r["ctrl_value"])
        keys.append((cm, cv, tb))
        mats.append(m)
        bits.append(frozenset(tb) | frozenset(b for b in range(cm.bit_length()) if (cm >> b) & 1))
    alive = [True] * n
    swallowed = {}
    # ---- pass 1: cancel runs whose product is the identity --------------------------------------------------
    tab = _Table()
    for i in range(n):
        eye = np.eye(mats[i].shape[0])
        ident = bool(core_is_identity[i]) if core_is_identity is not None else np.allclose(mats[i], eye)
        if ident:
            continue                                   # stays where it is, invisible to its neighbours in this pass
        run = tab.chain(bits[i], keys[i], keys, window)
        tab.push(i, bits[i])
        prod = mats[i]
        for j, p in enumerate(run):
            prod = mats[p] @ prod
            if np.allclose(prod, eye):
                tab.drop_tail(bits[i], j + 2)
                alive[i] = False
                for q in run[:j + 1]:
                    alive[q] = False
                break
    # ---- pass 2: merge runs while the product stays at least as fine as `emit` -----------------------------
    tab = _Table()
    for i in range(n):
        if not alive[i]:
            continue
        n_ctrl = bin(keys[i][0]).count("1")
        run = tab.chain(bits[i], keys[i], keys, window)
        prod = mats[i]
        absorbed = []
        for p in run:
            tentative = mats[p] @ prod
            if grain_of(tentative, n_ctrl) < emit:
                break
            prod = tentative
            alive[p] = False
            absorbed.append(p)
            absorbed += swallowed.pop(p, [])
        if absorbed:
            tab.drop_tail(bits[i], len([p for p in run if not alive[p]]))
            swallowed[i] = absorbed
        mats[i] = prod
        tab.push(i, bits[i])
    keep = [i for i in range(n) if alive[i]]
    return keep, {i: mats[i] for i in keep}, {i: swallowed.get(i, []) for i in keep}


def optimize_leaves(leaves, window=0, emit=GateGrain.CLIFFORD_T):
    """The pass on gate objects (this package's ``CtrlGate`` or the reference's, by duck typing) in matrix-product order:
    flatten to records, run :func:`peephole_records`, and hand back gates on the survivors' own qubits and roles."""
    from ..construct.cgate import CtrlGate
    from ..construct.qspace import qubit_key
    from ..program import flatten
    leaves = list(leaves)
    if not leaves:
        return []
    prog = flatten(leaves)
    recs, pool = prog.records(0), prog.pool()
    flags = [bool(np.allclose(np.asarray(g.matrix()), np.eye(np.asarray(g.matrix()).shape[0]))) for g in leaves]
    keep, mats, absorbed = peephole_records(recs, pool, window, emit, core_is_identity=flags)
    out = []
    n_bits = prog.n_qubits
    bit_of = {q: n_bits - 1 - i for i, q in enumerate(prog.register)}
    for i in keep:
        g = leaves[i]
        if not absorbed[i]:
            out.append(g)
            continue
        # the product's qubit order: the common one if all factors agree, else the sorted one (cgate.py:108-112)
        same_order = all(list(leaves[j].qspace) == list(g.qspace) for j in absorbed[i])
        roles = dict(zip(g.qspace, g.controls()))
        qspace = list(g.qspace) if same_order else sorted(g.qspace, key=qubit_key)
        targets = [q for q in qspace if q in set(g.target_qids())]
        m = mats[i]
        want = [bit_of[q] for q in targets]
        if want != sorted(want, reverse=True):   # key form is bit-descending: bring the matrix to the gate's own target order
            t = len(want)
            desc = sorted(want, reverse=True)
            idx = np.zeros(1 << t, dtype=np.int64)
            for v in range(1 << t):
                o = 0
                for jn, b in enumerate(want):
                    if (v >> (t - 1 - jn)) & 1:
                        o |= 1 << (t - 1 - desc.index(b))
                idx[v] = o
            m = m[np.ix_(idx, idx)]
        out.append(CtrlGate(m, [roles[q] for q in qspace], qspace))
    return out


class RecordPeephole:
    """``optimize(root)`` for a ``Bytecode`` tree, like the reference's optimiser objects: the tree's gate leaves are
    replaced by the survivors of :func:`optimize_leaves` under the same root (leaves that are not gates -- two-level
    matrices -- fence the pass: nothing merges across them)."""

    def __init__(self, window: int = 0, emit=GateGrain.CLIFFORD_T):
        assert window >= 0
        self.window = window
        self.emit = emit

    def optimize(self, root):
        from ..construct.bytecode import Bytecode, BytecodeIter
        leaves = [a.data for a in BytecodeIter(root) if a.is_leaf() and not a.skip]
        out, run = [], []
        for d in leaves:
            if hasattr(d, "controls"):
                run.append(d)
            else:
                out += optimize_leaves(run, self.window, self.emit) + [d]
                run = []
        out += optimize_leaves(run, self.window, self.emit)
        root.children = [Bytecode(d) for d in out]
        root.skip = False
        return root
