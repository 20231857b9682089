"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Never imported by the product package.

CPU (numpy) restatement of the reference's simulation path, used by ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs
of ``bench.py`` as the checker and the timed CPU baseline.

Parity status: PINNED.  ``oracle/gen_golden.py`` imports the real reference from
``/root/reference`` (shim: ``np.complexfloating = np.complex128``) and stores its
outputs under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks every
function below against those files.

Two independent restatements are provided so they can also check each other:
 * dense: ``inflate`` of one gate exactly as ``quompiler/construct/unitary.py:70-77``
   with the core from ``quompiler/construct/qontroller.py:112-120`` (``ctrl2core``),
   lifted to the full register and multiplied (``cgate.py:104-121`` without the
   object churn);
 * statevector: the mask-based action on a ``[2^N, B]`` array, SURVEY.md section 8(a).
"""
import itertools

import numpy as np

TARGET, CONTROL0, CONTROL1, IDLER = 1, 2, 4, 8


def qkey(q):
    """Register order: ordinary qubits before ancillas, then by id (qspace.py:31-36)."""
    return (bool(q.ancilla), int(q.qid))


def register_of(gates):
    qs = set()
    for g in gates:
        qs.update(g.qspace)
    return sorted(qs, key=qkey)


def ctrl2core(controls):
    """qontroller.py:112-120 -- big-endian product over each role's spanned bit values."""
    bases = [(0,) if int(c) == CONTROL0 else (1,) if int(c) == CONTROL1 else (0, 1) for c in controls]
    out = []
    for bits in itertools.product(*bases):
        v = 0
        for b in bits:
            v = (v << 1) | b
        out.append(v)
    return out


def core2control(bitlength, core):
    """qontroller.py:86-109."""
    roles = []
    for i in range(bitlength):  # i = bit position, LSB first
        vals = {(a >> i) & 1 for a in core}
        roles.append(TARGET if len(vals) == 2 else (CONTROL1 if (core[0] >> i) & 1 else CONTROL0))
    return tuple(roles[::-1])


def inflate_unitarym(dimension, core, matrix, phase=1.0):
    """unitary.py:70-77: identity with ``matrix*phase`` scattered into ``ix_(core, core)``."""
    out = np.eye(dimension, dtype=np.complex128)
    out[np.ix_(list(core), list(core))] = np.asarray(matrix, dtype=np.complex128) * phase
    return out


def inflate_gate(g):
    """CtrlGate.inflate (cgate.py:61-62) over the gate's own qspace order."""
    controls = [int(c) for c in g.controls()]
    return inflate_unitarym(1 << len(controls), ctrl2core(controls), g.matrix(), complex(g.phase()))


def lift(g, register):
    """Dense operator of one gate on the full ``register`` (first qubit = MSB)."""
    n = len(register)
    pos = {q: i for i, q in enumerate(register)}
    small = inflate_gate(g)
    k = len(g.qspace)
    gbits = [n - 1 - pos[q] for q in g.qspace]  # register bit of each gate qubit, gate-MSB first
    dim = 1 << n
    big = np.zeros((dim, dim), dtype=np.complex128)
    gmask = sum(1 << b for b in gbits)
    for col in range(dim):
        sub_c = 0
        for b in gbits:
            sub_c = (sub_c << 1) | ((col >> b) & 1)
        rest = col & ~gmask
        column = small[:, sub_c]
        for sub_r in np.nonzero(column)[0]:
            row = rest
            for j, b in enumerate(gbits):
                if (int(sub_r) >> (k - 1 - j)) & 1:
                    row |= 1 << b
            big[row, col] = column[sub_r]
    return big


def dense_product(gates_product_order, register=None):
    """``reduce(@)`` of the lifted gates: gates[0] is the left-most factor."""
    register = register_of(gates_product_order) if register is None else list(register)
    out = np.eye(1 << len(register), dtype=np.complex128)
    for g in gates_product_order:
        out = out @ lift(g, register)
    return out


def ancilla_mask(register):
    n = len(register)
    return sum(1 << (n - 1 - i) for i, q in enumerate(register) if q.ancilla)


def project_ancilla(mat, register):
    """<0|U|0> on every ancilla (cgate.py:289-305 with state |0>)."""
    am = ancilla_mask(register)
    keep = [i for i in range(mat.shape[0]) if not (i & am)]
    return mat[np.ix_(keep, keep)]


# ---- statevector restatement ---------------------------------------------------------------
def gate_bits(g, register):
    """(target bits MSB-first, ctrl_mask, ctrl_value) of ``g`` on ``register``."""
    n = len(register)
    pos = {q: i for i, q in enumerate(register)}
    tb, mask, val = [], 0, 0
    for q, c in zip(g.qspace, g.controls()):
        bit = n - 1 - pos[q]
        c = int(c)
        if c == CONTROL1:
            mask |= 1 << bit
            val |= 1 << bit
        elif c == CONTROL0:
            mask |= 1 << bit
        else:
            tb.append(bit)
    return tb, mask, val


def apply_bits(state, n, tbits, cmask, cval, m):
    """In place: for x with (x & cmask) == cval, mix the 2^t amplitudes x|targets with ``m``.

    ``state`` is ``[2^n, B]``.  A controlled X with exact 0/1 entries is applied as a
    swap so permutation circuits stay bit-exact, like the reference's exact products of
    0/1 matrices.
    """
    t = len(tbits)
    tmask = sum(1 << b for b in tbits)
    idx = np.arange(1 << n, dtype=np.int64)
    base = idx[((idx & cmask) == cval) & ((idx & tmask) == 0)]
    offs = np.zeros(1 << t, dtype=np.int64)
    for v in range(1 << t):
        for j, b in enumerate(tbits):
            if (v >> (t - 1 - j)) & 1:
                offs[v] |= 1 << b
    rows = base[None, :] + offs[:, None]          # [2^t, groups]
    m = np.asarray(m, dtype=np.complex128)
    if t == 1 and np.array_equal(m, np.array([[0, 1], [1, 0]])):
        tmp = state[rows[0]].copy()
        state[rows[0]] = state[rows[1]]
        state[rows[1]] = tmp
        return
    block = state[rows]                           # [2^t, groups, B]
    state[rows] = np.einsum('rc,cgb->rgb', m.astype(state.dtype), block)


def apply_gate(state, register, g):
    tb, cm, cv = gate_bits(g, register)
    apply_bits(state, len(register), tb, cm, cv, np.asarray(g.matrix(), dtype=np.complex128) * complex(g.phase()))


def simulate(gates_time_order, register, state):
    """Apply gates (first acts first) to a ``[2^N]`` or ``[2^N, B]`` array; returns a new array."""
    st = np.array(state, dtype=np.result_type(state, np.complex64))
    flat = st.ndim == 1
    st = st.reshape(st.shape[0], -1)
    for g in gates_time_order:
        apply_gate(st, register, g)
    return st[:, 0] if flat else st


def unitary_of(gates_product_order, register=None, drop_ancilla=True):
    """Statevector form of ``reduce(@).sorted().dela().inflate()``: identity columns with the
    ancillas in |0>, the LAST factor acting first (bytecode.py:47-66)."""
    register = register_of(gates_product_order) if register is None else list(register)
    n = len(register)
    am = ancilla_mask(register) if drop_ancilla else 0
    keep = [i for i in range(1 << n) if not (i & am)]
    st = np.zeros((1 << n, len(keep)), dtype=np.complex128)
    st[keep, np.arange(len(keep))] = 1.0
    st = simulate(list(gates_product_order)[::-1], register, st)
    return st[keep, :]


def apply_two_level(state, core, m):
    """Rows ``core`` of a ``[D, B]`` array mixed by ``m`` (UnitaryM action, unitary.py:75-76)."""
    core = list(core)
    state[core] = np.asarray(m, dtype=state.dtype) @ state[core]


def unitary_of_rows(unitaries_product_order, dimension):
    st = np.eye(dimension, dtype=np.complex128)
    for u in list(unitaries_product_order)[::-1]:
        apply_two_level(st, u.core, np.asarray(u.matrix) * u.phase)
    return st


def qft_reference(states):
    """fft_matrix(n) @ x == ifft(x, norm='ortho') (mgen.py:128-137: omega = exp(+2 pi i / N))."""
    return np.fft.ifft(states, axis=0, norm="ortho")


# ---- record-level runner (times the CPU baseline on exactly what the GPU executes) ----------
def run_records(recs, pool, state, n_bits):
    """Apply flattened ``qsv_gate_t`` records (time order) to a flat ``[2^n_bits]`` array in place."""
    st = state.reshape(-1, 1)
    for r in recs:
        kind = int(r["kind"])
        t = int(r["n_targets"])
        if kind == 4:
            off = int(r["matrix_offset"])
            raise NotImplementedError("row records are run with apply_two_level")
        tb = [int(b) for b in r["target_bits"][:t]]
        if kind == 1:
            m = np.array([[0, 1], [1, 0]], dtype=np.complex128)
        else:
            off = int(r["matrix_offset"])
            m = pool[off:off + (1 << (2 * t))].reshape(1 << t, 1 << t)
        apply_bits(st, n_bits, tb, int(r["ctrl_mask"]), int(r["ctrl_value"]), m)
    return state
