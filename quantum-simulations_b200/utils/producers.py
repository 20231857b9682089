"""Fast gate-stream producers.

The reference builds the stream for a dense N x N unitary in two steps:
 * ``mat2l_decompose`` (``quompiler/utils/mat2l_decompose.py:12-43``): Givens QR into two-level
   unitaries -- but it re-multiplies the full matrix by an inflated N x N factor for every rotation
   (line 36), O(N^5) overall, which makes N = 1024 (BASELINE config 4) unreachable;
 * ``cnot_decompose`` (``quompiler/utils/cnot_decompose.py:84-107``): each two-level matrix becomes a
   Gray-code chain (``gray.py:4-17``) of fully controlled X gates around one fully controlled U.

Here the same rotations (formulas of lines 31-33, skip rules of lines 24-30, the trailing 2x2 of
lines 37-42) are produced by updating only the two affected rows, O(N^3), and the SINGLET-level chain
is written directly as ``qsv_gate_t`` records.  Validated against the reference's own output in
``tests/golden/two_level.npz`` (dimension 8/3/6) and ``streams.npz`` (``r32``: 2034 SINGLET gates).
"""
import numpy as np

from .._lib import DENSE1, DIAG1, GATE_DTYPE, PERMX, ROWS2

_X = np.array([[0, 1], [1, 0]], dtype=np.complex128)


def _identity_lines(m, tol_kw):
    """Indexes whose row AND column equal the identity's (mat_utils.py:7-18), vectorised."""
    eye = np.eye(m.shape[0])
    close = np.isclose(m, eye, **tol_kw)
    return close.all(axis=0) & close.all(axis=1)


def two_level_stream(u: np.ndarray):
    """Givens QR of a unitary: list of ``(row_i, row_j, 2x2)`` in the reference's product order
    (``u == T_0 @ T_1 @ ... @ T_last``, each T two-level)."""
    matrix = np.array(u, dtype=np.complex128)
    N = matrix.shape[0]
    out = []
    for n in range(N - 1):
        if (~_identity_lines(matrix, {})).sum() <= 2:          # validm2l (mat_utils.py:38-46)
            break
        col = matrix[n + 1:, n]
        if np.allclose(col, 0):
            continue
        for i in range(n + 1, N):
            a, b = matrix[n, n], matrix[i, n]
            if np.isclose(b, 0) and np.isclose(a, 1):
                continue
            den = np.sqrt(np.conj(a) * a + np.conj(b) * b)
            c = np.array([[a / den, np.conj(b) / den], [b / den, -np.conj(a) / den]])
            out.append((n, i, c))
            rows = matrix[[n, i], :]
            matrix[[n, i], :] = c.conj().T @ rows             # only rows n and i change
    idl = _identity_lines(matrix, {})
    core = tuple(np.nonzero(~idl)[0].tolist())
    if not core:
        core = (0, 1)
    if len(core) == 1:                                        # keep it a 2x2 like coreindexes + [:2] padding would
        core = tuple(sorted(set(core) | {0 if core[0] != 0 else 1}))
    m2l = matrix[np.ix_(core, core)]
    if len(core) == 2 and not np.allclose(m2l, np.eye(2)):
        out.append((core[0], core[1], m2l))
    elif len(core) > 2:
        raise ValueError("input is not unitary enough: the QR residue is not two-level")
    return out


def two_level_records(stream):
    """ROWS2 records (time order = reversed product order) + matrix pool for a two-level stream."""
    n = len(stream)
    recs = np.zeros(n, dtype=GATE_DTYPE)
    pool = np.zeros(4 * n, dtype=np.complex128)
    for k, (i, j, c) in enumerate(reversed(stream)):
        recs[k]["kind"] = ROWS2
        recs[k]["row0"], recs[k]["row1"] = i, j
        recs[k]["matrix_offset"] = 4 * k
        pool[4 * k:4 * k + 4] = np.asarray(c).reshape(-1)
    return recs, pool


def gray_code(n1: int, n2: int):
    """gray.py:4-17."""
    res = [n1]
    for i in range(max(n1.bit_length(), n2.bit_length())):
        mask = 1 << i
        bit = n2 & mask
        if res[-1] & mask != bit:
            res.append((res[-1] ^ mask) | bit)
    return res


def singlet_records(stream, n_qubits: int, batch_bits: int = 0):
    """SINGLET-level records (fully controlled X / U, cnot_decompose.py:84-107) of a two-level stream,
    in circuit time order, bit positions already shifted by ``batch_bits``."""
    full = (1 << n_qubits) - 1
    recs = []
    mats = []
    for (i, j, c) in reversed(stream):                        # time order: last factor acts first
        code = gray_code(i, j)
        comps = list(zip(code, code[1:-1]))
        a, b = code[-2], code[-1]
        v = np.asarray(c) if a < b else _X @ np.asarray(c) @ _X
        chain = [("x", p, q) for p, q in comps] + [("u", a, b)] + [("x", p, q) for p, q in reversed(comps)]
        for kind, p, q in chain:
            diff = p ^ q
            tb = diff.bit_length() - 1
            rec = np.zeros((), dtype=GATE_DTYPE)
            rec["n_targets"] = 1
            rec["target_bits"][0] = tb + batch_bits
            rec["ctrl_mask"] = (full & ~diff) << batch_bits
            rec["ctrl_value"] = (p & ~diff) << batch_bits
            if kind == "x":
                rec["kind"] = PERMX
            else:
                rec["kind"] = DIAG1 if (v[0, 1] == 0 and v[1, 0] == 0) else DENSE1
                rec["matrix_offset"] = 4 * len(mats)
                mats.append(v.reshape(-1))
            recs.append(rec)
    recs = np.array(recs, dtype=GATE_DTYPE) if recs else np.zeros(0, dtype=GATE_DTYPE)
    pool = np.concatenate(mats) if mats else np.zeros(0, dtype=np.complex128)
    return recs, pool
