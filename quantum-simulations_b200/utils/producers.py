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


def two_level_stream_scan(u: np.ndarray, device=None):
    """The same stream as :func:`two_level_stream`, one COLUMN of the QR sweep at a time instead of one rotation at a time
    ("GPU Givens-QR is a natural follow-on").  Runs on torch tensors -- ``device="cuda"`` for the GPU,
    ``"cpu"`` for tests; the result is the same list of ``(row_i, row_j, 2x2)``.

    All rotations of column n share the pivot row n, and the pivot element after k of them is the prefix norm
    ``A_k = sqrt(|a_0|^2 + sum_{m<k} |b_m|^2)`` (``a_0 = M[n,n]``, ``b_m = M[m,n]``), so every 2x2 of the column
    (``mat2l_decompose.py:31-33``) is known from the column alone.  The rows follow in closed form too: the pivot row
    before rotation j is ``s_j = (conj(a_0) row_n + sum_{m<j} conj(b_m) row_m) / A_j`` -- one cumulative sum over the rows --
    and row j becomes ``(b_j s_j - a_j row_j) / den_j``.  Skip rules as in the reference (``mat2l_decompose.py:24-30``):
    a rotation with ``b ~ 0`` and ``a ~ 1`` is left out (it neither enters the sums nor the stream)."""
    import torch
    dev = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
    M = torch.as_tensor(np.array(u, dtype=np.complex128), device=dev).clone()
    N = M.shape[0]
    eye = torch.eye(N, dtype=torch.complex128, device=dev)
    rows_out, mats_out = [], []

    def identity_lines(m):
        close = torch.isclose(m, eye)
        return close.all(dim=0) & close.all(dim=1)

    for n in range(N - 1):
        if int((~identity_lines(M)).sum()) <= 2:               # validm2l (mat_utils.py:38-46)
            break
        b = M[n + 1:, n].clone()
        if bool(torch.isclose(b, torch.zeros_like(b)).all()):
            continue
        a0 = M[n, n].clone()
        b2 = (b.conj() * b).real
        zero_b = torch.isclose(b, torch.zeros_like(b))
        one = torch.ones((), dtype=torch.complex128, device=dev)
        # pass 1: pivots as if nothing were skipped (a skipped rotation has |b| <= 1e-8: it moves the prefix norm by ~1e-16)
        pref = torch.cumsum(b2, 0) - b2 + (a0.conj() * a0).real          # |a_0|^2 + sum_{m<j} |b_m|^2
        a_real = torch.sqrt(pref).to(torch.complex128)
        first_guess = torch.arange(b.numel(), device=dev) == 0
        a_j = torch.where(first_guess, a0, a_real)
        skip = zero_b & torch.isclose(a_j, one.expand_as(a_j))
        # pass 2: exact pivots with the skipped rotations left out; the first rotation that is NOT skipped sees a_0 itself
        keep = ~skip
        if not bool(keep.any()):
            continue
        b2k = torch.where(keep, b2, torch.zeros_like(b2))
        pref = torch.cumsum(b2k, 0) - b2k + (a0.conj() * a0).real
        kept_before = torch.cumsum(keep.to(torch.int64), 0) - keep.to(torch.int64)
        is_first = keep & (kept_before == 0)
        a_j = torch.where(is_first, a0, torch.sqrt(pref).to(torch.complex128))
        den = torch.sqrt((a_j.conj() * a_j).real + b2).to(torch.complex128)
        # 2x2 of every kept rotation (mat2l_decompose.py:31-33)
        c = torch.stack([torch.stack([a_j / den, b.conj() / den], -1), torch.stack([b / den, -a_j.conj() / den], -1)], -2)
        idx = torch.nonzero(keep).flatten()
        rows_out.append(torch.stack([torch.full_like(idx, n), idx + n + 1], -1))
        mats_out.append(c[idx])
        # rows: s_j = pivot row before rotation j, for every j at once
        R = M[n + 1:, :]
        bk = torch.where(keep, b, torch.zeros_like(b))
        contrib = bk.conj().unsqueeze(1) * R
        C = torch.cumsum(contrib, 0) - contrib                            # sum_{m<j} conj(b_m) row_m
        row_n = M[n, :].clone()
        A = torch.sqrt(pref).to(torch.complex128)                         # A_j (exact pivots)
        S = (a0.conj() * row_n.unsqueeze(0) + C) / A.unsqueeze(1)
        S = torch.where(is_first.unsqueeze(1) | (kept_before == 0).unsqueeze(1), row_n.unsqueeze(0).expand_as(S), S)
        newR = (b.unsqueeze(1) * S - a_j.unsqueeze(1) * R) / den.unsqueeze(1)
        last = idx[-1]
        new_pivot = (a_j[last].conj() * S[last] + b[last].conj() * R[last]) / den[last]     # (R is a view: read it before the update)
        M[n + 1:, :] = torch.where(keep.unsqueeze(1), newR, R)
        M[n, :] = new_pivot
    idl = identity_lines(M).cpu().numpy()
    Mh = M.cpu().numpy()
    out = []
    if rows_out:
        rr = torch.cat(rows_out).cpu().numpy()
        mm = torch.cat(mats_out).cpu().numpy()
        out = [(int(r[0]), int(r[1]), m) for r, m in zip(rr, mm)]
    core = tuple(np.nonzero(~idl)[0].tolist())
    if not core:
        core = (0, 1)
    if len(core) == 1:
        core = tuple(sorted(set(core) | {0 if core[0] != 0 else 1}))
    m2l = Mh[np.ix_(core, core)]
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


def singlet_records_fast(stream, n_qubits: int, batch_bits: int = 0):
    """:func:`singlet_records` without the Python loop over the stream: ops are grouped by the number of differing index
    bits p (the Gray chain of ``cnot_decompose.py:98-107`` flips them lowest first: p-1 fully controlled X, the fully
    controlled U, the same X again in reverse), and every group is written with array operations.  Same records, same
    order, same pool."""
    K = len(stream)
    if K == 0:
        return np.zeros(0, dtype=GATE_DTYPE), np.zeros(0, dtype=np.complex128)
    full = (1 << n_qubits) - 1
    ii = np.array([s[0] for s in stream][::-1], dtype=np.int64)            # time order: last factor acts first
    jj = np.array([s[1] for s in stream][::-1], dtype=np.int64)
    cc = np.array([np.asarray(s[2], dtype=np.complex128) for s in stream][::-1]).reshape(K, 2, 2)
    d = ii ^ jj
    p = np.zeros(K, dtype=np.int64)
    for b in range(n_qubits):
        p += (d >> b) & 1
    assert (p >= 1).all()
    count = 2 * (p - 1) + 1
    start = np.concatenate([[0], np.cumsum(count)])
    total = int(start[-1])
    recs = np.zeros(total, dtype=GATE_DTYPE)
    recs["n_targets"] = 1
    pool = np.zeros((K, 4), dtype=np.complex128)
    xcx = cc[:, ::-1, ::-1]                                                   # X c X
    for pp in np.unique(p):
        sel = np.nonzero(p == pp)[0]
        i0, dd = ii[sel], d[sel]
        # the differing bits of every op, lowest first: bits[t] (shape [len(sel)])
        bits = []
        rem = dd.copy()
        for _ in range(int(pp)):
            low = rem & -rem
            bits.append(low)
            rem = rem ^ low
        state = i0.copy()                                                     # code[t]
        states = [state.copy()]
        for t in range(int(pp)):
            state = state ^ bits[t]
            states.append(state.copy())
        base = start[sel]
        for t in range(int(pp)):                                              # step t: code[t] -> code[t+1], flips bits[t]
            diff = bits[t]
            tb = np.log2(diff.astype(np.float64)).astype(np.int64)
            mask = (full & ~diff) << batch_bits
            val = (states[t] & ~diff) << batch_bits
            if t < pp - 1:
                for pos in (base + t, base + count[sel] - 1 - t):             # X on the way in and on the way out
                    recs["kind"][pos] = PERMX
                    recs["target_bits"][pos, 0] = tb + batch_bits
                    recs["ctrl_mask"][pos] = mask
                    recs["ctrl_value"][pos] = val
            else:
                pos = base + t
                a, b = states[t], states[t + 1]
                v = np.where((a < b)[:, None, None], cc[sel], xcx[sel])
                diag = (v[:, 0, 1] == 0) & (v[:, 1, 0] == 0)
                recs["kind"][pos] = np.where(diag, DIAG1, DENSE1)
                recs["target_bits"][pos, 0] = tb + batch_bits
                recs["ctrl_mask"][pos] = mask
                recs["ctrl_value"][pos] = val
                recs["matrix_offset"][pos] = 4 * sel
                pool[sel] = v.reshape(-1, 4)
    return recs, pool.reshape(-1)
