"""Readers for tests/golden/*.npz (written by oracle/gen_golden.py from the real reference)."""
import os

import numpy as np

import quantum_simulations_b200 as qs

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_ROLE = {1: qs.QType.TARGET, 2: qs.QType.CONTROL0, 4: qs.QType.CONTROL1, 8: qs.QType.IDLER}


def load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def unpack_gates(d, prefix=None):
    """Rebuild CtrlGates (this package's mirror type) from the packed arrays."""
    k = (lambda s: f"{prefix}_{s}") if prefix else (lambda s: s)
    qid, anc, role, off, phase, pool = (d[k(s)] for s in ("qid", "anc", "role", "off", "phase", "pool"))
    gates = []
    for i in range(len(phase)):
        n = int((qid[i] >= 0).sum())
        qspace = [qs.Qubit(int(qid[i, j]), bool(anc[i, j])) for j in range(n)]
        controls = [_ROLE[int(role[i, j])] for j in range(n)]
        m = pool[off[i]:off[i + 1]]
        dim = int(round(np.sqrt(m.size)))
        gates.append(qs.CtrlGate(m.reshape(dim, dim).copy(), controls, qspace, phase=complex(phase[i])))
    return gates


def unpack_mats(d, off_key, pool_key):
    off, pool = d[off_key], d[pool_key]
    out = []
    for i in range(len(off) - 1):
        m = pool[off[i]:off[i + 1]]
        dim = int(round(np.sqrt(m.size)))
        out.append(m.reshape(dim, dim))
    return out


STREAMS = ["cfg1", "cfg2s", "cfg2c", "cyc16", "r4singlet", "r4ctrl_pruned", "r4principal", "r4univ_gate", "r4multi_target", "r32"]
