#!/usr/bin/env python
"""Small fused runs for compute-sanitizer (racecheck / memcheck): every kernel image on a few tiles.

  compute-sanitizer --tool racecheck python tools/sanitize_small.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("QSV_TILE_BITS", "9")


def main():
    import quantum_simulations_b200 as qs  # noqa: F401
    from oracle import sv_oracle as orc
    from quantum_simulations_b200.engine import apply_circuit
    from quantum_simulations_b200.program import Program
    from quantum_simulations_b200.utils.mgen import qft_program
    rng = np.random.default_rng(3)
    n = 12
    progs = {"qft (radix-8 TMA image incl. the permuting store)": qft_program(n)}
    p = Program([qs.Qubit(i) for i in range(n)])
    for _ in range(40):
        bits = rng.choice(n, size=3, replace=False)
        a = rng.standard_normal((2, 2)) + 1j * rng.standard_normal((2, 2))
        q, _ = np.linalg.qr(a)
        p.add_raw([int(bits[0])], (1 << int(bits[1])) | (1 << int(bits[2])), 1 << int(bits[1]), q)
    progs["controlled dense gates (generic stage loop / interpreter images)"] = p
    for name, prog in progs.items():
        x = rng.standard_normal((1 << n, 1)) + 1j * rng.standard_normal((1 << n, 1))
        x /= np.linalg.norm(x)
        got = apply_circuit(prog, x, fuse=True)[:, 0]
        ref = x[:, 0].copy()
        orc.run_records(prog.records(0), prog.pool(), ref, n)
        print(name, "max abs err", float(np.abs(got - ref).max()))
        assert np.abs(got - ref).max() < 1e-10


if __name__ == "__main__":
    main()
