#!/usr/bin/env python
"""Differential fuzz of the fused tile-sweep path against the one-kernel-per-gate path (both on the GPU; the latter is
checked against the CPU oracle by the test-suite).  Random circuits mixing every op kind the fusion pass knows:
plain / controlled dense gates, Hadamards, X/CNOT/Toffoli, SWAP triples, diagonal gates with 0, 1 and >= 2 controls,
on random register sizes and batch widths; long circuits exercise the per-launch op / table budgets.

  python tools/fuzz_fused.py [--cases 120] [--seed 1]        (QSV_TILE_BITS=9..12 selects the tile size)
"""
import argparse
import os
import random
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def random_program(rng, n, n_gates):
    from quantum_simulations_b200.construct.qspace import Qubit
    from quantum_simulations_b200.program import Program
    prog = Program([Qubit(i) for i in range(n)])
    h = np.array([[1, 1], [1, -1]]) / np.sqrt(2)
    x = np.array([[0, 1], [1, 0]], dtype=complex)

    def rand_u():
        a = rng.standard_normal((2, 2)) + 1j * rng.standard_normal((2, 2))
        q, r = np.linalg.qr(a)
        return q * (np.diag(r) / np.abs(np.diag(r)))

    for _ in range(n_gates):
        kind = rng.random()
        t = int(rng.integers(n))
        others = [b for b in range(n) if b != t]
        nc = int(rng.choice([0, 0, 0, 1, 1, 2, 3])) if others else 0
        nc = min(nc, len(others))
        ctrls = rng.choice(others, size=nc, replace=False).tolist() if nc else []
        mask = sum(1 << c for c in ctrls)
        val = sum((1 << c) for c in ctrls if rng.random() < 0.7)
        if kind < 0.25:
            prog.add_raw([t], 0, 0, h)
        elif kind < 0.40:
            prog.add_raw([t], mask, val, rand_u())
        elif kind < 0.50:
            prog.add_raw([t], mask, val, x)
        elif kind < 0.58 and others:
            a, b = t, int(rng.choice(others))           # SWAP as three CNOTs
            prog.add_raw([b], 1 << a, 1 << a, x)
            prog.add_raw([a], 1 << b, 1 << b, x)
            prog.add_raw([b], 1 << a, 1 << a, x)
        elif kind < 0.62:
            prog.add_raw([t], mask, val, np.array([[np.cos(.3), -np.sin(.3)], [np.sin(.3), np.cos(.3)]]))  # real, maybe controlled
        else:
            d = np.exp(1j * rng.uniform(0, 2 * np.pi, 2))
            if rng.random() < 0.5:
                d[0] = 1.0
            prog.add_raw([t], mask, val, np.diag(d))
    return prog


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=120)
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args()
    from quantum_simulations_b200.engine import apply_circuit, get_engine
    rng = np.random.default_rng(args.seed)
    worst = 0.0
    fused_total = 0
    for case in range(args.cases):
        n = int(rng.integers(4, 15))
        batch = int(rng.choice([1, 1, 2, 4, 8]))
        n_gates = int(rng.choice([5, 20, 60, 200, 800, 2500])) if n <= 11 else int(rng.choice([5, 20, 60, 200]))
        prog = random_program(rng, n, n_gates)
        x = rng.standard_normal((1 << n, batch)) + 1j * rng.standard_normal((1 << n, batch))
        x /= np.linalg.norm(x, axis=0)
        eng = get_engine()
        eng.reset_stats()
        a = apply_circuit(prog, x, fuse=True)
        st = eng.stats()
        fused_total += st["fused_sweeps"]
        b = apply_circuit(prog, x, fuse=False)
        err = float(np.abs(a - b).max())
        worst = max(worst, err)
        assert st["gates_applied"] == len(prog), (case, st, len(prog))
        assert err < 1e-10, f"case {case}: n={n} batch={batch} gates={len(prog)} err={err}"
    print(f"fuzz ok: {args.cases} cases, worst |fused - unfused| = {worst:.3e}, fused sweeps run = {fused_total}, "
          f"tile bits = {os.environ.get('QSV_TILE_BITS', 'default')}")


if __name__ == "__main__":
    main()
