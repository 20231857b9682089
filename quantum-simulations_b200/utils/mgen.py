"""Seeded workload generators for parity tests and benchmarks.

Restates the generators of ``quompiler/utils/mgen.py`` that feed the simulation
path -- ``random_unitary`` (27-37), ``cyclic_matrix`` (112-125), ``fft_matrix``
(128-137), ``random_state`` (148-155), ``random_control`` (60-76),
``random_CtrlGate`` (52-55), ``random_ctrlgate`` (140-145) -- drawing the same
numbers from the same global ``random`` / ``numpy.random`` streams, plus the QFT
gate circuit of ``tests/qompiler/utils/test_mgen.py:84-99`` generalised to n qubits.
"""
import random

import numpy as np

from ..construct.cgate import CtrlGate
from ..construct.qspace import Qubit
from ..construct.types import QType, UnivGate


def random_unitary(n) -> np.ndarray:
    a = np.random.randn(n, n) + 1j * np.random.randn(n, n)
    q, r = np.linalg.qr(a)
    d = np.diag(r) / np.abs(np.diag(r))
    return q @ np.diag(d)


def random_state(dimension) -> np.ndarray:
    arr = np.random.randn(dimension) + 1j * np.random.randn(dimension)
    return arr / np.linalg.norm(arr)


def random_phase():
    return np.exp(1j * np.random.uniform(0, 2 * np.pi))


def cyclic_matrix(n, i=0, j=None, c=1) -> np.ndarray:
    j = n if j is None else j
    idx = list(range(n))
    perm = idx[:i] + np.roll(idx[i:j], c).tolist() + idx[j:]
    return np.diag([1] * n)[perm]


def fft_matrix(n) -> np.ndarray:
    N = 1 << n
    omega = np.exp(2j * np.pi / N)
    j, k = np.meshgrid(np.arange(N), np.arange(N))
    return (omega ** ((j * k) % N)) / np.sqrt(N)


def random_control(k, t=None):
    if t is None:
        return random.choices([QType.CONTROL1, QType.CONTROL0, QType.TARGET], k=k)
    assert 0 <= t <= k
    result = random.choices([QType.CONTROL1, QType.CONTROL0], k=k)
    for i in random.sample(range(k), t):
        result[i] = QType.TARGET
    return tuple(result)


def random_CtrlGate(controls, qubits=None) -> CtrlGate:
    from scipy.stats import unitary_group
    n = list(controls).count(QType.TARGET)
    return CtrlGate(unitary_group.rvs(1 << n), controls, qubits)


def random_ctrlgate(ctrnum, targetnum, qnum=None) -> CtrlGate:
    controls = random_control(ctrnum, targetnum)
    if qnum is None:
        return random_CtrlGate(controls)
    qubits = [Qubit(q) for q in random.sample(range(qnum), ctrnum)]
    return random_CtrlGate(controls, qubits)


def qft_circuit(n: int):
    """QFT on qubits q0..q(n-1) as CtrlGates in circuit TIME order.

    For i = 0..n-1: H(q_i), then for j > i a phase gate diag(1, exp(2 pi i / 2^(j-i+1))) on
    target q_i controlled by q_j; finally floor(n/2) swaps q_i <-> q_(n-1-i), each as three
    CNOTs.  Equals ``fft_matrix(n)`` (test_mgen.py:84-99 is the n = 3 instance).
    """
    q = [Qubit(i) for i in range(n)]
    T, C1 = QType.TARGET, QType.CONTROL1
    gates = []
    for i in range(n):
        gates.append(CtrlGate(UnivGate.H, [T], [q[i]]))
        for j in range(i + 1, n):
            ph = np.exp(2j * np.pi / (1 << (j - i + 1)))
            gates.append(CtrlGate(np.array([[1, 0], [0, ph]]), [T, C1], [q[i], q[j]]))
    for i in range(n // 2):
        a, b = q[i], q[n - 1 - i]
        gates.append(CtrlGate(UnivGate.X, [C1, T], [a, b]))
        gates.append(CtrlGate(UnivGate.X, [T, C1], [a, b]))
        gates.append(CtrlGate(UnivGate.X, [C1, T], [a, b]))
    return gates


def qft_program(n: int):
    """The same QFT-n circuit as :func:`qft_circuit`, flattened directly with EXACT phases.

    Going through ``CtrlGate`` snaps any matrix within ``np.allclose`` tolerance of a named gate to
    that gate (reference behaviour, cgate.py:33-37), which turns the controlled phases with
    angle < ~1e-5 (j - i >= 19) into identities; benchmarks and large-n FFT checks want the exact
    transform, so they use this generator.
    """
    from ..program import Program
    prog = Program([Qubit(i) for i in range(n)])
    h = np.array(UnivGate.H)
    x = np.array(UnivGate.X)
    bit = lambda i: n - 1 - i
    for i in range(n):
        prog.add_raw([bit(i)], 0, 0, h)
        for j in range(i + 1, n):
            ph = np.exp(2j * np.pi / (1 << (j - i + 1)))
            prog.add_raw([bit(i)], 1 << bit(j), 1 << bit(j), np.array([[1, 0], [0, ph]]))
    for i in range(n // 2):
        a, b = bit(i), bit(n - 1 - i)
        prog.add_raw([b], 1 << a, 1 << a, x)
        prog.add_raw([a], 1 << b, 1 << b, x)
        prog.add_raw([b], 1 << a, 1 << a, x)
    return prog
