"""``QasmBuilder``: OpenQASM 3 text behind the ``CircuitBuilder`` plug-in API.

Plays the role of the reference's ``QiskitBuilder`` (``quompiler/circuits/qiskit_factory/qiskit_builder.py:18-93``)
for the one thing the demos do with its result -- ``qasm.dumps(circuit)`` in
``demo/compile_unitary_demo.py:59-61`` -- without qiskit (absent from this image; the Qiskit leg of the reference's
tests asserts shapes only, so PARITY IS UNPINNED for it).  What pins this exporter instead: the text
parses back (``parse_qasm``) into the gate program it was built from, and that program, run by the CUDA engine or
the oracle, reproduces the circuit's unitary (tests/test_qasm.py).

Mapping (same cases as ``QiskitBuilder.build_gate``):
 * ordinary qubits -> ``qubit[n] q``, ancillas -> ``qubit[m] a``, both in sorted register order (two registers
   like ``QuantumRegister`` + ``AncillaRegister``, qiskit_builder.py:28-42);
 * standard gates (``is_std``) -> ``id x y z h s t sdg tdg`` of ``stdgates.inc`` (``map_gate``, 71-93);
 * principal-axis rotations (``is_principal``) -> ``rx / ry / rz(angle)`` (19, 50-53);
 * any other one-target unitary -> ``U(theta, phi, lambda)`` plus ``gphase`` (qiskit: ``UnitaryGate(matrix)``);
 * every control becomes one gate modifier, ``ctrl @`` for CONTROL1 and ``negctrl @`` for CONTROL0, in the
   gate's own qubit order (qiskit: ``.control(n, ctrl_state=...)``, 57-63);
 * the gate's phase factor (``CtrlGate.phase()``, cgate.py:40-41) multiplies only the controlled block, so it is
   emitted as a ``gphase`` under the same modifiers (on an uncontrolled gate it is a true global phase and is
   still written, to keep the text an exact description of the operator).
Convention for ``U``: OpenQASM 3.0, ``U(t,p,l) = [[cos(t/2), -e^{il} sin(t/2)], [e^{ip} sin(t/2), e^{i(p+l)} cos(t/2)]]``.
``UnitaryM`` gates are ignored with a warning, as the Cirq and Qiskit builders do (cirq_builder.py:57).
Gates with two or more targets cannot be written without a decomposition and raise ``NotImplementedError``.
"""
import cmath
import math
import re
import warnings

import numpy as np

from ..construct.qspace import Qubit, qubit_key
from ..construct.types import CONTROL0, CONTROL1, UnivGate
from ..program import Program
from .qbuilder import CircuitBuilder

_STD_NAMES = {"I": "id", "X": "x", "Y": "y", "Z": "z", "H": "h", "S": "s", "T": "t", "SD": "sdg", "TD": "tdg"}
_EPS = 1e-14


def zyz_angles(m):
    """(theta, phi, lam, gamma) with ``m == exp(i gamma) * U(theta, phi, lam)`` for a 2x2 unitary ``m``."""
    m = np.asarray(m, dtype=np.complex128)
    a, b, c, d = m[0, 0], m[0, 1], m[1, 0], m[1, 1]
    theta = 2.0 * math.atan2(abs(c), abs(a))
    if abs(a) > _EPS and abs(c) > _EPS:
        gamma = cmath.phase(a)
        phi = cmath.phase(c) - gamma
        lam = cmath.phase(-b) - gamma
    elif abs(c) <= _EPS:      # diagonal: U = diag(1, e^{i(phi+lam)})
        gamma = cmath.phase(a)
        phi = 0.0
        lam = cmath.phase(d) - gamma
        theta = 0.0
    else:                     # anti-diagonal: U = [[0, -e^{il}], [e^{ip}, 0]]
        lam = 0.0
        gamma = cmath.phase(-b)
        phi = cmath.phase(c) - gamma
        theta = math.pi
    return theta, phi, lam, gamma


def u_matrix(theta, phi, lam):
    c, s = math.cos(theta / 2), math.sin(theta / 2)
    return np.array([[c, -cmath.exp(1j * lam) * s], [cmath.exp(1j * phi) * s, cmath.exp(1j * (phi + lam)) * c]])


def _num(x: float) -> str:
    return repr(float(x))  # shortest text that round-trips the double


class QasmCircuit:
    """What ``QasmBuilder.finish`` returns: the program text plus the register it is written over."""

    def __init__(self, lines, qubits):
        self.qubits = list(qubits)
        self._body = list(lines)

    def all_qubits(self):
        return list(self.qubits)

    def __len__(self):
        return sum(1 for ln in self._body if "gphase(" not in ln)

    def qasm(self) -> str:
        nq = sum(1 for q in self.qubits if not q.ancilla)
        na = len(self.qubits) - nq
        head = ["OPENQASM 3.0;", 'include "stdgates.inc";']
        if nq:
            head.append(f"qubit[{nq}] q;")
        if na:
            head.append(f"qubit[{na}] a;")
        return "\n".join(head + self._body) + "\n"

    __str__ = qasm

    def program(self, register=None, ancilla_as_plain=False) -> Program:
        """The gate program the text describes (parsed back from the text), index bits following ``register``."""
        return parse_qasm(self.qasm(), self.qubits, register, ancilla_as_plain)

    def unitary(self, qubit_order=None, drop_ancilla=False, fuse=True) -> np.ndarray:
        """Dense operator of the TEXT, evaluated by the CUDA engine -- the check ``verify_circuit`` implementations
        make on a built circuit (``tests/qompiler/circuits/test_cirq_circuit.py:39``).  With ``drop_ancilla`` the
        result is <0|U|0> on the ancillas, else ancillas count as ordinary qubits."""
        from ..engine import simulate_columns
        prog = self.program(qubit_order, ancilla_as_plain=not drop_ancilla)
        n_free = prog.n_qubits - bin(prog.ancilla_mask).count("1")
        return simulate_columns(prog, 0, n_free, fuse=fuse, drop_ancilla=True)


class QasmBuilder(CircuitBuilder):
    def __init__(self):
        self._qubits = set()
        self._gates = []

    def register(self, qspace) -> None:
        self._qubits.update(qspace)  # may be called repeatedly with overlapping sets (circuit_test_template.py:93-95)

    def all_qubits(self) -> list:
        return sorted(self._qubits, key=qubit_key)

    def build_gate(self, m) -> None:
        if not hasattr(m, "controls"):
            warnings.warn(f"Warning: gate of type {type(m)} is ignored.")
            return
        self._qubits.update(m.qspace)
        self._gates.append(m)

    def finish(self, optimized=False) -> QasmCircuit:
        qubits = self.all_qubits()
        name = {}
        iq = ia = 0
        for q in qubits:
            if q.ancilla:
                name[q] = f"a[{ia}]"
                ia += 1
            else:
                name[q] = f"q[{iq}]"
                iq += 1
        lines = []
        for g in self._gates:
            lines.extend(self._emit(g, name))
        return QasmCircuit(lines, qubits)

    @staticmethod
    def _emit(g, name):
        mods, ctrl_ops, targets = [], [], []
        for q, c in zip(g.qspace, g.controls()):
            code = int(c)
            if code == CONTROL1:
                mods.append("ctrl @ ")
                ctrl_ops.append(name[q])
            elif code == CONTROL0:
                mods.append("negctrl @ ")
                ctrl_ops.append(name[q])
            else:
                targets.append(name[q])
        if len(targets) != 1:
            raise NotImplementedError(f"OpenQASM export needs one-target gates (decompose first); this one has {len(targets)}")
        mod = "".join(mods)
        ops = ", ".join(ctrl_ops + targets)
        phase = complex(g.phase())
        gamma = 0.0
        if g.is_std():
            call = _STD_NAMES[g.gate.name]
            gamma = cmath.phase(phase)
        elif g.is_principal():
            axis = g.gate.axis.principal.name.lower()
            call = f"r{axis}({_num(g.gate.angle)})"
            gamma = cmath.phase(phase)
        else:
            theta, phi, lam, gamma = zyz_angles(np.asarray(g.matrix(), dtype=np.complex128) * phase)
            call = f"U({_num(theta)}, {_num(phi)}, {_num(lam)})"
        out = []
        if abs(gamma) > 0.0:
            out.append(f"{mod}gphase({_num(gamma)})" + (f" {', '.join(ctrl_ops)};" if ctrl_ops else ";"))
        out.append(f"{mod}{call} {ops};")
        return out


# ---- reading the text back ------------------------------------------------------------------------------------------
_STD_MATS = {v: np.array(UnivGate[k]) for k, v in _STD_NAMES.items()}
_PAULI = {"x": np.array([[0, 1], [1, 0]], dtype=complex), "y": np.array([[0, -1j], [1j, 0]]), "z": np.array([[1, 0], [0, -1]], dtype=complex)}
_STMT = re.compile(r"^((?:(?:neg)?ctrl\s*@\s*)*)([A-Za-z_][A-Za-z0-9_]*)\s*(?:\(([^)]*)\))?\s*(.*)$")


def _operand_bits(text, n_ordinary, position, n):
    bits = []
    for tok in (t.strip() for t in text.split(",") if t.strip()):
        mm = re.fullmatch(r"([qa])\[(\d+)\]", tok)
        if not mm:
            raise ValueError(f"cannot parse operand {tok!r}")
        slot = int(mm.group(2)) + (n_ordinary if mm.group(1) == "a" else 0)  # sorted register: ordinary qubits first
        bits.append(n - 1 - position[slot])
    return bits


def parse_qasm(text, qubits=None, register=None, ancilla_as_plain=False) -> Program:
    """Parse the OpenQASM 3 subset ``QasmBuilder`` writes into a ``Program`` (time order as written).

    ``qubits``: the circuit's sorted register, ordinary qubits first (defaults to ``q0..`` then ancillas ``a0..``,
    sized by the declarations); ``register``: the order the program's index bits follow (defaults to ``qubits``);
    ``ancilla_as_plain``: build the program over ordinary stand-ins so that ancillas are simulated as qubits."""
    nq = na = 0
    body = []
    for raw in text.split(";"):
        ln = " ".join(x for x in (p.split("//")[0].strip() for p in raw.splitlines()) if x)
        if not ln or ln.startswith("OPENQASM") or ln.startswith("include"):
            continue
        mm = re.fullmatch(r"qubit\[(\d+)\]\s+([qa])", ln)
        if mm:
            if mm.group(2) == "q":
                nq = int(mm.group(1))
            else:
                na = int(mm.group(1))
            continue
        body.append(ln)
    if qubits is None:
        qubits = [Qubit(i) for i in range(nq)] + [Qubit(i, True) for i in range(na)]
    qubits = list(qubits)
    if len(qubits) != nq + na:
        raise ValueError("register does not match the declarations of the text")
    register = list(register) if register is not None else qubits
    if set(register) != set(qubits) or len(register) != len(qubits):
        raise ValueError("qubit order must be a permutation of the circuit's qubits")
    n = len(qubits)
    where = {q: i for i, q in enumerate(register)}
    position = [where[q] for q in qubits]
    prog = Program([Qubit(i) for i in range(n)] if ancilla_as_plain else register)
    for ln in body:
        mm = _STMT.match(ln)
        if not mm:
            raise ValueError(f"cannot parse statement {ln!r}")
        mods = re.findall(r"(neg)?ctrl\s*@", mm.group(1))
        gname, args, ops = mm.group(2), mm.group(3), mm.group(4)
        argv = [float(a) for a in args.split(",")] if args else []
        bits = _operand_bits(ops, nq, position, n)
        n_ctrl = len(mods)
        mask = value = 0
        for neg, b in zip(mods, bits[:n_ctrl]):
            mask |= 1 << b
            if not neg:
                value |= 1 << b
        rest = bits[n_ctrl:]
        if gname == "gphase":
            if rest:
                raise ValueError(f"gphase takes no operand beyond its controls: {ln!r}")
            e = cmath.exp(1j * argv[0])
            if n_ctrl == 0:  # a true global phase: kept, so that the program is the operator the text describes exactly
                prog.add_raw([0], 0, 0, np.diag([e, e]))
                continue
            # a phase on the controlled block == diag(1, e) (or diag(e, 1)) on the last control, controlled by the others
            t = bits[n_ctrl - 1]
            d = np.array([1.0, e]) if value & (1 << t) else np.array([e, 1.0])
            prog.add_raw([t], mask & ~(1 << t), value & ~(1 << t), np.diag(d))
            continue
        if len(rest) != 1:
            raise ValueError(f"one target expected: {ln!r}")
        if gname in _STD_MATS:
            m = _STD_MATS[gname]
        elif gname in ("rx", "ry", "rz"):
            m = math.cos(argv[0] / 2) * np.eye(2) - 1j * math.sin(argv[0] / 2) * _PAULI[gname[1]]
        elif gname == "U":
            m = u_matrix(*argv)
        else:
            raise ValueError(f"unknown gate {gname!r}")
        prog.add_raw(rest, mask, value, m)
    return prog
