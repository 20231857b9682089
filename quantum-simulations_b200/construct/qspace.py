"""Qubit identity and register ordering.

Mirrors the reference's ``quompiler/construct/qspace.py:5-44``: a qubit is an
``(qid, ancilla)`` pair, ordinary qubits sort before ancillas and then by id,
and prints as ``q3`` / ``a102``.  The sorted register defines the index bits of
the state vector (first qubit = most significant bit).
"""
from functools import total_ordering


@total_ordering
class Qubit:
    __slots__ = ("qid", "ancilla", "_h")

    def __init__(self, qid: int, ancilla: bool = False):
        assert qid >= 0, f'Qubit {qid} must be >= 0'
        self.qid = int(qid)
        self.ancilla = bool(ancilla)
        # same value as the reference (qspace.py:27-28) so mixed dict lookups agree; cached: gate streams hash
        # their qubits millions of times
        self._h = hash(self.qid) + hash(self.ancilla)

    def _key(self):
        return (self.ancilla, self.qid)

    def __eq__(self, other):
        if other.__class__ is Qubit:
            return self.qid == other.qid and self.ancilla == other.ancilla
        if not _is_qubit(other):
            return NotImplemented
        return self._key() == (bool(other.ancilla), int(other.qid))

    def __lt__(self, other):
        if not _is_qubit(other):
            return NotImplemented
        return self._key() < (bool(other.ancilla), int(other.qid))

    def __hash__(self):
        return self._h

    def __repr__(self):
        return f"{'a' if self.ancilla else 'q'}{self.qid}"

    __str__ = __repr__

    def to_dict(self):
        return {"qid": self.qid, "ancilla": self.ancilla}

    def __reduce__(self):
        return (Qubit, (self.qid, self.ancilla))

    def __setstate__(self, state):
        # a Qubit pickled by the reference (a dataclass: plain attribute dict) loading into this class
        if isinstance(state, tuple):
            state = state[1] or state[0]
        self.qid = int(state["qid"])
        self.ancilla = bool(state["ancilla"])
        self._h = hash(self.qid) + hash(self.ancilla)


def _is_qubit(x) -> bool:
    return hasattr(x, "qid") and hasattr(x, "ancilla")


def qubit_key(q):
    """Sort key usable for our Qubit and for the reference's dataclass alike."""
    return (bool(q.ancilla), int(q.qid))
