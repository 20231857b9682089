"""Grain of a gate if it were emitted as is (mirror of ``quompiler/utils/granularity.py:8-49``)."""
from ..construct.types import GateGrain, UnivGate


def granularity(obj) -> GateGrain:
    if hasattr(obj, "controls"):
        return granularity_gate(obj)
    if hasattr(obj, "core") and hasattr(obj, "dimension"):
        return GateGrain.TWO_LEVEL if obj.is2l() else GateGrain.UNITARY
    return GateGrain.INVALID


def granularity_gate(g) -> GateGrain:
    if g.is_std():
        if 1 < len(g.control_qids()):
            return GateGrain.SINGLET
        if g.gate in UnivGate.cliffordt():
            return GateGrain.CLIFFORD_T
        return GateGrain.CTRL_PRUNED
    if g.is_principal():
        return GateGrain.PRINCIPAL
    if not g.issinglet():
        return GateGrain.MULTI_TARGET
    if len(g.control_qids()) < 2:
        return GateGrain.CTRL_PRUNED
    return GateGrain.SINGLET
