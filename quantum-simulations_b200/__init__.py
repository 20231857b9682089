"""quantum_simulations_b200 -- a B200 (sm_100a) statevector engine behind quompiler's
``CircuitBuilder`` plug-in API.  See DESIGN.md; the C ABI is ``include/qsv.h``."""
from .construct.bytecode import Bytecode, BytecodeIter, BytecodeRevIter
from .construct.cgate import CtrlGate
from .construct.chain import GateChain
from .construct.qspace import Qubit
from .construct.su2gate import RGate
from .construct.types import GateGrain, PrincipalAxis, QompilePlatform, QType, UnivGate
from .construct.unitary import UnitaryM
from .circuits.b200_builder import B200Builder, B200Circuit
from .circuits.qasm_builder import QasmBuilder, QasmCircuit, parse_qasm
from .circuits.qbuilder import CircuitBuilder
from .circuits.qfactory import get_builder, get_render
from .circuits.render import QRenderer

__all__ = ["Bytecode", "BytecodeIter", "BytecodeRevIter", "CtrlGate", "GateChain", "Qubit", "RGate", "GateGrain",
           "PrincipalAxis", "QompilePlatform", "QType", "UnivGate", "UnitaryM", "B200Builder", "B200Circuit",
           "CircuitBuilder", "get_builder", "get_render", "QRenderer", "QasmBuilder", "QasmCircuit", "parse_qasm"]
