"""Back-end dispatch on ``QompilePlatform`` (reference: ``quompiler/circuits/qfactory.py:36-43``).

Only the new target is constructed here; the Cirq/Qiskit/Quimb back-ends live in
third-party packages that are out of scope (and absent from this image), so asking
for them raises the same ``ValueError`` the reference raises for an unknown platform.
"""
from ..construct.types import QompilePlatform
from .b200_builder import B200Builder
from .qasm_builder import QasmBuilder
from .render import QRenderer


def get_builder(target):
    name = getattr(target, "name", str(target))
    if name == QompilePlatform.B200.name:
        return B200Builder()
    if name == QompilePlatform.QASM.name:
        return QasmBuilder()
    raise ValueError(f"Unsupported platform {target}")


def get_render(target) -> QRenderer:
    return QRenderer(get_builder(target))
