"""Replay a compiled Bytecode tree into a ``CircuitBuilder``.

Same protocol as ``QRenderer.render`` (``quompiler/circuits/render.py:12-32``):
register the sorted qubit set found in the tree, then walk it right-to-left
(circuit time order), sending leaves to ``build_gate`` and inner nodes to
``build_group``; nodes flagged ``skip`` are ignored.  Accepts a ``.qco`` path
(pickle, ``quompiler/utils/file_io.py:9-21``) or an in-memory tree.
"""
import pickle

from ..construct.bytecode import BytecodeRevIter
from ..construct.qspace import qubit_key

CODE_FILE_EXT = ".qco"


def read_code(path):
    with open(path, "rb") as f:
        code = pickle.load(f)
    if not hasattr(code, "children") or not hasattr(code, "data"):
        raise RuntimeError(f"{path} does not hold a Bytecode tree")
    return code


def write_code(path, code):
    with open(path, "wb") as f:
        pickle.dump(code, f)


def collect_qspace(code):
    qs = set()
    for node in BytecodeRevIter(code):
        if hasattr(node.data, "qspace"):
            qs.update(node.data.qspace)
    return sorted(qs, key=qubit_key)


class QRenderer:
    def __init__(self, builder):
        self.builder = builder

    def render(self, codefile):
        code = read_code(codefile) if isinstance(codefile, (str, bytes)) or hasattr(codefile, "__fspath__") else codefile
        self.builder.register(collect_qspace(code))
        for node in BytecodeRevIter(code):
            if node.skip:
                continue
            if node.is_leaf():
                self.builder.build_gate(node.data)
            else:
                self.builder.build_group(node.data)
        return self.builder.finish()
