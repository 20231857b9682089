"""Reading the reference's compiled-code files (``.qco``) without the reference installed.

``Qompiler.output`` pickles the Bytecode tree (``quompiler/utils/file_io.py:9-21``), so the file embeds the
module paths ``quompiler.construct.{bytecode,cgate,unitary,qspace,types,su2gate}``.  When the real package is
importable those pickles load as the reference's own objects (which the builder accepts by duck typing).  When it
is not -- e.g. on a GPU box that only has this repo -- ``install_quompiler_alias()`` registers this package's
mirror modules under those names so that the same file loads into the mirror classes (their ``__setstate__``
methods translate the reference's attribute layout).
"""
import importlib
import sys
import types

_MAP = {
    "quompiler.construct.bytecode": "quantum_simulations_b200.construct.bytecode",
    "quompiler.construct.cgate": "quantum_simulations_b200.construct.cgate",
    "quompiler.construct.unitary": "quantum_simulations_b200.construct.unitary",
    "quompiler.construct.qspace": "quantum_simulations_b200.construct.qspace",
    "quompiler.construct.types": "quantum_simulations_b200.construct.types",
    "quompiler.construct.su2gate": "quantum_simulations_b200.construct.su2gate",
    "quompiler.construct.qontroller": "quantum_simulations_b200.construct.qontroller",
}


def install_quompiler_alias(force: bool = False) -> bool:
    """Make ``import quompiler.construct.*`` resolve to the mirror modules.  Returns True if installed;
    does nothing (returns False) when the real quompiler is importable, unless ``force``."""
    if not force:
        try:
            import quompiler.construct.cgate  # noqa: F401
            return False
        except Exception:
            pass
    root = sys.modules.get("quompiler") or types.ModuleType("quompiler")
    root.__path__ = []
    sys.modules["quompiler"] = root
    sub = types.ModuleType("quompiler.construct")
    sub.__path__ = []
    sys.modules["quompiler.construct"] = sub
    root.construct = sub
    for alias, real in _MAP.items():
        mod = importlib.import_module(real)
        sys.modules[alias] = mod
        setattr(sub, alias.rsplit(".", 1)[1], mod)
    return True


def read_qco(path):
    """Load a ``.qco`` file into a Bytecode tree (reference objects if quompiler is installed, mirror objects otherwise)."""
    install_quompiler_alias()
    from .circuits.render import read_code
    return read_code(path)
