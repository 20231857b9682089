"""Import alias for the package directory ``quantum-simulations_b200/``.

The directory name mandated by the repo layout contains a hyphen, which Python
cannot import directly.  Importing this module loads that directory as the
package ``quantum_simulations_b200`` (sub-modules resolve through its
``__path__`` as usual) and replaces this stub in ``sys.modules``.
"""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "quantum-simulations_b200")
_spec = importlib.util.spec_from_file_location(
    __name__, os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
