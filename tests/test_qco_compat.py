"""Compiled-code files written by the reference (tests/golden/*.qco, pickled Bytecode trees) load without the
reference installed and render into the B200 builder."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

_CHILD = r'''
import sys
sys.path.insert(0, %(root)r)
assert not any("reference" in p for p in sys.path)
import numpy as np
import quantum_simulations_b200 as qs
from quantum_simulations_b200.compat import install_quompiler_alias, read_qco
assert install_quompiler_alias() is True            # the real quompiler is not importable here
from oracle import sv_oracle as orc
d = np.load(%(inputs)r)
for name in ("qco_r4", "qco_cyc8", "qco_r8p"):
    code = read_qco(%(golden)r + "/" + name + ".qco")
    assert type(code).__module__.startswith("quantum_simulations_b200")
    b = qs.B200Builder()
    circ = qs.QRenderer(b).render(code)
    assert all(type(g).__module__.startswith("quantum_simulations_b200") for g in circ.gates)
    prog = circ.program()
    n, am = prog.n_qubits, prog.ancilla_mask
    keep = [i for i in range(1 << n) if not (i & am)]
    k = len(keep).bit_length() - 1
    st = np.zeros((1 << n, 1 << k), dtype=np.complex128)
    st[keep, np.arange(len(keep))] = 1
    flat = st.reshape(-1)
    orc.run_records(prog.records(k), prog.pool(), flat, n + k)
    got = flat.reshape(1 << n, 1 << k)[keep, :]
    err = np.abs(got - d[name + "_input"]).max()
    assert err < 1e-10, (name, err)
    if name == "qco_cyc8":
        assert np.array_equal(got, d[name + "_input"])
    print(name, len(circ), [repr(q) for q in circ.all_qubits()], err)
print("ok")
'''


def test_reference_qco_files_load_through_the_alias_and_render():
    code = _CHILD % {"root": ROOT, "inputs": os.path.join(GOLDEN, "qco_inputs.npz"), "golden": GOLDEN}
    env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, env=env, cwd="/tmp")
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stdout[-1500:] + out.stderr[-3000:]


@pytest.mark.gpu
def test_qco_files_on_the_gpu():
    """Same files, evaluated by the CUDA engine (<0|U|0> on the ancillas == the compiled unitary)."""
    import quantum_simulations_b200 as qs
    from quantum_simulations_b200.compat import read_qco
    d = np.load(os.path.join(GOLDEN, "qco_inputs.npz"))
    for name in ("qco_r4", "qco_cyc8", "qco_r8p"):
        b = qs.B200Builder()
        circ = qs.QRenderer(b).render(read_qco(os.path.join(GOLDEN, name + ".qco")))
        for fuse in (True, False):
            got = circ.unitary(drop_ancilla=True, fuse=fuse)
            if name == "qco_cyc8":
                assert np.array_equal(got, d[name + "_input"])
            else:
                assert np.abs(got - d[name + "_input"]).max() < 1e-10
