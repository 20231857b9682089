import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from quantum_simulations_b200.utils.mgen import qft_program, random_state
from quantum_simulations_b200.engine import apply_circuit, get_engine
from quantum_simulations_b200.program import Program
from quantum_simulations_b200.construct.qspace import Qubit
from quantum_simulations_b200.construct.types import UnivGate
np.random.seed(1)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 11
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 4
x = np.stack([random_state(1 << n) for _ in range(nb)], axis=1)
for bit in (n - 1, n - 2, 5, 0):
    p2 = Program([Qubit(i) for i in range(n)])
    p2.add_raw([bit], 0, 0, np.array(UnivGate.H))
    a = apply_circuit(p2, x, fuse=True); b = apply_circuit(p2, x, fuse=False)
    d = np.abs(a - b)
    print("H on bit", bit, "fused-vs-unfused", d.max(), "fused-vs-input", np.abs(a - x).max(), "bad rows", np.unique(np.nonzero(d > 1e-12)[0]).size, "of", 1 << n,
          "bad cols", np.unique(np.nonzero(d > 1e-12)[1]))
