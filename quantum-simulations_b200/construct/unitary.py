"""``UnitaryM``: identity of order ``dimension`` except on the ``core`` rows/columns.

Mirror of ``quompiler/construct/unitary.py:20-110`` (same constructor, same
validation errors, ``phase`` multiplies only the core block, line 76).  The dense
work the reference does in ``inflate`` (70-77) and ``__matmul__`` (56-62) is not
done on the host here: ``@`` builds a lazy :class:`GateChain` and
``inflate()`` / ``np.array(...)`` run the chain on the CUDA engine.
"""
import numpy as np


class UnitaryM:
    def __init__(self, dimension: int, core, matrix, phase=1):
        matrix = np.asarray(matrix)
        assert np.isclose(abs(phase), 1), 'phase factor must be normalized.'
        s = matrix.shape
        assert len(s) == 2, f'Matrix must be 2D array but got {s}.'
        assert s[0] == s[1], f'Matrix must be square but got {s}.'
        assert np.allclose(matrix @ matrix.conj().T, np.eye(s[0])), f'Matrix is not unitary {matrix}'
        assert dimension >= s[0], 'Dimension must be greater than or equal to the dimension of the core matrix.'
        assert len(core) == s[0], 'The number of indexes must match the size of the expansion matrix.'
        assert len(set(core)) == len(core), 'The indexes in core must be unique.'
        self.dimension = int(dimension)
        self.core = tuple(int(c) for c in core)
        self.matrix = matrix
        self.phase = phase

    def order(self) -> int:
        return self.dimension

    def __getitem__(self, index):
        return self.matrix[index]

    def __setitem__(self, index, value):
        self.matrix[index] = value

    def __matmul__(self, other):
        from .chain import GateChain
        if isinstance(other, GateChain):
            return GateChain([self]) @ other
        if not isinstance(other, UnitaryM):
            return NotImplemented
        if self.dimension != other.dimension:
            raise ValueError('matmul: Input operands have dimension mismatch.')
        if self.core == other.core:  # small host-side product, as unitary.py:59-60
            return UnitaryM(self.dimension, self.core, self.matrix @ other.matrix, self.phase * other.phase)
        return GateChain([self, other])

    def inflate(self) -> np.ndarray:
        from .chain import GateChain
        return GateChain([self]).inflate()

    def __array__(self, dtype=None, copy=None):
        m = self.inflate()
        return m if dtype is None else m.astype(dtype)

    def isid(self) -> bool:
        return bool(np.allclose(self.matrix, np.eye(self.matrix.shape[0])))

    def is2l(self) -> bool:
        return len(self.core) <= 2

    def issinglet(self) -> bool:
        d = self.dimension
        if d & (d - 1) or len(self.core) != 2:
            return False
        x = self.core[0] ^ self.core[1]
        return x & (x - 1) == 0

    def herm(self) -> 'UnitaryM':
        return UnitaryM(self.dimension, self.core, np.conj(self.matrix.T), np.conj(self.phase))

    def __repr__(self):
        return f'UnitaryU{{dimension={self.dimension}, core={self.core}, matrix={self.matrix}}}'
