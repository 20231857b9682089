"""Tree of compiled gates and its two traversal orders.

Mirror of ``quompiler/construct/bytecode.py:13-88``.  ``BytecodeIter`` walks
pre-order left-to-right, which is *matrix-product* order (first leaf is the
left-most factor, so it acts LAST on a state); ``BytecodeRevIter`` walks
right-to-left, which is circuit *time* order and is what the renderer feeds to a
``CircuitBuilder`` (``quompiler/circuits/render.py:25``).
"""


class Bytecode:
    def __init__(self, data, children=None, metadata=None, skip=False):
        self.data = data
        self.children = list(children) if children else []
        self.metadata = metadata or {}
        self.skip = skip

    def add_child(self, child: 'Bytecode'):
        self.children.append(child)

    def is_leaf(self) -> bool:
        return not self.children

    def herm(self) -> 'Bytecode':
        import numpy as np
        data = np.conj(self.data.T) if isinstance(self.data, np.ndarray) else self.data.herm()
        return Bytecode(data, [c.herm() for c in reversed(self.children)])

    def __repr__(self):
        return f'Bytecode{{{self.skip},{self.data!r}}}'


def _walk(root, leftmost_first: bool):
    stack = [root]
    while stack:
        node = stack.pop()
        stack.extend(reversed(node.children) if leftmost_first else node.children)
        yield node


class BytecodeIter:
    def __init__(self, root):
        self._it = _walk(root, True)

    def __iter__(self):
        return self

    def __next__(self):
        return next(self._it)


class BytecodeRevIter:
    def __init__(self, root):
        self._it = _walk(root, False)

    def __iter__(self):
        return self

    def __next__(self):
        return next(self._it)
