"""``SlidingWindowCombiner`` (mirror of ``quompiler/optimize/basic_optimizer.py:11-72``): annihilate, then consolidate,
then prune the skipped nodes.  HXH -> Z, TT -> S, SS -> Z, XX -> nothing, ..."""
from ..construct.types import GateGrain
from .window import AnnihilateOperator, ConsolidateOperator


class SlidingWindowCombiner:
    def __init__(self, window: int = 0, emit=GateGrain.CLIFFORD_T):
        """``window`` = how many predecessors a leaf may combine with (0 = unbounded); ``emit`` = the coarsest grain a
        combined gate may have."""
        self.operators = [AnnihilateOperator(window, emit), ConsolidateOperator(window, emit)]

    def optimize(self, root):
        for op in self.operators:
            self.iterate(root, op)
        prune(root)
        root.skip = False
        return root

    @staticmethod
    def iterate(root, operator):
        stack = [root]
        while stack:
            node = stack.pop()
            if node.is_leaf() and not node.skip:
                operator.run(node)
            stack.extend(reversed(node.children))


def prune(node):
    """Drop skipped nodes; an inner node left with one child is replaced by it (basic_optimizer.py:52-72)."""
    if node.skip:
        return None
    if node.is_leaf():
        return node
    children = []
    for child in node.children:
        pruned = prune(child)
        if pruned:
            children.append(pruned)
    if not children:
        return None
    if len(children) == 1:
        return children[0]
    node.children = children
    return node
