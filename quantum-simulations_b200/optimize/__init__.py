"""Window optimiser over compiled gate trees, with its equivalence check run on the GPU."""
from .basic_optimizer import SlidingWindowCombiner, prune
from .verify import equal_on_device, optimize_checked
from .window import AnnihilateOperator, ConsolidateOperator

__all__ = ["SlidingWindowCombiner", "prune", "AnnihilateOperator", "ConsolidateOperator", "equal_on_device", "optimize_checked"]
