"""Peephole optimiser over flattened gate records, with its equivalence check run on the GPU."""
from .peephole import RecordPeephole, grain_of, optimize_leaves, peephole_records
from .verify import equal_on_device, optimize_checked

__all__ = ["RecordPeephole", "optimize_leaves", "peephole_records", "grain_of", "equal_on_device", "optimize_checked"]
