"""Index conventions of controlled gates (big-endian: first control = most significant bit).

Restates the two helpers of ``quompiler/construct/qontroller.py`` that define
the bit conventions of the simulation path: ``ctrl2core`` (112-120) and
``core2control`` (86-109).
"""
from .types import QType, CONTROL0, CONTROL1, TARGET


def ctrl2core(controls) -> list:
    """Basis indexes a gate with this role sequence acts on, ascending in the
    free (TARGET/IDLER) bits, with control bits pinned to their activation value."""
    k = len(controls)
    fixed = 0
    free = []
    for pos, c in enumerate(controls):
        bit = k - 1 - pos
        code = int(c)
        if code == CONTROL1:
            fixed |= 1 << bit
        elif code != CONTROL0:
            free.append(bit)
    out = []
    for v in range(1 << len(free)):
        idx = fixed
        for j, bit in enumerate(free):  # free[0] is the most significant free bit
            if (v >> (len(free) - 1 - j)) & 1:
                idx |= 1 << bit
        out.append(idx)
    return out


def core2control(bitlength: int, core) -> tuple:
    """Role sequence spanned by a set of indexes: bits shared by all of them become
    CONTROL0/CONTROL1 (by their common value), bits that differ become TARGET."""
    core = list(core)
    assert core, 'Core cannot be empty'
    assert max(core) < (1 << bitlength), 'Invalid core: index exceeds the register'
    all_and = core[0]
    all_or = core[0]
    for c in core[1:]:
        all_and &= c
        all_or |= c
    roles = []
    for bit in range(bitlength - 1, -1, -1):
        m = 1 << bit
        if (all_or & m) and not (all_and & m):
            roles.append(QType.TARGET)
        else:
            roles.append(QType.CONTROL1 if all_and & m else QType.CONTROL0)
    return tuple(roles)
