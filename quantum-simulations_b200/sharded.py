"""State-vector sharding across GPUs by the top index bits.

With ``N = 2^g`` ranks the ``2^n`` amplitudes are split by their top ``g`` index bits: rank ``r``
holds the ``2^(n-g)`` amplitudes whose top bits equal ``r``.  A gate is local unless its target sits
on one of those *global* bits; then

 * a control on a global bit costs nothing: the rank either takes part or skips the gate;
 * a diagonal gate with a global target costs nothing: the rank's bit picks ``d0`` or ``d1`` and the
   gate becomes a scalar phase on the (locally) controlled subspace;
 * any other gate needs ONE exchange step: all g global bits trade places with the top g local bits
   (an NCCL all-to-all over NVLink through ``torch.distributed``; with one global bit this is the
   pairwise half-state swap with partner ``rank ^ 1``, which ``TorchExchanger.exchange`` also offers for
   a single bit), after which the qubits are local; a bit-position table tracks the remap and the
   evicted bits are chosen by furthest next use.

The planner below is host logic and identical on every rank (the exchange sequence must be); it cuts
the gate stream into local segments separated by exchanges and restores the canonical layout at the
end.  Local segments run on the CUDA engine (fused tile sweeps); the same plan can be driven by any
executor, which is how the world-size-2 ``gloo`` tests exercise it on CPU.

There is no reference counterpart: the reference is single-process numpy.
"""
import json
import os
import sys

import numpy as np

from ._lib import DIAG1, GATE_DTYPE, PERMX, ROWS2


# ---------------------------------------------------------------------------------------------
# planning (rank independent)
# ---------------------------------------------------------------------------------------------
class ShardPlan:
    """steps: ("local", [record indexes]) | ("xall",) | ("exchange", global_bit) | ("permute", perm_list)."""

    def __init__(self, n_bits, g_bits):
        self.n_bits = n_bits
        self.g_bits = g_bits
        self.n_loc = n_bits - g_bits
        self.steps = []
        self.where_at = []  # bit-position table valid for each "local" step (logical bit -> physical position)
        self.relabelled = 0  # source gates absorbed as SWAP relabels

    @property
    def n_exchanges(self):
        return sum(1 for s in self.steps if s[0] in ("exchange", "xall"))


def plan_sharded(recs, n_bits, g_bits, restore=True, relabel_swaps=True) -> ShardPlan:
    """Cut the gate stream into local segments separated by exchange steps.

    Exchange step ("xall"): the g global bit positions trade places with the top g local positions
    (global i <-> local n_loc-g+i) -- an all-to-all in which every rank keeps 1/2^g of its shard.
    Which logical bits get evicted is decided by furthest next use (Belady): before an exchange a
    local bit permutation (one gate-free tile sweep) parks the g local bits that will not be needed as
    a non-diagonal target for the longest time in the top local positions.
    """
    plan = ShardPlan(n_bits, g_bits)
    n_loc, g = plan.n_loc, g_bits
    if n_loc < max(1, 2 * g):
        raise ValueError("need at least 2*log2(world) local bits per rank")
    top = list(range(n_loc - g, n_loc))
    where = list(range(n_bits))  # logical bit -> physical position
    who = list(range(n_bits))    # physical position -> logical bit
    cur = []

    def flush():
        if cur:
            plan.steps.append(("local", list(cur)))
            plan.where_at.append(list(where))
            cur.clear()

    def swap_positions(p1, p2):
        b1, b2 = who[p1], who[p2]
        who[p1], who[p2] = b2, b1
        where[b1], where[b2] = p2, p1

    def park_at_top(bits):
        """Local permutation putting the given logical bits (all local) at the top-g local positions."""
        bits = list(bits)
        assert len(bits) == g and all(where[b] < n_loc for b in bits)
        if set(who[p] for p in top) == set(bits):
            return
        flush()
        perm = list(range(n_loc))          # perm[i] = destination of the index bit now at position i
        free_top = [p for p in top if who[p] not in bits]
        for b in bits:
            p = where[b]
            if p in top:
                continue
            q = free_top.pop()
            perm[p], perm[q] = q, p
            swap_positions(p, q)
        plan.steps.append(("permute", perm))

    def xall():
        flush()
        plan.steps.append(("xall",))
        for i in range(g):
            swap_positions(n_loc + i, n_loc - g + i)

    def cnot(r):
        if int(r["kind"]) != PERMX or int(r["n_targets"]) != 1:
            return None
        cm = int(r["ctrl_mask"])
        if bin(cm).count("1") != 1 or int(r["ctrl_value"]) != cm:
            return None
        return cm.bit_length() - 1, int(r["target_bits"][0])

    # classify records once: SWAP triples (relabels) and the non-diagonal target uses of every logical bit
    n_recs = len(recs)
    swap_at = {}
    i = 0
    while relabel_swaps and i + 2 < n_recs:
        c0, c1, c2 = cnot(recs[i]), cnot(recs[i + 1]), cnot(recs[i + 2])
        if c0 and c1 and c2 and c1 == (c0[1], c0[0]) and c2 == c0:
            swap_at[i] = c0
            i += 3
        else:
            i += 1
    uses = [[] for _ in range(n_bits)]     # per logical bit: record indexes where it is a non-diagonal target
    # (uses are recorded against the logical bit named in the record; relabels rename data, see next_use)
    skip = set()
    for i0, (a, b) in swap_at.items():
        skip.update((i0, i0 + 1, i0 + 2))
    for i, r in enumerate(recs):
        if i in skip or int(r["kind"]) in (DIAG1, ROWS2):
            continue
        for tb in r["target_bits"][:int(r["n_targets"])]:
            uses[int(tb)].append(i)

    import bisect

    def next_use(b, i):
        """Index of the next record (>= i) that needs logical bit b as a non-diagonal target.  SWAP relabels
        between now and then rename the data, so follow them: the data now called b is called tau(b) later."""
        name = b
        # walk forward through relabels in order; between relabels look up uses of the current name
        pos_i = i
        for s_i in sorted(k_ for k_ in swap_at if k_ >= i):
            lst = uses[name]
            j = bisect.bisect_left(lst, pos_i)
            if j < len(lst) and lst[j] < s_i:
                return lst[j]
            a, b2 = swap_at[s_i]
            if name == a:
                name = b2
            elif name == b2:
                name = a
            pos_i = s_i
        lst = uses[name]
        j = bisect.bisect_left(lst, pos_i)
        return lst[j] if j < len(lst) else n_recs + 1

    i = 0
    while i < n_recs:
        r = recs[i]
        if i in swap_at:
            flush()  # gates already collected were mapped with the table as it was
            a, b = swap_at[i]
            swap_positions(where[a], where[b])
            plan.relabelled += 3
            i += 3
            continue
        kind = int(r["kind"])
        if kind == ROWS2:
            raise NotImplementedError("row (two-level) records are column-sharded, not state-sharded")
        t = int(r["n_targets"])
        tbits = [int(b) for b in r["target_bits"][:t]]
        if kind != DIAG1 and any(where[tb] >= n_loc for tb in tbits):
            # evict the g local bits whose next non-diagonal use is furthest away (never the gate's own targets)
            cand = [who[p] for p in range(n_loc) if who[p] not in tbits]
            cand.sort(key=lambda b: -next_use(b, i))
            park_at_top(cand[:g])
            xall()
            assert all(where[tb] < n_loc for tb in tbits)
        cur.append(i)
        i += 1
    flush()
    if restore:
        want_global = list(range(n_loc, n_bits))
        if any(who[p] != p for p in want_global):
            if any(where[b] >= n_loc for b in want_global):
                # some of the bits that belong on global positions are global already (maybe misplaced): bring
                # everything local first, evicting bits that do not belong up there
                victims = [who[p] for p in range(n_loc) if who[p] not in want_global][:g]
                park_at_top(victims)
                xall()
            park_in_order = [n_loc + i for i in range(g)]
            # top local position n_loc-g+i must hold logical bit n_loc+i before the exchange
            if [who[p] for p in top] != park_in_order:
                flush()
                perm = list(range(n_loc))
                # build the permutation position by position
                dest = {}
                for i_, b in enumerate(park_in_order):
                    dest[where[b]] = top[i_]
                taken = set(dest.values())
                free = [p for p in range(n_loc) if p not in taken]
                movers = [p for p in range(n_loc) if p not in dest]
                # keep everything else in place when possible
                remaining = [p for p in movers if p in free]
                for p in remaining:
                    dest[p] = p
                    free.remove(p)
                for p in movers:
                    if p not in dest:
                        dest[p] = free.pop()
                perm = [dest[p] for p in range(n_loc)]
                newwho = list(who)
                for p in range(n_loc):
                    newwho[perm[p]] = who[p]
                for p in range(n_loc):
                    who[p] = newwho[p]
                    where[who[p]] = p
                plan.steps.append(("permute", perm))
            xall()
        if who[:n_loc] != list(range(n_loc)):
            flush()
            plan.steps.append(("permute", [who[p] for p in range(n_loc)]))
            for p in range(n_loc):
                who[p] = p
                where[p] = p
    plan.final_where = list(where)
    return plan


def localize_records(recs, pool, idxs, where, n_loc, rank):
    """Records of one local segment as this rank executes them (bit positions -> local physical
    positions, global controls resolved against the rank, global diagonal targets -> scalar phases)."""
    out = []
    extra = []
    pool_len = len(pool)
    for i in idxs:
        r = recs[i]
        kind = int(r["kind"])
        t = int(r["n_targets"])
        cm, cv = int(r["ctrl_mask"]), int(r["ctrl_value"])
        nm = nv = 0
        skip = False
        b = 0
        while cm >> b:
            if (cm >> b) & 1:
                p = where[b]
                want = (cv >> b) & 1
                if p >= n_loc:
                    if ((rank >> (p - n_loc)) & 1) != want:
                        skip = True
                        break
                else:
                    nm |= 1 << p
                    nv |= want << p
            b += 1
        if skip:
            continue
        rec = np.zeros((), dtype=GATE_DTYPE)
        rec["kind"] = kind
        rec["n_targets"] = t
        rec["ctrl_mask"], rec["ctrl_value"] = nm, nv
        rec["matrix_offset"] = r["matrix_offset"]
        tpos = [where[int(x)] for x in r["target_bits"][:t]]
        if kind == DIAG1 and tpos[0] >= n_loc:
            off = int(r["matrix_offset"])
            side = (rank >> (tpos[0] - n_loc)) & 1
            d = pool[off + 3] if side else pool[off]
            if d == 1.0:
                continue
            free = next(p for p in range(n_loc) if not (nm >> p) & 1)
            rec["target_bits"][0] = free
            rec["matrix_offset"] = pool_len + 4 * len(extra)
            extra.append(np.array([d, 0, 0, d], dtype=np.complex128))
        else:
            assert all(p < n_loc for p in tpos), "planner left a non-diagonal target on a global bit"
            rec["target_bits"][:t] = tpos
        out.append(rec)
    recs_out = np.array(out, dtype=GATE_DTYPE) if out else np.zeros(0, dtype=GATE_DTYPE)
    pool_out = np.concatenate([pool] + extra) if extra else pool
    return recs_out, pool_out


# ---------------------------------------------------------------------------------------------
# execution
# ---------------------------------------------------------------------------------------------
class TorchExchanger:
    """Pairwise half-state swap through torch.distributed (NCCL on GPU tensors, gloo on CPU tensors)."""

    def __init__(self, rank, world):
        import torch.distributed as dist
        self.dist = dist
        self.rank = rank
        self.world = world
        self.bytes_sent = 0

    def exchange(self, state, spare, gbit):
        import torch
        dist = self.dist
        partner = self.rank ^ (1 << gbit)
        half = state.numel() // 2
        mine = (self.rank >> gbit) & 1
        send = state[half:] if mine == 0 else state[:half]
        recv = spare[:half]
        ops = [dist.P2POp(dist.isend, torch.view_as_real(send), partner), dist.P2POp(dist.irecv, torch.view_as_real(recv), partner)]
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        send.copy_(recv)
        self.bytes_sent += send.numel() * send.element_size()
        return state

    def exchange_all(self, state, spare):
        if self.dist.get_backend() == "gloo":
            # gloo has no all_to_all: emulate with pairwise chunk swaps (CPU tests only)
            import torch
            w = self.world
            inp = torch.view_as_real(state).view(w, -1)
            out = torch.view_as_real(spare).view(w, -1)
            out[self.rank].copy_(inp[self.rank])
            ops = []
            for peer in range(w):
                if peer != self.rank:
                    ops.append(self.dist.P2POp(self.dist.isend, inp[peer], peer))
                    ops.append(self.dist.P2POp(self.dist.irecv, out[peer], peer))
            for req in self.dist.batch_isend_irecv(ops):
                req.wait()
            res = (spare, state)
        else:
            res = _exchange_all(self.dist, state, spare, self.world)
        self.bytes_sent += state.numel() * state.element_size() * (self.world - 1) // self.world
        return res


def _exchange_all(dist, state, spare, world):
    """All global bits <-> the top local bits: chunk c of this rank goes to rank c (NCCL/gloo all-to-all)."""
    import torch
    inp = torch.view_as_real(state).view(world, -1)
    out = torch.view_as_real(spare).view(world, -1)
    dist.all_to_all_single(out, inp)
    return spare, state


class SymmExchanger:
    """Exchange step as ONE CUDA kernel over peer memory (no NCCL): the gate-free permuting sweep that parks the
    victims is given every rank's destination buffer (torch symmetric memory = peer-mapped allocations) and stores
    each amplitude straight into the rank that will own it (``qsv_permute_exchange``).  Two signal-pad barriers
    bracket the kernel: all ranks have stopped using the destination buffers / all stores have landed."""

    def __init__(self, rank, world, n_amps, device):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        self.rank, self.world = rank, world
        self.g = world.bit_length() - 1
        self.bytes_sent = 0
        group = dist.group.WORLD
        self.bufs = []
        self.peers = {}
        self.handles = {}
        for _ in range(2):
            raw = symm_mem.empty(2 * n_amps, dtype=torch.float64, device=device)
            hdl = symm_mem.rendezvous(raw, group)
            t = torch.view_as_complex(raw.view(-1, 2))
            self.bufs.append(t)
            self.peers[t.data_ptr()] = [int(p) for p in hdl.buffer_ptrs]
            self.handles[t.data_ptr()] = hdl
        self._raw = None

    def buffers(self):
        return self.bufs[0], self.bufs[1]

    def permute_exchange(self, executor, state, spare, n_loc, perm):
        import ctypes as C
        from ._lib import check
        eng = executor.any_engine()
        peers = self.peers[spare.data_ptr()]
        arr = (C.c_void_p * self.world)(*peers)
        p = np.ascontiguousarray(perm, dtype=np.int32)
        hdl = self.handles[spare.data_ptr()]
        hdl.barrier(channel=0)      # every rank has finished whatever still touched its destination buffer
        check(eng.lib.qsv_permute_exchange(eng._ctx, C.c_void_p(state.data_ptr()), arr, self.g, self.rank, n_loc,
                                           p.ctypes.data_as(C.c_void_p), eng._stream()), eng._ctx)
        hdl.barrier(channel=1)      # every rank's stores have landed
        self.bytes_sent += state.numel() * state.element_size() * (self.world - 1) // self.world
        return spare, state


def _segment_exchange(self, executor, key, state, spare):
    """A whole local segment whose LAST sweep is the exchange: that sweep's store writes every amplitude into the rank
    that owns it afterwards, over peer memory (``qsv_run_exchange``) -- the segment's last gates and the all-to-all are
    ONE kernel, and no separate pass over the shard is spent on the exchange.  The plan was built so that every earlier
    sweep works in place: the destination (this rank's ``spare``) is untouched until the peers' stores arrive."""
    eng, _ = executor.engines[key]
    peers = self.peers[spare.data_ptr()]
    hdl = self.handles[spare.data_ptr()]
    hdl.barrier(channel=0)      # every rank has finished whatever still touched its destination buffer
    eng.run_exchange(state, spare, peers, self.g, self.rank)
    hdl.barrier(channel=1)      # every rank's stores have landed
    self.bytes_sent += state.numel() * state.element_size() * (self.world - 1) // self.world
    return spare, state


SymmExchanger.segment_exchange = _segment_exchange


class GpuExecutor:
    """Runs local segments on the CUDA engine; one libqsv context (with its fused plan) per segment."""

    def __init__(self, device, fuse=True, tile_bits=0):
        self.device = device
        self.fuse = fuse
        self.tile_bits = tile_bits
        self.engines = {}
        self.launches = 0
        self.fused_sweeps_per_apply = 0   # fused compute sweeps of all prepared segments together (one apply runs each once)
        self.sweeps_of = {}               # segment key -> fused sweeps of its plan
        self.copy_back = {}               # segment key -> True: this rank copies the result back to the buffer it started in

    def prepare(self, key, recs, pool, n_loc, final_pos=None, carry_exchange=False):
        """``final_pos`` (logical bit -> physical position): the layout the segment must end in -- the parking permutation
        of the exchange that follows rides on the segment's last sweep.  ``carry_exchange``: that sweep can store into the
        peers' buffers.  Returns the parity of the plan's out-of-place sweeps (1 = the result changes buffers)."""
        from .engine import Engine
        eng = Engine(self.device, np.complex128)
        eng.upload_records(recs, pool, n_loc, 0)
        fused = False
        flips = 0
        if self.fuse and final_pos is not None:
            self.sweeps_of[key] = int(eng.fuse_to_layout(final_pos, carry_exchange, self.tile_bits))
            self.fused_sweeps_per_apply += self.sweeps_of[key]
            fused = True
            n_sw, n_oop = eng.plan_info()
            assert n_oop == 1, "a segment that ends in a requested layout stores out of place exactly once, in its last sweep"
            flips = 1
        elif self.fuse and len(recs):
            # (no per-gate downgrade here: a segment the fusion pass refuses is a planning error at these sizes)
            self.sweeps_of[key] = int(eng.fuse(self.tile_bits))
            self.fused_sweeps_per_apply += self.sweeps_of[key]
            fused = True
            flips = eng.plan_info()[1] & 1
        self.engines[key] = (eng, fused)
        return flips

    def agree_on_buffers(self, keys, flips, all_reduce_min):
        """A segment's result lands in the other buffer when its plan has an odd number of out-of-place sweeps, and the
        plans are per rank: a rank whose global control is not satisfied drops gates the others keep (their SWAP
        relabels with them).  The fused exchange stores into EVERY rank's spare buffer, so all ranks must agree on
        which one that is.  ``all_reduce_min`` reduces an int list over the ranks; where the ranks disagree, the odd
        ones copy their result back (one extra pass on those ranks only) so that every rank ends where it started."""
        agreed = all_reduce_min(list(flips))
        for key, mine, low in zip(keys, flips, agreed):
            self.copy_back[key] = bool(mine and not low)

    def run_segment(self, key, state, spare):
        eng, fused = self.engines[key]
        res = eng.run(state, spare if fused else None)
        if res is not state and self.copy_back.get(key):
            state.copy_(res)
            return state, spare
        return (res, state) if res is not state else (state, spare)

    def permute(self, state, spare, n_loc, perm):
        import ctypes as C
        from ._lib import check
        eng = self.any_engine()
        p = np.ascontiguousarray(perm, dtype=np.int32)
        check(eng.lib.qsv_permute_bits(eng._ctx, C.c_void_p(state.data_ptr()), C.c_void_p(spare.data_ptr()), n_loc,
                                       p.ctypes.data_as(C.c_void_p), eng._stream()), eng._ctx)
        return spare, state

    def any_engine(self):
        if not self.engines:
            from .engine import Engine
            self.engines["_perm"] = (Engine(self.device, np.complex128), False)
        return next(iter(self.engines.values()))[0]

    def total_stats(self):
        tot = {}
        for eng, _ in self.engines.values():
            for k, v in eng.stats().items():
                tot[k] = tot.get(k, 0) + v
        return tot

    def reset_stats(self):
        for eng, _ in self.engines.values():
            eng.reset_stats()


class ShardedCircuit:
    """A planned, rank-localised circuit ready to be applied repeatedly to this rank's shard.

    ``restore=False`` leaves the result in the layout the last segment ended in and reports it as
    ``final_where`` (logical bit -> physical position, global positions last): QFT then needs two exchange steps
    instead of three -- its final SWAP layer becomes output ordering."""

    def __init__(self, recs, pool, n_bits, rank, world, executor, exchanger, restore=True, all_reduce_min=None, merge_exchanges=True):
        g = world.bit_length() - 1
        assert 1 << g == world, "world size must be a power of two"
        self.plan = plan_sharded(recs, n_bits, g, restore=restore)
        self.final_where = list(self.plan.final_where)
        self.rank, self.world = rank, world
        self.executor, self.exchanger = executor, exchanger
        self.n_loc = self.plan.n_loc
        self.n_gates = len(recs)
        li = 0
        self._keys = []
        flips = []
        # A local segment followed by (park permutation,) exchange: the permutation rides on the segment's last sweep, and
        # with the peer-memory engine that sweep's store IS the exchange.  self._merged[k] = number of plan steps the
        # segment at step index k swallows (0: nothing merged).
        self._merged = {}
        can_layout = merge_exchanges and hasattr(executor, "agree_on_buffers") and getattr(executor, "fuse", False)
        steps = self.plan.steps
        for k, s in enumerate(steps):
            if s[0] == "local":
                lrecs, lpool = localize_records(recs, pool, s[1], self.plan.where_at[li], self.n_loc, rank)
                final_pos = None
                if can_layout and k + 1 < len(steps):
                    if steps[k + 1][0] == "permute" and k + 2 < len(steps) and steps[k + 2][0] == "xall":
                        final_pos, swallowed = list(steps[k + 1][1]), (2 if hasattr(exchanger, "segment_exchange") else 1)
                    elif steps[k + 1][0] == "xall" and hasattr(exchanger, "segment_exchange"):
                        final_pos, swallowed = list(range(self.n_loc)), 1
                if final_pos is not None:
                    carry = hasattr(exchanger, "segment_exchange")
                    flips.append(int(executor.prepare(li, lrecs, lpool, self.n_loc, final_pos=final_pos, carry_exchange=carry)))
                    self._merged[k] = (swallowed, carry)
                else:
                    flips.append(int(executor.prepare(li, lrecs, lpool, self.n_loc) or 0))
                self._keys.append(li)
                li += 1
        if all_reduce_min is not None and hasattr(executor, "agree_on_buffers"):
            executor.agree_on_buffers(self._keys, flips, all_reduce_min)
        # sweeps whose time shows up under "local" in a per-step split (the exchange-carrying segments are timed as a whole)
        carrying = {self._keys[i] for i, k_ in enumerate(k2 for k2, s2 in enumerate(steps) if s2[0] == "local")
                    if self._merged.get(k_, (0, False))[1]}
        so = getattr(executor, "sweeps_of", {})
        self.plain_local_sweeps = sum(v for key, v in so.items() if key not in carrying)
        self.carrying_sweeps = sum(v for key, v in so.items() if key in carrying)

    def apply(self, state, spare, timers=None):
        """Returns (state, spare) -- the buffers may have traded places.  ``timers`` (a list) collects one
        ``(step kind, start event, end event)`` per plan step, recorded on the current CUDA stream."""
        li = 0
        fused_x = hasattr(self.exchanger, "permute_exchange")
        steps = self.plan.steps
        k = 0
        while k < len(steps):
            s = steps[k]
            merged = self._merged.get(k) if s[0] == "local" else None
            k += 1
            if merged:
                swallowed, carry = merged
                if timers is not None:
                    import torch
                    e0 = torch.cuda.Event(enable_timing=True)
                    e0.record()
                if carry:      # compute + exchange in one kernel (peer stores)
                    state, spare = self.exchanger.segment_exchange(self.executor, self._keys[li], state, spare)
                else:          # the parking permutation rode on the segment's last sweep; the exchange itself follows
                    state, spare = self.executor.run_segment(self._keys[li], state, spare)
                li += 1
                k += swallowed
                if timers is not None:
                    e1 = torch.cuda.Event(enable_timing=True)
                    e1.record()
                    timers.append(("local+xall" if carry else "local", e0, e1))
                continue
            if fused_x and s[0] in ("permute", "xall"):
                # (local permutation, exchange) or a bare exchange -> one kernel storing into the peers' buffers
                if s[0] == "permute" and k < len(steps) and steps[k][0] == "xall":
                    perm = s[1]
                    k += 1
                    s = ("pxall", perm)
                elif s[0] == "xall":
                    s = ("pxall", list(range(self.n_loc)))
            if timers is not None:
                import torch
                e0 = torch.cuda.Event(enable_timing=True)
                e0.record()
            if s[0] == "pxall":
                state, spare = self.exchanger.permute_exchange(self.executor, state, spare, self.n_loc, s[1])
            elif s[0] == "local":
                state, spare = self.executor.run_segment(self._keys[li], state, spare)
                li += 1
            elif s[0] == "exchange":
                state = self.exchanger.exchange(state, spare, s[1])
            elif s[0] == "xall":
                state, spare = self.exchanger.exchange_all(state, spare)
            else:
                state, spare = self.executor.permute(state, spare, self.n_loc, s[1])
            if timers is not None:
                e1 = torch.cuda.Event(enable_timing=True)
                e1.record()
                timers.append(("xall" if s[0] == "pxall" else s[0], e0, e1))
        return state, spare


# ---------------------------------------------------------------------------------------------
# benchmark entry (bench.py --gpus N under torchrun)
# ---------------------------------------------------------------------------------------------
def _run_sharded(B, args, n, steps, warmup, e2e_steps, rank, world, local_rank):
    """One sharded workload (QFT-n over `world` GPUs): timed steps, per-step-kind split, self-check, e2e.  Every rank
    runs it; the returned record is complete on rank 0."""
    import torch
    import torch.distributed as dist
    g = world.bit_length() - 1
    prog, recs, pool = B.qft_records(n)
    n_loc = n - g
    amps = 1 << n_loc
    dev = f"cuda:{local_rank}"
    # exchange engine: the fused peer-memory kernel when torch symmetric memory works here, else NCCL all-to-all
    exch = None
    exch_kind = "nccl all_to_all"
    # measured in round 2 (profiles/r2e_*, r2g_*): with the exchange riding on a compute sweep the peer-memory engine beats
    # NCCL at every world size on QFT-30 (25.6 / 16.0 / 9.3 ms against 38.8 / 20.1 / 11.2 on 2 / 4 / 8 GPUs)
    want = getattr(args, "exchange", "auto")
    auto_fused = world <= 4 or amps <= int(os.environ.get("QSV_FUSED_MAX_AMPS", str(1 << 62)))
    if want == "fused" or (want == "auto" and auto_fused):
        try:
            exch = SymmExchanger(rank, world, amps, torch.device(dev))
            exch_kind = "fused permute+exchange kernel over peer memory (torch symmetric memory)"
        except Exception as e:  # noqa: BLE001
            if args.exchange == "fused":
                raise
            if rank == 0:
                print(f"[bench] symmetric memory unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
            exch = None
    ok = torch.tensor([1 if exch is not None else 0], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if int(ok.item()) == 0:
        exch = None
        exch_kind = "nccl all_to_all"
    if exch is not None:
        state, spare = exch.buffers()
    else:
        state = torch.empty(amps, dtype=torch.complex128, device=dev)
        spare = torch.empty(amps, dtype=torch.complex128, device=dev)
    gen = torch.Generator(device=dev)
    gen.manual_seed(42 + rank)
    view = torch.view_as_real(state)
    chunk = 1 << 26
    for s in range(0, amps, chunk):
        view[s:s + chunk].normal_(generator=gen)
    nrm = torch.linalg.vector_norm(state) ** 2
    dist.all_reduce(nrm)
    state.div_(torch.sqrt(nrm))
    execu = GpuExecutor(local_rank, fuse=not args.no_fuse, tile_bits=args.tile_bits)
    if exch is None:
        exch = TorchExchanger(rank, world)

    def all_reduce_min(values):
        t = torch.tensor(values if values else [0], device=dev, dtype=torch.int32)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return [int(v) for v in t.tolist()][:len(values)]

    restore = bool(getattr(args, "restore_layout", False))
    circ = ShardedCircuit(recs, pool, n, rank, world, execu, exch, restore=restore, all_reduce_min=all_reduce_min,
                          merge_exchanges=not getattr(args, "no_merge", False))
    for _ in range(warmup):
        state, spare = circ.apply(state, spare)
    torch.cuda.synchronize()
    dist.barrier()
    execu.reset_stats()
    exch.bytes_sent = 0
    sampler = B.ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        state, spare = circ.apply(state, spare)
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clocks = sampler.stop() if sampler else None
    st = execu.total_stats()
    tot = torch.tensor([st.get("amp_bytes", 0), st.get("kernel_launches", 0), exch.bytes_sent], device=dev, dtype=torch.float64)
    dist.all_reduce(tot)
    # one more, untimed, step with CUDA events around every plan step: where the time goes, and the NVLink rate
    timers = []
    for eng_, fused_ in execu.engines.values():
        if fused_:
            eng_.set_profiling(True)
    dist.barrier()
    state, spare = circ.apply(state, spare, timers)
    torch.cuda.synchronize()
    # this rank's fused sweeps one by one, per local segment in plan order (the last sweep of a carrying segment is the exchange)
    sweeps_ms = {}
    for key_, (eng_, fused_) in execu.engines.items():
        if fused_:
            sweeps_ms[str(key_)] = [round(v, 3) for v in eng_.sweep_profile()]
            eng_.set_profiling(False)
    split = {}
    for kind, a, b in timers:
        split[kind] = split.get(kind, 0.0) + a.elapsed_time(b)
    sp_t = torch.tensor([split.get("local", 0.0), split.get("permute", 0.0), split.get("xall", 0.0) + split.get("exchange", 0.0),
                         split.get("local+xall", 0.0)], device=dev, dtype=torch.float64)
    dist.all_reduce(sp_t, op=dist.ReduceOp.MAX)
    # ---- correctness at this size, outside the timed region: QFT of a basis state against the closed form on every
    # amplitude of every shard (the layout the run ends in is read through the plan's bit-position table), plus the norm
    check = None
    if not getattr(args, "no_check", False) and B.workload_kind() == "qft":
        x0 = B.CHECK_X & ((1 << n) - 1)
        state.zero_()
        if (x0 >> n_loc) == rank:
            state[x0 & (amps - 1)] = 1.0
        dist.barrier()
        state, spare = circ.apply(state, spare)
        torch.cuda.synchronize()
        err = B.qft_basis_error(state, n, x0, base_index=rank << n_loc, where=circ.final_where)
        nrm2 = torch.linalg.vector_norm(state) ** 2
        both = torch.tensor([err], device=dev, dtype=torch.float64)
        dist.all_reduce(both, op=dist.ReduceOp.MAX)
        dist.all_reduce(nrm2)
        err = float(both.item())
        check = {"tolerance": 1e-10, "check": f"QFT|x0={x0:#x}> against the closed form on all 2^{n} amplitudes (every rank its shard, max over ranks); norm",
                 "max_abs_err": err, "max_rel_err": err * 2.0 ** (n / 2), "norm_after": float(nrm2.item()),
                 "layout": "canonical" if circ.final_where == list(range(n)) else "permuted: read through the plan's bit-position table",
                 "ok": bool(err <= 1e-10 and err * 2.0 ** (n / 2) <= 1e-9 and abs(float(nrm2.item()) - 1) <= 1e-9)}
    # ---- end to end through host buffers: every rank uploads its shard from pinned memory, applies the circuit and
    # downloads the result, every step.  Steps are independent jobs: the upload of step i+1 and the download of step
    # i-1 run on their own streams (staging buffers on the device) under the compute of step i; max over ranks
    e2e = None
    if e2e_steps > 0:
        host_in = torch.empty(amps, dtype=torch.complex128, pin_memory=True)
        host_out = torch.empty(amps, dtype=torch.complex128, pin_memory=True)
        host_in.copy_(state)
        stage_in = torch.empty(amps, dtype=torch.complex128, device=dev)
        stage_out = torch.empty(amps, dtype=torch.complex128, device=dev)
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        main = torch.cuda.current_stream()
        piece = (256 << 20) // 16     # host copies in 256 MiB pieces: measured faster with both directions busy (profiles/r2x_pcie_probe.json)

        def copy_in_pieces(dst_t, src_t):
            for o in range(0, amps, piece):
                dst_t[o:o + piece].copy_(src_t[o:o + piece], non_blocking=True)

        torch.cuda.synchronize()
        dist.barrier()

        def e2e_run(count):
            nonlocal state, spare
            ev_in = ev_taken = ev_out = None
            for i in range(count + 1):
                if i > 0:          # step i-1: its upload has landed in the staging buffer -> state
                    main.wait_event(ev_in)
                    state.copy_(stage_in, non_blocking=True)
                    ev_taken = torch.cuda.Event()
                    ev_taken.record(main)
                if i < count:      # upload of step i, under the compute of step i-1
                    with torch.cuda.stream(s_in):
                        if ev_taken is not None:
                            s_in.wait_event(ev_taken)
                        copy_in_pieces(stage_in, host_in)
                        ev_in = torch.cuda.Event()
                        ev_in.record(s_in)
                if i > 0:
                    state, spare = circ.apply(state, spare)
                    if ev_out is not None:
                        main.wait_event(ev_out)                # the previous download has left the staging buffer
                    stage_out.copy_(state, non_blocking=True)
                    ev_res = torch.cuda.Event()
                    ev_res.record(main)
                    with torch.cuda.stream(s_out):
                        s_out.wait_event(ev_res)
                        copy_in_pieces(host_out, stage_out)
                        ev_out = torch.cuda.Event()
                        ev_out.record(s_out)
            main.wait_stream(s_out)

        e2e_run(1)
        torch.cuda.synchronize()
        dist.barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        s_in.wait_event(a0)
        e2e_run(e2e_steps)
        a1.record()
        torch.cuda.synchronize()
        tt = torch.tensor([a0.elapsed_time(a1) / e2e_steps], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms = float(tt.item())
        e2e = {"value": circ.n_gates / (e2e_ms / 1e3), "unit": "gates/s", "h2d_bytes_per_step": amps * 16 * world,
               "d2h_bytes_per_step": amps * 16 * world, "ms_per_step": e2e_ms, "steps": e2e_steps,
               "note": "per rank: upload of step i+1 and download of step i-1 on their own streams under the compute of step i"}
        del host_in, host_out, stage_in, stage_out
    rec = None
    if rank == 0:
        sec = ms / 1e3
        peak, peak_src = B.measured_peak()
        amp_gbps = float(tot[0]) / sec / 1e9
        n_ex = circ.plan.n_exchanges
        # the local compute sweeps on their own (rank 0's plan, the slowest rank's time): what the north-star target
        # "fraction of HBM bandwidth per local-gate sweep" is about -- the aggregate figure below includes exchange time
        local_roof = None
        n_loc_sweeps = int(getattr(circ, "plain_local_sweeps", 0))
        if n_loc_sweeps and float(sp_t[0]) > 0:
            ach = n_loc_sweeps * 2 * amps * 16 / (float(sp_t[0]) * 1e-3) / 1e9
            local_roof = {"bound": "hbm", "kernel": "k_sweep8 / k_sweep, local compute sweeps of one step", "sweeps_per_step": n_loc_sweeps,
                          "ms_per_step": float(sp_t[0]), "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                          "frac_of_nominal_8TBps": ach / 8000.0}
        # NVLink rate: from the exchanges that run as steps of their own (a segment whose last sweep IS the exchange also
        # computes, so its time says nothing about the wire); bytes per exchange per rank = shard x (world-1)/world
        n_carried = sum(1 for v in circ._merged.values() if v[1])
        per_x = amps * 16 * (world - 1) / world
        x_ms = float(sp_t[2])
        sent = per_x * (n_ex - n_carried)
        rec = {"metric": "gates_per_sec", "value": steps * circ.n_gates / sec, "unit": "gates/s", "n_gpus": world,
               "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True,
               "scaling": "strong", "vs_baseline": None, "dtype": "c128", "data": "synthetic",
               "config": B.workload_config(n, circ.n_gates),
               "impl_config": {"sharding": f"top {g} qubits over {world} GPUs", "exchanges_per_step": n_ex, "exchange": exch_kind,
                               "local_amplitudes": amps, "restore_layout": restore, "final_bit_positions": circ.final_where,
                               "exchange_rides_on_last_sweep": bool(circ._merged),
                               "l2": "local shard far larger than the 126 MB L2" if amps * 16 > (1 << 28) else "local shard fits L2"},
               "amp_GBps": amp_gbps, "hbm_frac_of_measured": amp_gbps / (peak * world),
               "hbm_frac_of_nominal_8TBps": amp_gbps / (8000.0 * world),
               "roofline": {"bound": "hbm", "kernel": "fused tile sweeps, aggregate over ranks incl. exchange time",
                            "achieved": amp_gbps / world, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                            "frac": amp_gbps / world / peak, "traffic": None, "traffic_source": None},
               "local_sweep_roofline": local_roof,
               "nvlink": {"bytes_sent_per_rank_per_step": per_x * n_ex, "exchanges_per_step": n_ex, "exchanges_carried_by_a_compute_sweep": n_carried,
                          "standalone_exchange_bytes_per_rank": sent, "exchange_ms_per_step": x_ms,
                          "achieved_GBps_per_direction_per_gpu": sent / (x_ms * 1e-3) / 1e9 if x_ms > 0 else None,
                          "peak_GBps": 770.0, "peak_source": "measured peer copy per direction (B200_PROFILING.md); nominal 900",
                          "frac": (sent / (x_ms * 1e-3) / 1e9 / 770.0) if x_ms > 0 else None},
               "ms_split_one_step": {"local_sweeps": float(sp_t[0]), "local_permutes": float(sp_t[1]), "exchanges": float(sp_t[2]),
                                     "segments_whose_last_sweep_is_the_exchange": float(sp_t[3]),
                                     "sweeps_in_those_segments": int(getattr(circ, "carrying_sweeps", 0)),
                                     "rank0_sweep_ms_by_segment": sweeps_ms},
               "cpu_baseline": None, "e2e": e2e, "gpu_launches": int(tot[1]), "clocks": clocks,
               "check": check, "max_abs_err": check["max_abs_err"] if check else None}
    del state, spare, circ, execu, exch
    torch.cuda.empty_cache()
    return rec, (check is None or check["ok"])


def bench_main(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1:
        raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench as B
    if B.workload_kind() != "qft":
        raise SystemExit("the sharded bench runs the QFT workload (--circuit qft)")
    n = args.qubits or 30
    line, ok = _run_sharded(B, args, n, args.steps, args.warmup, args.e2e_steps, rank, world, local_rank)
    # BASELINE config 5 rides along on 8 GPUs: QFT-34 complex128 (256 GiB state, 32 GiB per GPU) on the same line
    if world == 8 and not args.qubits and not getattr(args, "no_qft34", False):
        dist.barrier()
        sub, ok34 = _run_sharded(B, args, 34, min(args.steps, 3), 1, 0, rank, world, local_rank)
        ok = ok and ok34
        if rank == 0:
            keep = ("value", "unit", "steps", "warmup", "ms_per_step", "config", "impl_config", "amp_GBps", "hbm_frac_of_measured",
                    "local_sweep_roofline", "nvlink", "ms_split_one_step", "check", "max_abs_err", "gpu_launches")
            line["qft34"] = {k: sub[k] for k in keep}
    flag = torch.tensor([1 if ok else 0], device=f"cuda:{local_rank}")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps(line))
        if not int(flag.item()):
            print(f"[bench] SELF-CHECK FAILED: {line.get('check')} {line.get('qft34', {}).get('check')}", file=sys.stderr)
    dist.destroy_process_group()
    if not int(flag.item()):
        sys.exit(3)
