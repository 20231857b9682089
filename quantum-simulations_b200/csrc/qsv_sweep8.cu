// Radix-8 image of the fused tile sweep (IMG_FFT8): FFT-shaped launches with three register bits per stage, tiles moved
// in AND out by the TMA engine.
//
// Why a second image.  The radix-16 FFT image (k_sweep<.., IMG_FFT>, qsv_sweep.cu) holds 16 complex128 amplitudes per
// thread, needs 168 registers and therefore runs 3 CTAs x 4 warps = 12 warps per SM; ncu shows it waiting on fixed-latency
// dependencies with too few warps to cover them (profiles/r2a_ncu_full_fft_qft30.csv: 17 % of the warp slots occupied,
// 1.45 "wait" + 1.0 "barrier" stall cycles per issued instruction).  Here a thread holds 8 amplitudes, a CTA has 256
// threads under an 85-register cap, and three CTAs give 24 warps per SM on the same 32 KB double-buffered tiles.  A sweep
// with up to six butterfly targets is two stages either way (4 + 2 there, 3 + 3 here).
//
// Tile movement.  ONE thread issues cp.async.bulk.tensor (SASS UTMALDG) with a tensor map that describes the tile's bit
// runs as a box of up to five dimensions (TmaTile, qsv_sweep_ops.cuh); an mbarrier counts the bytes in; no thread computes
// a load address or spends an LSU slot on the copy.  The box lands in the engine's 128-byte swizzle, which the stage
// accesses follow (swz128): conflict-free 128-bit LDS/STS whenever a quarter-warp's three lowest thread bits fall on
// distinct XOR classes of the slot number -- true for the planner's usual stage choices (register bits all above tile bit
// 5, or exactly tile bits 0-2 / 3-5).  Every tile leaves through a TMA store (UTMASTG) issued by the same thread: measured
// on B200, QFT-30 sweeps took 7.2-7.3 ms with 128-bit register stores and 6.7 ms with the engine's store (profiles/
// r2_ab_sweep8.txt).  A bit-permuting sweep (SWAP relabels folded into the store) has a second tensor map over the OUTPUT
// positions of the tile's bits; its last stage writes the tile back in that order (one more barrier) and the engine does the
// scatter.
//
// Per tile: wait for the tile's mbarrier, ONE CTA barrier (which also publishes the previous tile, fenced, to the storing
// thread and the scalars computed a tile ahead), stage, barrier, stage.  The elected thread then stores the previous tile,
// waits until the engine has read it, and loads the tile after next into the same buffer.
//
// The per-stage arithmetic (fstage8_compute / fstage8_put) is in qsv_sweep_ops.cuh, shared with the host check.
#include <cuda.h>
#include <stdlib.h>

#include <algorithm>

#include "qsv_common.cuh"
#include "qsv_sweep_ops.cuh"

struct TmaCoords {  // device copy of a TmaTile's coordinate recipe
    int rank, n_inst;
    uint8_t shift[5], bits[5];
    uint32_t inst_amps;
    uint64_t inst_goff[TMA_MAX_INST];
};

struct Sweep8Params {
    SweepDesc d;
    SwOp ops[MAX_OPS8];
    TmaCoords in, out;
    int permuting;            // the last stage re-orders the tile for a bit-permuting store
    SegMap loc_perm;          // input-local index -> output-local index (bit permutation)
    uint16_t boff_out[RB8];   // output-order slot offset of the last stage's register bit q
    int l2_hint;              // A/B aid: 1 = evict_first on loads, 2 = on stores, 3 = both
    int generic;              // the launch is not FFT-shaped: stages are DC_STAGE ops followed by generic ops
    int l2_prefetch;          // tiles ahead of the load proper that are prefetched into L2 (0 = off)
    int mid_layout;           // two-stage permuting launch through the intermediate swz() layout (Sweep8Extra)
    SegMap tmap_last;
    uint16_t boff_mid_w[RB8], boff_last_r[RB8];
    // d.xg > 0 (the store is the exchange, see SweepDesc): the tile, complete in OUTPUT order in shared memory, is not
    // handed to the engine but stored by the threads, each amplitude into the rank that owns it afterwards.
    uint64_t x_bit_off[KMAX8];  // output index offset of bit u of the slot number (= the u-th lowest output tile bit)
};

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t grid_x8() {
    uint32_t gx;
    asm volatile("mov.u32 %0, %%nctaid.x;" : "=r"(gx));
    return gx;
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "W8_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra W8_DONE;\n"
        "bra W8_WAIT;\n"
        "W8_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// one box: global -> shared, completion counted on the mbarrier.  (No multicast: every CTA fetches its own tile.)
// `policy` != 0: an L2 cache-policy word (createpolicy) rides along -- A/B aid, see sweep8_launch.
__device__ __forceinline__ void tma_load(const CUtensorMap* tm, void* smem_dst, uint64_t* bar, int rank, int c1, int c2, int c3, int c4,
                                         uint64_t policy = 0) {
    const uint32_t d = smem_u32(smem_dst), b = smem_u32(bar);
    const uint64_t m = reinterpret_cast<uint64_t>(tm);
    if (policy && rank == 3) {
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(d),
            "l"(m), "r"(b), "r"(0), "r"(c1), "r"(c2), "l"(policy)
            : "memory");
        return;
    }
    switch (rank) {
        case 2:
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(d),
                         "l"(m), "r"(b), "r"(0), "r"(c1)
                         : "memory");
            break;
        case 3:
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(d),
                         "l"(m), "r"(b), "r"(0), "r"(c1), "r"(c2)
                         : "memory");
            break;
        case 4:
            asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
                             d),
                         "l"(m), "r"(b), "r"(0), "r"(c1), "r"(c2), "r"(c3)
                         : "memory");
            break;
        default:
            asm volatile(
                "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(d),
                "l"(m), "r"(b), "r"(0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
                : "memory");
            break;
    }
}
// one box: global -> L2 only (no shared-memory buffer is needed yet: issued a tile earlier than the load proper)
__device__ __forceinline__ void tma_prefetch_l2(const CUtensorMap* tm, int rank, int c1, int c2, int c3, int c4) {
    const uint64_t m = reinterpret_cast<uint64_t>(tm);
    switch (rank) {
        case 2: asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(m), "r"(0), "r"(c1) : "memory"); break;
        case 3: asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(m), "r"(0), "r"(c1), "r"(c2) : "memory"); break;
        case 4:
            asm volatile("cp.async.bulk.prefetch.tensor.4d.L2.global.tile [%0, {%1, %2, %3, %4}];" ::"l"(m), "r"(0), "r"(c1), "r"(c2), "r"(c3) : "memory");
            break;
        default:
            asm volatile("cp.async.bulk.prefetch.tensor.5d.L2.global.tile [%0, {%1, %2, %3, %4, %5}];" ::"l"(m), "r"(0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
                         : "memory");
            break;
    }
}
// one box: shared -> global (bulk async group)
__device__ __forceinline__ void tma_store(const CUtensorMap* tm, const void* smem_src, int rank, int c1, int c2, int c3, int c4, uint64_t policy = 0) {
    const uint32_t s = smem_u32(smem_src);
    const uint64_t m = reinterpret_cast<uint64_t>(tm);
    if (policy && rank == 3) {
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group.L2::cache_hint [%0, {%2, %3, %4}], [%1], %5;" ::"l"(m), "r"(s), "r"(0),
                     "r"(c1), "r"(c2), "l"(policy)
                     : "memory");
        return;
    }
    switch (rank) {
        case 2:
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(m), "r"(s), "r"(0), "r"(c1) : "memory");
            break;
        case 3:
            asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(m), "r"(s), "r"(0), "r"(c1), "r"(c2)
                         : "memory");
            break;
        case 4:
            asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(m), "r"(s), "r"(0), "r"(c1),
                         "r"(c2), "r"(c3)
                         : "memory");
            break;
        default:
            asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::"l"(m), "r"(s), "r"(0),
                         "r"(c1), "r"(c2), "r"(c3), "r"(c4)
                         : "memory");
            break;
    }
}

}  // namespace

__global__ void __launch_bounds__(256, 3) k_sweep8(const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_out,
                                                   const double* __restrict__ aux, const __grid_constant__ Sweep8Params P) {
    typedef double2 C;
    typedef double R;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const SweepDesc& D = P.d;
    const uint32_t tile_bytes = 16u << D.k;
    C* const buf0 = reinterpret_cast<C*>(smem_raw);
    double2* const smE0 = reinterpret_cast<double2*>(smem_raw + 2 * tile_bytes);  // [2][MAX_GSLOT8] per-tile scalars by tile parity
    double2* const smT = smE0 + 2 * MAX_GSLOT8;
    uint32_t* const park = reinterpret_cast<uint32_t*>(smT + MAX_TAB);  // [PARK8_WORDS][blockDim]: slot of register 0 per stage, read / write side
    uint64_t* const mbar = reinterpret_cast<uint64_t*>(park + PARK8_WORDS * blockDim.x);
    const double2* aux2 = reinterpret_cast<const double2*>(aux);

    const uint32_t tid = threadIdx.x;
    const uint32_t n_tiles = 1u << (D.n_bits - D.k);
    const int n_ops = D.n_ops;
    // tile-independent per-thread constants: where the thread's register 0 sits in each stage (and in the output order)
    auto slot0 = [&](int i) {
        const OpView o{&P.ops[i]};
        return expand_rb8(tid, o.rpos(0), o.rpos(1), o.rpos(2));
    };
    for (int i = 0; i < n_ops && i < PARK8_OPS && !P.generic; ++i) park[i * blockDim.x + tid] = swz128(slot0(i));
    if (P.mid_layout) {  // (two stages) first stage writes swz(), last stage reads swz() with its lanes on the output order's low bits
        const uint32_t bl1 = seg_apply(tid, P.tmap_last);
        park[3 * blockDim.x + tid] = swz(slot0(0));
        const OpView o1{&P.ops[1]};  // (the phase tables of the stage are indexed by the thread number of the standard assignment)
        park[1 * blockDim.x + tid] = swz(bl1) | (compress_rb8(bl1, o1.rpos(0), o1.rpos(1), o1.rpos(2)) << 16);
        park[2 * blockDim.x + tid] = swz128(seg_apply(bl1, P.loc_perm));
    } else if (P.permuting) {
        park[2 * blockDim.x + tid] = swz128(seg_apply(slot0(n_ops - 1), P.loc_perm));
    }
    for (uint32_t i = tid; i < D.tab_len; i += blockDim.x) smT[i] = aux2[D.tab_off + i];
    if (tid == 0) {
        if (smem_u32(smem_raw) & 1023u) __trap();  // the 128-byte swizzle pattern is anchored at 1 KB
        mbar_init(&mbar[0], 1);
        mbar_init(&mbar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_async_smem();
    }
    // per-tile phase scalars: product of 16-entry chunk tables over the tile id, written one tile AHEAD (parity buffers)
    auto scalars = [&](uint32_t t, double2* smE, uint32_t gs) {
        const double2* gt = aux2 + D.gtab_off + (size_t)gs * D.g_chunks * 16;
        double2 e = gt[t & 15u];
        for (int c = 1; c < D.g_chunks; ++c) e = cmul(e, gt[c * 16 + ((t >> (4 * c)) & 15u)]);
        smE[gs] = e;
    };
    if (tid < (uint32_t)D.n_gslots && blockIdx.x < n_tiles) scalars(blockIdx.x, smE0, tid);
    __syncthreads();

    // coordinates of a box: dimension d >= 1 starts at the tile-id bits of its index range (the tile bits there are zero)
    auto box_coords = [&](const TmaCoords& T, uint64_t p, int& c1, int& c2, int& c3, int& c4) {
        c1 = (int)((p >> T.shift[1]) & ((1ull << T.bits[1]) - 1ull));
        c2 = T.rank > 2 ? (int)((p >> T.shift[2]) & ((1ull << T.bits[2]) - 1ull)) : 0;
        c3 = T.rank > 3 ? (int)((p >> T.shift[3]) & ((1ull << T.bits[3]) - 1ull)) : 0;
        c4 = T.rank > 4 ? (int)((p >> T.shift[4]) & ((1ull << T.bits[4]) - 1ull)) : 0;
    };
    uint64_t pol = 0;
    if (P.l2_hint) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    const uint64_t pol_ld = (P.l2_hint & 1) ? pol : 0, pol_st = (P.l2_hint & 2) ? pol : 0;
    auto issue_load = [&](uint32_t t, int b) {  // one thread
        const uint64_t phys = seg_apply64x(t, D.in_glob);
        mbar_expect_tx(&mbar[b], tile_bytes);
        for (int e = 0; e < P.in.n_inst; ++e) {
            int c1, c2, c3, c4;
            box_coords(P.in, phys | P.in.inst_goff[e], c1, c2, c3, c4);
            tma_load(&tm_in, smem_raw + (size_t)b * tile_bytes + (size_t)e * P.in.inst_amps * sizeof(C), &mbar[b], P.in.rank, c1, c2, c3, c4, pol_ld);
        }
    };
    auto issue_prefetch = [&](uint32_t t) {  // one thread
        const uint64_t phys = seg_apply64x(t, D.in_glob);
        for (int e = 0; e < P.in.n_inst; ++e) {
            int c1, c2, c3, c4;
            box_coords(P.in, phys | P.in.inst_goff[e], c1, c2, c3, c4);
            tma_prefetch_l2(&tm_in, P.in.rank, c1, c2, c3, c4);
        }
    };
    auto issue_store = [&](uint32_t t, int b) {  // one thread; the tile in buffer b is complete and fenced
        const uint64_t phys = seg_apply64x(t, D.out_glob);
        for (int e = 0; e < P.out.n_inst; ++e) {
            int c1, c2, c3, c4;
            box_coords(P.out, phys | P.out.inst_goff[e], c1, c2, c3, c4);
            tma_store(&tm_out, smem_raw + (size_t)b * tile_bytes + (size_t)e * P.out.inst_amps * sizeof(C), P.out.rank, c1, c2, c3, c4, pol_st);
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    };

    if (tid == 0 && blockIdx.x < n_tiles) issue_load(blockIdx.x, 0);
    uint32_t it = 0, t_prev = 0;
    for (uint32_t t = blockIdx.x; t < n_tiles; t_prev = t, t += grid_x8(), ++it) {
        const int cur = it & 1;
        C* tile = buf0 + ((size_t)cur << D.k);
        const double2* smE = smE0 + cur * MAX_GSLOT8;
        mbar_wait(&mbar[cur], (it >> 1) & 1u);
        __syncthreads();  // this tile's scalars are visible; every thread has finished (and fenced) the previous tile
        const uint32_t tn = t + grid_x8();
        const bool more = tn > t && tn < n_tiles;
        if (tid == 0) {
            // the previous tile (other buffer) goes out, and as soon as the engine has read it the next one comes in
            if (it && !D.xg) {
                issue_store(t_prev, cur ^ 1);
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            if (more) issue_load(tn, cur ^ 1);
            if (P.l2_prefetch) {
                const uint64_t tp = (uint64_t)t + (uint64_t)(1 + P.l2_prefetch) * grid_x8();
                if (tp < n_tiles) issue_prefetch((uint32_t)tp);
            }
        } else if (more && tid >= blockDim.x - (uint32_t)D.n_gslots) {
            scalars(tn, smE0 + (cur ^ 1) * MAX_GSLOT8, blockDim.x - 1 - tid);
        }
        if (P.mid_layout) {
            C w[NREG8];
            const OpView op0{&P.ops[0]}, op1{&P.ops[1]};
            fstage8_compute<C, R>(op0, tile, smT, smE, tid, park[tid], op0.boff(0), op0.boff(1), op0.boff(2), w);
            __syncthreads();  // (each stage of this mode writes other threads' slots: everyone reads first)
            fstage8_put<C>(tile, park[3 * blockDim.x + tid], P.boff_mid_w[0], P.boff_mid_w[1], P.boff_mid_w[2], w);
            __syncthreads();
            const uint32_t pk = park[blockDim.x + tid];
            fstage8_compute<C, R>(op1, tile, smT, smE, pk >> 16, pk & 0xFFFFu, P.boff_last_r[0], P.boff_last_r[1], P.boff_last_r[2], w);
            __syncthreads();
            fstage8_put<C>(tile, park[2 * blockDim.x + tid], P.boff_out[0], P.boff_out[1], P.boff_out[2], w);
        } else if (P.generic) {
            // launches that are not FFT-shaped: a stage is a DC_STAGE op followed by the ops it holds in registers
            for (int i = 0; i < n_ops;) {
                const OpView sop{&P.ops[i]};
                const uint32_t bl = expand_rb8(tid, sop.st_rpos(0), sop.st_rpos(1), sop.st_rpos(2));
                const uint32_t a0 = swz128(bl), b0 = sop.st_roffsw(1), b1 = sop.st_roffsw(2), b2 = sop.st_roffsw(4);
                C w[NREG8];
                {
                    uint32_t a[NREG8];
                    slots8(a, a0, b0, b1, b2);
#pragma unroll
                    for (int j = 0; j < NREG8; ++j) w[j] = tile[a[j]];
                }
                for (++i; i < n_ops; ++i) {
                    const OpView op{&P.ops[i]};
                    const uint32_t type = op.type();
                    if (type == DC_STAGE) break;
                    const uint64_t gm = op.gmask();
                    if (gm && ((uint64_t)t & gm) != op.gval()) continue;  // control on a tile-id bit not satisfied for this tile
                    generic_op8<C, R>(w, op, type, (bl & op.lmask()) == op.lval(), smT, smE, tid);
                }
                fstage8_put<C>(tile, a0, b0, b1, b2, w);
                if (i < n_ops) __syncthreads();
            }
        } else {
            for (int i = 0; i < n_ops; ++i) {
                const OpView op{&P.ops[i]};
                const uint32_t a0 = i < PARK8_OPS ? park[i * blockDim.x + tid] : swz128(slot0(i));
                C w[NREG8];
                fstage8_compute<C, R>(op, tile, smT, smE, tid, a0, op.boff(0), op.boff(1), op.boff(2), w);
                if (i == n_ops - 1 && P.permuting) {
                    __syncthreads();  // the output order uses other threads' slots: everyone has read first
                    fstage8_put<C>(tile, park[2 * blockDim.x + tid], P.boff_out[0], P.boff_out[1], P.boff_out[2], w);
                } else {
                    fstage8_put<C>(tile, a0, op.boff(0), op.boff(1), op.boff(2), w);
                }
                if (i < n_ops - 1) __syncthreads();
            }
        }
        if (D.xg) {
            // The store IS the exchange: the top xg bits of an amplitude's output index name the rank that owns it next, and
            // this rank's id takes their place in the address there.  Slot l of the output-ordered tile is stored by thread
            // l mod blockDim: a warp writes 512 contiguous bytes of one peer per instruction.
            __syncthreads();
            const int tb = D.k - RB8;
            uint64_t x_tid_off = 0;  // where the thread's slots tid + (j << tb) go, less the j part (recomputed per tile: a
            for (int u = 0; u < tb; ++u)  // value kept across the stages would cost the other launches a register pair)
                if ((tid >> u) & 1u) x_tid_off |= P.x_bit_off[u];
            const int sh = D.n_bits - D.xg;
            const uint64_t lowmask = (1ull << sh) - 1ull, mine = (uint64_t)D.xrank << sh;
            const uint64_t gb = seg_apply64x(t, D.out_glob) | x_tid_off;
            const uint64_t o0 = P.x_bit_off[tb], o1 = P.x_bit_off[tb + 1], o2 = P.x_bit_off[tb + 2];
            const uint32_t s0 = swz128(tid);
            C w[NREG8];
#pragma unroll
            for (int j = 0; j < NREG8; ++j) w[j] = tile[s0 | ((uint32_t)j << tb)];
#pragma unroll
            for (int j = 0; j < NREG8; ++j) {
                const uint64_t y = gb | ((j & 1) ? o0 : 0ull) | ((j & 2) ? o1 : 0ull) | ((j & 4) ? o2 : 0ull);
                C* pd = reinterpret_cast<C*>(D.peer[y >> sh]);
                pd[mine | (y & lowmask)] = w[j];
            }
        }
        fence_async_smem();  // this thread's tile writes are visible to the engine once it has passed the next barrier
    }
    __syncthreads();
    if (tid == 0 && it && !D.xg) {
        issue_store(t_prev, (it - 1) & 1);
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

// ================================================================================================
// host
// ================================================================================================
typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn get_encode() {
    static encode_tiled_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<encode_tiled_fn>(p);
    }
    return fn;
}

static int encode_tile_map(qsv_ctx* ctx, const TmaTile& T, void* base, CUtensorMap* out) {
    encode_tiled_fn enc = get_encode();
    if (!enc) return qsv_fail(ctx, QSV_E_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t dims[5], strides[4];
    cuuint32_t box[5], estr[5];
    for (int d = 0; d < T.rank; ++d) {
        dims[d] = T.size[d];
        box[d] = T.box[d];
        estr[d] = 1;
        if (d) strides[d - 1] = T.stride_bytes[d];
    }
    const CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, (cuuint32_t)T.rank, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return qsv_fail(ctx, QSV_E_CUDA, "cuTensorMapEncodeTiled failed (CUresult %d, rank %d)", (int)r, T.rank);
    return QSV_OK;
}

int sweep8_prepare(qsv_ctx* ctx, int k) {
    const size_t smem = SWEEP8_SMEM(k, 1 << (k - RB8));
    cudaError_t e = cudaFuncSetAttribute(k_sweep8, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return qsv_fail(ctx, QSV_E_CUDA, "cudaFuncSetAttribute(k_sweep8, %zu B smem) failed: %s", smem, cudaGetErrorString(e));
    return QSV_OK;
}

static void fill_coords(TmaCoords& c, const TmaTile& T) {
    c.rank = T.rank;
    c.n_inst = T.n_inst;
    c.inst_amps = T.inst_amps;
    for (int d = 0; d < 5; ++d) {
        c.shift[d] = T.shift[d];
        c.bits[d] = T.bits[d];
    }
    for (int e = 0; e < TMA_MAX_INST; ++e) c.inst_goff[e] = T.inst_goff[e];
}

// Everything a launch passes by value, kept per plan step: a repeated run on the same buffers (every bench step, every
// segment of a sharded step) re-uses the 30 KB parameter block and the two encoded tensor maps instead of rebuilding them.
struct Sweep8Launch {
    Sweep8Params P;
    CUtensorMap tm_in, tm_out;
    const void* src = nullptr;
    void* dst = nullptr;
    unsigned grid = 0;
};

void sweep8_free_cache(void* cache) { delete static_cast<Sweep8Launch*>(cache); }

// xg > 0: the launch carries the exchange (peers[0 .. 2^xg) = every rank's destination buffer, xrank = this rank).
int sweep8_launch(qsv_ctx* ctx, int n, int k, const void* src, void* dst, const double* d_aux, const PlanStep& st, cudaStream_t stream, int xg,
                  int xrank, const uint64_t* peers) {
    if (ctx->dtype != QSV_C128 || !st.tma) return qsv_fail(ctx, QSV_E_INVALID, "internal: radix-8 launch without a TMA tile description");
    if (n - k > 31) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "unsupported: %d tile-id bits (max 31)", n - k);
    const int threads = 1 << (k - RB8);
    const size_t smem = SWEEP8_SMEM(k, threads);
    PlanStep& mst = const_cast<PlanStep&>(st);
    if (Sweep8Launch* L = static_cast<Sweep8Launch*>(mst.launch_cache)) {
        if (L->src != src || L->dst != dst) {  // same plan, other buffers (ping-pong, another state): only the tensor maps change
            int rc = encode_tile_map(ctx, st.tma->in, const_cast<void*>(src), &L->tm_in);
            if (!rc) rc = encode_tile_map(ctx, st.tma->out, dst, &L->tm_out);
            if (rc) return rc;
            L->src = src;
            L->dst = dst;
        }
        L->P.d.xg = xg;
        L->P.d.xrank = xrank;
        for (int i = 0; i < (xg > 0 ? 1 << xg : 0); ++i) L->P.d.peer[i] = peers[i];
        k_sweep8<<<L->grid, threads, smem, stream>>>(L->tm_in, L->tm_out, d_aux, L->P);
        QSV_CUDA(ctx, cudaGetLastError());
        return QSV_OK;
    }
    const Sweep8Extra& X = *st.tma;
    Sweep8Launch* L = new Sweep8Launch();
    Sweep8Params* P = &L->P;
    memset(P, 0, sizeof(*P));
    P->d = st.params->d;
    for (int i = 0; i < P->d.n_ops && i < MAX_OPS8; ++i) P->ops[i] = st.params->ops[i];
    fill_coords(P->in, X.in);
    fill_coords(P->out, X.out);
    P->permuting = X.permuting;
    P->loc_perm = X.loc_perm;
    static const int l2pf = getenv("QSV_L2_PREFETCH") ? atoi(getenv("QSV_L2_PREFETCH")) : 0;  // A/B aid: measured slower, see profiles/r2_ab_sweep8.txt
    P->l2_prefetch = l2pf;
    static const int l2_hint = getenv("QSV_L2_HINT") ? atoi(getenv("QSV_L2_HINT")) : 0;  // A/B aid
    P->l2_hint = l2_hint;
    P->generic = P->d.n_ops > 0 && P->ops[0].type == DC_STAGE;
    P->mid_layout = X.mid_layout;
    P->tmap_last = X.tmap_last;
    for (int q = 0; q < RB8; ++q) {
        P->boff_out[q] = X.boff_out[q];
        P->boff_mid_w[q] = X.boff_mid_w[q];
        P->boff_last_r[q] = X.boff_last_r[q];
    }
    int rc = encode_tile_map(ctx, X.in, const_cast<void*>(src), &L->tm_in);
    if (!rc) rc = encode_tile_map(ctx, X.out, dst, &L->tm_out);
    if (rc) {
        delete L;
        return rc;
    }
    L->src = src;
    L->dst = dst;
    int sm_count = 148;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, ctx->device);
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_sweep8, threads, smem);
    occ = std::max(occ, 1);
    const uint64_t n_tiles = 1ull << (n - k);
    L->grid = (unsigned)std::min<uint64_t>(n_tiles, (uint64_t)sm_count * occ);
    for (int u = 0; u < k; ++u) P->x_bit_off[u] = 1ull << X.out_pos[u];
    P->d.xg = xg;
    P->d.xrank = xrank;
    for (int i = 0; i < (xg > 0 ? 1 << xg : 0); ++i) P->d.peer[i] = peers[i];
    mst.launch_cache = L;
    k_sweep8<<<L->grid, threads, smem, stream>>>(L->tm_in, L->tm_out, d_aux, L->P);
    QSV_CUDA(ctx, cudaGetLastError());
    return QSV_OK;
}
