// Shared by qsv_sweep.cu (kernel), qsv_fuse.cu (fusion pass) and qsv_hostcheck.cu (host interpretation, test infrastructure):
// launch budgets, the op record and sweep descriptor the kernel reads from parameter space, and the per-op arithmetic,
// which is __host__ __device__ so that the SAME source runs inside k_sweep and inside the host check.
#pragma once
#include <stddef.h>

#include "qsv_common.cuh"

#define KMAX 12          // max local (tile) bits: 2^12 x 16 B = 64 KB of shared memory
#define RB 4             // register bits per stage
#define NREG 16
#define MAX_OPS 320      // ops per launch (kernel parameter space, 32 KB limit)
#define MAX_GSLOT 48     // per-tile phase scalars per launch (double-buffered in shared memory by tile parity)
#define MAX_TAB 384      // double2 entries of phase tables per launch (6 KB of shared memory)
#define PARK_WORDS 3     // 64-bit per-thread tile-copy constants parked in shared memory (registers are at the 3-CTA/SM cap)
#define SWEEP_SMEM(amp_bytes, k, threads) \
    (2 * ((size_t)(amp_bytes) << (k)) + (2 * MAX_GSLOT + MAX_TAB) * sizeof(double2) + (size_t)PARK_WORDS * (threads) * sizeof(uint64_t))
#define TSLOT_LEN 32     // one thread-phase slot = 2 chunks x 16 entries
#define MAX_SEG 12
// Radix-8 image (IMG_FFT8, qsv_sweep8.cu): 3 register bits per stage, 8 amplitudes per thread, tiles moved by TMA
#define RB8 3
#define KMAX8 11         // largest tile of the radix-8 image (2^11 amplitudes = 256 threads x 8)
#define NREG8 8
#define MAX_OPS8 300     // ops per launch of the radix-8 image (kernel parameter space; an FFT-shaped launch is one op per stage)
#define MAX_GSLOT8 16    // per-tile phase scalars per launch of the radix-8 image
#define PARK8_OPS 2      // stages whose per-thread slot constants are parked in shared memory (later ones recompute them)
#define PARK8_WORDS 4    // + the write-side constants of a permuting launch: [2] last stage (output order), [3] first stage (mid layout)
#define TMA_MAX_INST 16  // cp.async.bulk.tensor instructions per tile when the tile's bit runs need more than 5 dimensions
#define SWEEP8_SMEM(k, threads) \
    (2 * ((size_t)16 << (k)) + (2 * MAX_GSLOT8 + MAX_TAB) * sizeof(double2) + (size_t)(threads) * 4 * PARK8_WORDS + 32)

enum : uint8_t { OP_STAGE = 0, OP_DENSE = 1, OP_PERMX = 2, OP_MPHASE = 3, OP_W = 4, OP_TWID = 5, OP_HADD = 6, OP_HW = 7, OP_FSTAGE = 8 };
// Dispatch codes (SwOp::type as stored): the host resolves op kind, target register bit, real/complex
// and controlled/uncontrolled into ONE small integer so that the device decodes an op with a single
// jump-table branch.
enum : uint8_t {
    // codes 0..18 are the lean kernel's set and are dense so that its switch compiles to one jump table
    DC_STAGE = 0,
    DC_DENSE_REAL_NC = 1,    // +tgt (4 codes each); NC = no control of any kind
    DC_DENSE_CPLX_NC = 5,
    DC_HADD = 9,             // +tgt: unnormalised Hadamard butterfly (a0+a1, a0-a1); the 1/sqrt2 is folded into a later OP_TWID
    DC_W = 13,               // register-register phase table: v[j] *= W[j]
    DC_TWID = 14,            // deferred per-register-bit phases (thread / tile dependent) + thread-wide factor + real scale
    DC_HW = 15,              // +tgt: unnormalised Hadamard butterfly immediately followed by a register-register phase table
    DC_FSTAGE = 19,          // +shape (10 codes): a whole register stage of the "FFT shape" as ONE straight-line op (see
                             // fstage_math); runs only in the IMG_FFT kernel image, whose launches hold nothing else
    DC_FIRST_FULL = 29,      // codes from here on exist only in the full kernel image
    DC_DENSE_CPLX_C = 29,    // +tgt: controlled (register / thread / tile controls)
    DC_PERMX = 33,           // +tgt
    DC_MPHASE = 37,
};
#define NONE16 0xFFFFu
#define NONE32 0xFFFFFFFFu

struct __align__(16) SwOp {  // 96 bytes = 6 x 16 B, read into registers with six 128-bit constant loads
    uint8_t type, tgt, rmask, rval;  // tgt: register bit (0..3); rmask/rval: control over the register index
    uint16_t lmask, lval;            // control over the thread's local bits (register bits excluded)
    uint64_t gmask, gval;            // control over the tile id
    union {
        double m[8];                 // OP_DENSE: 2x2 (re,im) row-major; OP_MPHASE: m[0..1] = factor
        struct {                     // OP_STAGE
            uint16_t roffsw[NREG];   // swizzled shared-memory slot offset of register j
            uint8_t rpos[RB];        // local bit positions of the register bits, ascending
            uint8_t flags;
        } st;
        struct {                     // OP_W
            uint16_t w_off;          // W table offset (16 entries) in the shared tables
            uint16_t wmask;          // registers j with W[j] != 1
        } w;
        struct {                     // OP_TWID
            double scale;            // real factor applied to every register (1.0 = none)
            uint16_t tslot[RB + 1];  // thread-phase slot (offset into the shared tables) per listed bit, [RB] = thread-wide factor
            uint16_t gslot[RB + 1];  // per-tile scalar slot per listed bit, [RB] = thread-wide factor
            uint8_t abit[RB];        // register bit index of each listed slot
            uint8_t n_active;
        } tw;
        struct {                     // OP_FSTAGE: the first 33 bytes are laid out exactly like `tw`
            double scale;
            uint16_t tslot[RB + 1];
            uint16_t gslot[RB + 1];
            uint8_t abit[RB];
            uint8_t n_active;
            uint8_t hmask;           // @33: number of butterflies NH: register bits 3, 2, ... (3 - NH + 1) get one, in that order
            uint8_t flags;           // @34: 1 = per-bit deferred phases, 2 = thread-wide factor / scale, 4 = that factor is complex
            uint8_t pad_;
            uint16_t w_off[3];       // @36: 16-entry phase table applied right after the butterfly of bit 3 / 2 / 1 (always valid in
                                     //      the shapes with tables); it touches only registers with bit A set, indexed by the bits below A
            uint8_t rpos[RB];        // @42: local bit positions of the register bits, ascending (for the thread-index expansion)
            uint16_t boff[RB];       // @46: swizzled shared-memory slot offset of register bit q (register bits are NOT in position order)
        } fs;
    };
};
static_assert(sizeof(SwOp) == 96, "SwOp layout");
static_assert(offsetof(SwOp, tw) == 24 && offsetof(SwOp, fs.hmask) == 24 + 33 && offsetof(SwOp, fs.w_off) == 24 + 36 &&
                  offsetof(SwOp, fs.rpos) == 24 + 42 && offsetof(SwOp, fs.boff) == 24 + 46 && offsetof(SwOp, tw.abit) == offsetof(SwOp, fs.abit),
              "OP_FSTAGE payload offsets are hard-coded in the kernel");

struct SegMap {  // x -> sum_s ((x >> src) & ((1<<len)-1)) << dst
    uint8_t n, src[MAX_SEG], len[MAX_SEG], dst[MAX_SEG];
};

struct SegMap64 {  // same over up to 64 source bits (tile id -> global offset)
    uint8_t n, src[48], len[48], dst[48];
};

struct SweepDesc {
    int n_bits, k, n_ops, n_gslots, n_glob, tb;  // tb = log2(blockDim)
    uint32_t gtab_off;              // double2 units: per per-tile scalar slot, g_chunks x 16 entries
    int g_chunks;                   // ceil(n_glob / 4), at least 1
    SegMap in_loc, out_loc, st_enum;  // local index -> physical offset (load / store), store enumeration
    SegMap64 in_glob, out_glob;     // tile id -> physical base offset (load / store)
    // All index maps are bitwise-linear, so the tile copy splits every address into a per-thread part
    // (computed once per kernel from the maps above) and a per-slot part (below, host-computed):
    // slot i of thread tid handles local index  tid | (i << tb)  on load and  st_enum(tid | (i << tb))  on store.
    uint64_t ld_goff[NREG], st_goff[NREG];  // global offset contributed by slot i
    uint16_t ld_soff[NREG], st_soff[NREG];  // swizzled shared-memory slot contributed by slot i
    // When the first (last) stage keeps only high local bits in registers, the tile is loaded
    // (stored) straight into (from) that stage's register layout and skips one trip through
    // shared memory: slot j is then register j and the thread part is the stage's base index.
    uint8_t direct_load, direct_store;
    uint8_t first_rpos[RB], last_rpos[RB];
    uint32_t tab_off, tab_len;              // shared phase tables: aux offset / length in double2 units
    // Fused exchange (multi-GPU): when xg > 0 the store does not go to `dst` but straight into the peers' buffers over
    // NVLink -- the top xg bits of the (permuted) output index select the destination rank, and this rank's id takes
    // their place in the address there.  That is the all-to-all "global bits <-> top local bits" done by the sweep's
    // own stores: no NCCL call and no extra pass over the shard.
    int xg, xrank;
    uint64_t peer[8];                       // device address of every rank's destination buffer (peer-mapped)
};

struct SweepParams {
    SweepDesc d;
    SwOp ops[MAX_OPS];
};
static_assert(sizeof(SweepParams) < 32700, "kernel parameter space");

__host__ __device__ __forceinline__ uint32_t seg_apply(uint32_t x, const SegMap& m) {
    uint32_t r = 0;
    for (int s = 0; s < m.n; ++s) r |= ((x >> m.src[s]) & ((1u << m.len[s]) - 1u)) << m.dst[s];
    return r;
}
__host__ __device__ __forceinline__ uint64_t seg_apply64(uint32_t x, const SegMap& m) {
    uint64_t r = 0;
    for (int s = 0; s < m.n; ++s) r |= (uint64_t)((x >> m.src[s]) & ((1u << m.len[s]) - 1u)) << m.dst[s];
    return r;
}
__host__ __device__ __forceinline__ uint64_t seg_apply64x(uint64_t x, const SegMap64& m) {
    uint64_t r = 0;
    for (int s = 0; s < m.n; ++s) r |= ((x >> m.src[s]) & ((1ull << m.len[s]) - 1ull)) << m.dst[s];
    return r;
}
__host__ __device__ __forceinline__ uint32_t swz(uint32_t l) { return l ^ (((l >> 3) ^ (l >> 6) ^ (l >> 9)) & 7u); }
// The TMA engine's SWIZZLE_128B on 16-byte amplitudes: inside every 1 KB block (8 rows of 128 B) the 16-byte chunk index is
// XOR-ed with the row number -- slot of local index l in a tile buffer aligned to 1 KB.  Like swz() it is linear over GF(2).
__host__ __device__ __forceinline__ uint32_t swz128(uint32_t l) { return l ^ ((l >> 3) & 7u); }

// How one tile of a sweep is described to the TMA engine (complex128 only).  The state is viewed as a tensor of up to five
// dimensions over FLOAT64 elements: dimension 0 is (re, im) x index bits 0..2 = one 128-byte row; every further dimension d
// covers the index bits [shift[d], shift[d] + bits[d]), whose lowest log2(box[d]) bits are tile bits (the box spans them)
// and whose other bits come from the tile id (the start coordinate).  Tile bits that do not fit five dimensions are walked
// by n_inst instructions, each with its own index offset and shared-memory offset.  Box elements land in shared memory in
// ascending order of the tile bits' positions (bit i of the slot number before the swizzle = i-th lowest tile bit).
struct TmaTile {
    int rank, n_inst;
    uint32_t box[5];
    uint64_t size[5], stride_bytes[5];
    uint8_t shift[5], bits[5];
    uint64_t inst_goff[TMA_MAX_INST];   // index offset (amplitudes) of instruction e; its box goes to tile + e * inst_amps
    uint32_t inst_amps;
};

// What a radix-8 launch needs beyond the SweepDesc: the engine's view of the tile on load and on store (different for a
// bit-permuting sweep: the tile goes out in the order of its OUTPUT positions) and, for the permuting case, how the last
// stage re-orders the tile in shared memory -- input-local index -> output-local index is a bit permutation (loc_perm), so
// the output slots keep the XOR structure: slot of register 0 plus one offset per register bit (boff_out).
//
// A plain re-ordering store would leave the lanes spread over the INPUT order's low bits, i.e. over high bits of the output
// order: an 8-way bank conflict on every store (measured: 8.5 ms for the permuting QFT-30 sweep against 5.9 ms for the
// others).  Two-stage launches therefore pass through an intermediate layout (mid_layout): the first stage writes the tile
// back in input order but in swz() -- which folds tile bits 3..11 into the bank bits, not only 3..5 -- and the last stage
// assigns its threads so that the lanes walk the bits that are LOW in the output order (tmap_last).  Its loads are then
// conflict-free in swz() and its stores conflict-free in the engine's swizzle of the output order; each stage reads
// everything before it writes (one more barrier).
struct Sweep8Extra {
    TmaTile in, out;
    int permuting;
    SegMap loc_perm;
    uint16_t boff_out[RB8];
    int mid_layout;
    SegMap tmap_last;             // thread id -> input-local index of its register 0 in the last stage
    uint16_t boff_mid_w[RB8];     // swz() slot offset of the first stage's register bit q (its write side)
    uint16_t boff_last_r[RB8];    // swz() slot offset of the last stage's register bit q (its read side)
    uint8_t out_pos[KMAX];         // index bit (output layout) of slot-number bit u of the tile as it leaves: ascending
};

// An op held in registers: header words + 64-byte payload (all accesses statically indexed).
struct OpR {
    uint32_t w0, w1;       // w0 = type | tgt<<8 | rmask<<16 | rval<<24 ; w1 = lmask | lval<<16
    uint64_t gmask, gval;
    uint64_t p[8];         // payload: matrix doubles / stage table / run slots
    __host__ __device__ __forceinline__ uint32_t type() const { return w0 & 0xFFu; }
    __host__ __device__ __forceinline__ uint32_t tgt() const { return (w0 >> 8) & 0xFFu; }
    __host__ __device__ __forceinline__ uint32_t rmask() const { return (w0 >> 16) & 0xFFu; }
    __host__ __device__ __forceinline__ uint32_t rval() const { return w0 >> 24; }
    __host__ __device__ __forceinline__ uint32_t lmask() const { return w1 & 0xFFFFu; }
    __host__ __device__ __forceinline__ uint32_t lval() const { return w1 >> 16; }
    __host__ __device__ __forceinline__ double m(int i) const {
        double d;
        memcpy(&d, &p[i], sizeof(d));  // (bit cast; compiles to nothing on the device)
        return d;
    }
    // OP_STAGE payload: roffsw[16] (u16) then rpos[4] (u8)
    __host__ __device__ __forceinline__ uint32_t roffsw(int j) const { return (uint32_t)(p[j >> 2] >> (16 * (j & 3))) & 0xFFFFu; }
    __host__ __device__ __forceinline__ uint32_t rpos(int q) const { return (uint32_t)(p[4] >> (8 * q)) & 0xFFu; }
    // generic little-endian field access into the payload (offsets are compile-time constants)
    __host__ __device__ __forceinline__ uint32_t u16_at(int off) const { return (uint32_t)(p[off >> 3] >> (8 * (off & 7))) & 0xFFFFu; }
    __host__ __device__ __forceinline__ uint32_t u8_at(int off) const { return (uint32_t)(p[off >> 3] >> (8 * (off & 7))) & 0xFFu; }
    // OP_W payload: w_off @0, wmask @2.  OP_TWID payload: scale @0, tslot[s] @8+2s, gslot[s] @8+2(RB+1)+2s, abit[s] @8+4(RB+1)+s, n_active after
    __host__ __device__ __forceinline__ uint32_t tw_tslot(int s_) const { return u16_at(8 + 2 * s_); }
    __host__ __device__ __forceinline__ uint32_t tw_gslot(int s_) const { return u16_at(8 + 2 * (RB + 1) + 2 * s_); }
    __host__ __device__ __forceinline__ uint32_t tw_abit(int s_) const { return u8_at(8 + 4 * (RB + 1) + s_); }
    __host__ __device__ __forceinline__ uint32_t tw_n() const { return u8_at(8 + 4 * (RB + 1) + RB); }
};

__host__ __device__ __forceinline__ OpR fetch_op(const SwOp* __restrict__ ops, int i) {
    const ulonglong2* w = reinterpret_cast<const ulonglong2*>(ops + i);
    const ulonglong2 a = w[0], b = w[1], c = w[2], d = w[3], e = w[4], f = w[5];
    OpR r;
    r.w0 = (uint32_t)a.x;
    r.w1 = (uint32_t)(a.x >> 32);
    r.gmask = a.y;
    r.gval = b.x;
    r.p[0] = b.y; r.p[1] = c.x; r.p[2] = c.y; r.p[3] = d.x; r.p[4] = d.y; r.p[5] = e.x; r.p[6] = e.y; r.p[7] = f.x;
    return r;
}

template <typename C, bool CTRL, int A>
__host__ __device__ __forceinline__ void op_dense(C (&v)[NREG], const OpR& op, bool tpred) {
    typedef decltype(v[0].x) R;
    const C m00 = cmake<C, R>((R)op.m(0), (R)op.m(1)), m01 = cmake<C, R>((R)op.m(2), (R)op.m(3));
    const C m10 = cmake<C, R>((R)op.m(4), (R)op.m(5)), m11 = cmake<C, R>((R)op.m(6), (R)op.m(7));
    const uint32_t rm = op.rmask(), rv = op.rval();
#pragma unroll
    for (int j0 = 0; j0 < NREG; ++j0)
        if (!(j0 & (1 << A))) {
            const int j1 = j0 | (1 << A);
            if (!CTRL || (((j0 & rm) == rv) && tpred)) {
                const C a0 = v[j0], a1 = v[j1];
                v[j0] = cfma(m01, a1, cmul(m00, a0));
                v[j1] = cfma(m11, a1, cmul(m10, a0));
            }
        }
}

// real-entry matrices (H, rotations about y, ...): half the multiplies
template <typename C, bool CTRL, int A>
__host__ __device__ __forceinline__ void op_dense_real(C (&v)[NREG], const OpR& op, bool tpred) {
    typedef decltype(v[0].x) R;
    const R m00 = (R)op.m(0), m01 = (R)op.m(2), m10 = (R)op.m(4), m11 = (R)op.m(6);
    const uint32_t rm = op.rmask(), rv = op.rval();
#pragma unroll
    for (int j0 = 0; j0 < NREG; ++j0)
        if (!(j0 & (1 << A))) {
            const int j1 = j0 | (1 << A);
            if (!CTRL || (((j0 & rm) == rv) && tpred)) {
                const C a0 = v[j0], a1 = v[j1];
                v[j0].x = fma(m01, a1.x, m00 * a0.x);
                v[j0].y = fma(m01, a1.y, m00 * a0.y);
                v[j1].x = fma(m11, a1.x, m10 * a0.x);
                v[j1].y = fma(m11, a1.y, m10 * a0.y);
            }
        }
}

template <typename C, int A>
__host__ __device__ __forceinline__ void op_permx(C (&v)[NREG], const OpR& op, bool tpred) {
    const uint32_t rm = op.rmask(), rv = op.rval();
#pragma unroll
    for (int j0 = 0; j0 < NREG; ++j0)
        if (!(j0 & (1 << A))) {
            const int j1 = j0 | (1 << A);
            if (((j0 & rm) == rv) && tpred) {
                const C t = v[j0];
                v[j0] = v[j1];
                v[j1] = t;
            }
        }
}

__host__ __device__ __forceinline__ uint32_t expand_rb(uint32_t g, const uint8_t* rpos) {
#pragma unroll
    for (int q = 0; q < RB; ++q) {
        const uint32_t p = rpos[q];
        g = ((g >> p) << (p + 1)) | (g & ((1u << p) - 1u));
    }
    return g;
}

// unnormalised Hadamard butterfly on register bit A: (a0, a1) -> (a0 + a1, a0 - a1)
template <typename C, int A>
__host__ __device__ __forceinline__ void op_hadd(C (&v)[NREG]) {
#pragma unroll
    for (int j0 = 0; j0 < NREG; ++j0)
        if (!(j0 & (1 << A))) {
            const int j1 = j0 | (1 << A);
            const C a0 = v[j0], a1 = v[j1];
            v[j0].x = a0.x + a1.x;
            v[j0].y = a0.y + a1.y;
            v[j1].x = a0.x - a1.x;
            v[j1].y = a0.y - a1.y;
        }
}

// v[j] *= e for every register j that has bit A set
template <typename C, int A>
__host__ __device__ __forceinline__ void mul_bit(C (&v)[NREG], const double2 e) {
    typedef decltype(v[0].x) R;
    const C ee = cmake<C, R>((R)e.x, (R)e.y);
#pragma unroll
    for (int j = 0; j < NREG; ++j)
        if (j & (1 << A)) v[j] = cmul(v[j], ee);
}

// Deferred phases of a stage (OP_TWID and the tail of OP_FSTAGE): thread-wide complex factor / real scale, then one factor per
// register bit -- each the product of a per-tile scalar (smE) and two 16-entry tables over the 4-bit chunks of the thread id.
template <typename C, typename R>
__host__ __device__ __forceinline__ void twid_math(C (&v)[NREG], const OpR& op, const double2* __restrict__ smT, const double2* __restrict__ smE,
                                          const uint32_t tid) {
    const uint32_t c_lo = tid & 15u, c_hi = 16u + (tid >> 4);
    const double scale = op.m(0);
    const uint32_t ts0 = op.tw_tslot(RB), gs0 = op.tw_gslot(RB);
    if (ts0 != NONE16 || gs0 != NONE16) {  // thread-wide complex factor (times the scale)
        double2 c0 = make_double2(scale, 0.0);
        if (gs0 != NONE16) {
            const double2 g0 = smE[gs0];
            c0 = make_double2(scale * g0.x, scale * g0.y);
        }
        if (ts0 != NONE16) c0 = cmul(c0, cmul(smT[ts0 + c_lo], smT[ts0 + c_hi]));
        const C cc = cmake<C, R>((R)c0.x, (R)c0.y);
#pragma unroll
        for (int j = 0; j < NREG; ++j) v[j] = cmul(v[j], cc);
    } else if (scale != 1.0) {
        const R sc = (R)scale;
#pragma unroll
        for (int j = 0; j < NREG; ++j) {
            v[j].x *= sc;
            v[j].y *= sc;
        }
    }
    // slot s_ holds the factor of register bit s_ (abit[s_] == 0xFF: nothing pending on that bit)
#define TWID_BIT(B)                                                             \
    if (op.tw_abit(B) != 0xFFu) {                                               \
        double2 e = make_double2(1.0, 0.0);                                     \
        const uint32_t gsl = op.tw_gslot(B), tsl = op.tw_tslot(B);              \
        if (gsl != NONE16) e = smE[gsl];                                        \
        if (tsl != NONE16) e = cmul(e, cmul(smT[tsl + c_lo], smT[tsl + c_hi])); \
        mul_bit<C, B>(v, e);                                                    \
    }
    TWID_BIT(0)
    TWID_BIT(1)
    TWID_BIT(2)
    TWID_BIT(3)
#undef TWID_BIT
}

// OP_FSTAGE arithmetic, one straight line per SHAPE so that no branch ever joins two register allocations:
//   NH    butterflies on the TOP NH register bits, in the order 3, 2, ... (the host relabels register bits to make it so);
//   HAS_W each butterfly on bit A > 0 is followed by its register-register phase table, which multiplies only registers
//         with bit A set and depends only on the register bits below A (the internal twiddles of a radix-16 FFT
//         butterfly): 7 + 2*3 + 4*1 multiplies, statically indexed (absent tables point at an all-ones table);
//   then the deferred phases: unconditionally on the NH butterfly bits (absent factors point at all-ones slots),
//   behind a uniform branch for the thread-wide factor / scale and for the other register bits (rare).
template <typename C, typename R, int B>
__host__ __device__ __forceinline__ void twid_bit_always(C (&v)[NREG], const OpR& op, const double2* __restrict__ smT, const double2* __restrict__ smE,
                                                const uint32_t c_lo, const uint32_t c_hi) {
    const uint32_t gsl = op.tw_gslot(B), tsl = op.tw_tslot(B);
    const double2 e = cmul(smE[gsl], cmul(smT[tsl + c_lo], smT[tsl + c_hi]));
    mul_bit<C, B>(v, e);
}
template <typename C, typename R, int B>
__host__ __device__ __forceinline__ void twid_bit_maybe(C (&v)[NREG], const OpR& op, const double2* __restrict__ smT, const double2* __restrict__ smE,
                                               const uint32_t c_lo, const uint32_t c_hi) {
    if (op.tw_abit(B) != 0xFFu) twid_bit_always<C, R, B>(v, op, smT, smE, c_lo, c_hi);
}

template <typename C, typename R, int NH, bool HAS_W>
__host__ __device__ __forceinline__ void fstage_math(C (&v)[NREG], const OpR& op, const double2* __restrict__ smT, const double2* __restrict__ smE,
                                            const uint32_t tid) {
    if constexpr (NH >= 1) {
        op_hadd<C, 3>(v);
        if constexpr (HAS_W) {
            const double2* w = smT + op.u16_at(36) + 8;
#pragma unroll
            for (int m = 1; m < 8; ++m) {
                const double2 e = w[m];
                v[8 + m] = cmul(v[8 + m], cmake<C, R>((R)e.x, (R)e.y));
            }
        }
    }
    if constexpr (NH >= 2) {
        op_hadd<C, 2>(v);
        if constexpr (HAS_W) {
            const double2* w = smT + op.u16_at(38) + 4;
#pragma unroll
            for (int m = 1; m < 4; ++m) {
                const double2 e = w[m];
                const C ee = cmake<C, R>((R)e.x, (R)e.y);
                v[4 + m] = cmul(v[4 + m], ee);
                v[12 + m] = cmul(v[12 + m], ee);
            }
        }
    }
    if constexpr (NH >= 3) {
        op_hadd<C, 1>(v);
        if constexpr (HAS_W) {
            const double2 e = smT[op.u16_at(40) + 3];
            const C ee = cmake<C, R>((R)e.x, (R)e.y);
            v[3] = cmul(v[3], ee);
            v[7] = cmul(v[7], ee);
            v[11] = cmul(v[11], ee);
            v[15] = cmul(v[15], ee);
        }
    }
    if constexpr (NH >= 4) op_hadd<C, 0>(v);
    // ---- deferred phases ----
    const uint32_t c_lo = tid & 15u, c_hi = 16u + (tid >> 4);
    const uint32_t fl = op.u8_at(34);
    if (fl & 2u) {  // thread-wide complex factor and / or real scale
        const double scale = op.m(0);
        const uint32_t ts0 = op.tw_tslot(RB), gs0 = op.tw_gslot(RB);
        if (fl & 4u) {
            const double2 g0 = smE[gs0];
            const double2 c0 = cmul(make_double2(scale * g0.x, scale * g0.y), cmul(smT[ts0 + c_lo], smT[ts0 + c_hi]));
            const C cc = cmake<C, R>((R)c0.x, (R)c0.y);
#pragma unroll
            for (int j = 0; j < NREG; ++j) v[j] = cmul(v[j], cc);
        } else {
            const R sc = (R)scale;
#pragma unroll
            for (int j = 0; j < NREG; ++j) {
                v[j].x *= sc;
                v[j].y *= sc;
            }
        }
    }
    if (fl & 1u) {
        if constexpr (NH >= 1) twid_bit_always<C, R, 3>(v, op, smT, smE, c_lo, c_hi); else twid_bit_maybe<C, R, 3>(v, op, smT, smE, c_lo, c_hi);
        if constexpr (NH >= 2) twid_bit_always<C, R, 2>(v, op, smT, smE, c_lo, c_hi); else twid_bit_maybe<C, R, 2>(v, op, smT, smE, c_lo, c_hi);
        if constexpr (NH >= 3) twid_bit_always<C, R, 1>(v, op, smT, smE, c_lo, c_hi); else twid_bit_maybe<C, R, 1>(v, op, smT, smE, c_lo, c_hi);
        if constexpr (NH >= 4) twid_bit_always<C, R, 0>(v, op, smT, smE, c_lo, c_hi); else twid_bit_maybe<C, R, 0>(v, op, smT, smE, c_lo, c_hi);
    }
}

// One whole OP_FSTAGE: shared memory -> registers -> fstage_math -> shared memory, or straight to HBM when this is the
// launch's last op and the stage's register bits allow a coalesced store.  Returns true when it stored to HBM.
template <typename C, typename R, int NH, bool HAS_W>
__host__ __device__ __forceinline__ bool fstage_op(const OpR& op, C* __restrict__ tile, const double2* __restrict__ smT,
                                          const double2* __restrict__ smE, const uint32_t tid, const bool active, const bool to_hbm,
                                          C* __restrict__ dst, const uint64_t gbs, const uint64_t (&st_goff)[NREG]) {
    // (threads beyond the tile's groups -- tiles smaller than a warp's worth -- redo thread 0's loads and never store:
    // an unconditional load keeps the registers defined on every path, which the register allocator needs)
    uint32_t bl = active ? tid : 0u;
#pragma unroll
    for (int q = 0; q < RB; ++q) {
        const uint32_t pz = op.u8_at(42 + q);
        bl = ((bl >> pz) << (pz + 1)) | (bl & ((1u << pz) - 1u));
    }
    // swizzle and register offsets are XOR-linear: slot of register j = slot of register 0 ^ offsets of j's set bits
    uint32_t a[NREG];
    a[0] = swz(bl);
#pragma unroll
    for (int j = 1; j < NREG; ++j) a[j] = a[j & (j - 1)] ^ op.u16_at(46 + 2 * ((j & 1) ? 0 : (j & 2) ? 1 : (j & 4) ? 2 : 3));
    C w[NREG];
#pragma unroll
    for (int j = 0; j < NREG; ++j) w[j] = tile[a[j]];
    fstage_math<C, R, NH, HAS_W>(w, op, smT, smE, tid);
    if (to_hbm) {
        if (active) {
#pragma unroll
            for (int j = 0; j < NREG; ++j) dst[gbs | st_goff[j]] = w[j];
        }
        return true;
    }
    if (active) {
        // the 16 slot numbers are rebuilt from slot 0 (15 XORs) rather than kept in registers across the arithmetic:
        // the opaque copy stops the compiler from re-using the ones computed for the load
        uint32_t b[NREG];
        b[0] = a[0];
#ifdef __CUDA_ARCH__
        asm volatile("" : "+r"(b[0]));
#endif
#pragma unroll
        for (int j = 1; j < NREG; ++j) b[j] = b[j & (j - 1)] ^ op.u16_at(46 + 2 * ((j & 1) ? 0 : (j & 2) ? 1 : (j & 4) ? 2 : 3));
#pragma unroll
        for (int j = 0; j < NREG; ++j) tile[b[j]] = w[j];
    }
#ifdef __CUDA_ARCH__
    __syncthreads();
#endif
    return false;
}

// ---- radix-8 stages (IMG_FFT8) ---------------------------------------------------------------------------------------
// The same OP_FSTAGE record with three register bits: butterflies on register bits 2, 1, 0 in that order, the table after
// the butterfly on bit A multiplies the registers with bit A set by a factor that depends on the bits below A (3 + 2
// multiplies), then the deferred phases of bits 2, 1, 0.  8 amplitudes = 32 registers per thread, so that 256-thread CTAs
// stay under 85 registers and three of them (24 warps) share an SM.
template <typename C, int N, int A>
__host__ __device__ __forceinline__ void hadd_n(C (&v)[N]) {
#pragma unroll
    for (int j0 = 0; j0 < N; ++j0)
        if (!(j0 & (1 << A))) {
            const int j1 = j0 | (1 << A);
            const C a0 = v[j0], a1 = v[j1];
            v[j0].x = a0.x + a1.x;
            v[j0].y = a0.y + a1.y;
            v[j1].x = a0.x - a1.x;
            v[j1].y = a0.y - a1.y;
        }
}
template <typename C, int N, int A>
__host__ __device__ __forceinline__ void mul_bit_n(C (&v)[N], const double2 e) {
    typedef decltype(v[0].x) R;
    const C ee = cmake<C, R>((R)e.x, (R)e.y);
#pragma unroll
    for (int j = 0; j < N; ++j)
        if (j & (1 << A)) v[j] = cmul(v[j], ee);
}

// An op record read where it lies.  In k_sweep8 the records sit in kernel-parameter (constant-bank) space, so every field is
// ONE load of its own width -- no 96-byte fetch, no shift-and-mask extraction.  Offsets are those of OpR (payload-relative).
struct OpView {
    const SwOp* o;
    __host__ __device__ __forceinline__ const unsigned char* pay() const { return reinterpret_cast<const unsigned char*>(o) + 24; }
    __host__ __device__ __forceinline__ uint32_t type() const { return o->type; }
    __host__ __device__ __forceinline__ uint32_t u16_at(int off) const { return *reinterpret_cast<const uint16_t*>(pay() + off); }
    __host__ __device__ __forceinline__ uint32_t u8_at(int off) const { return pay()[off]; }
    __host__ __device__ __forceinline__ double m(int i) const { return *reinterpret_cast<const double*>(pay() + 8 * i); }
    __host__ __device__ __forceinline__ uint32_t tw_tslot(int s_) const { return u16_at(8 + 2 * s_); }
    __host__ __device__ __forceinline__ uint32_t tw_gslot(int s_) const { return u16_at(8 + 2 * (RB + 1) + 2 * s_); }
    __host__ __device__ __forceinline__ uint32_t tw_abit(int s_) const { return u8_at(8 + 4 * (RB + 1) + s_); }
    __host__ __device__ __forceinline__ uint32_t rpos(int q) const { return u8_at(42 + q); }
    __host__ __device__ __forceinline__ uint32_t boff(int q) const { return u16_at(46 + 2 * q); }
    // header of a generic op (SwOp bytes 0..23) and the payload of a stage op (roffsw[16] u16, rpos[4] u8)
    __host__ __device__ __forceinline__ uint32_t tgt() const { return o->tgt; }
    __host__ __device__ __forceinline__ uint32_t rmask() const { return o->rmask; }
    __host__ __device__ __forceinline__ uint32_t rval() const { return o->rval; }
    __host__ __device__ __forceinline__ uint32_t lmask() const { return o->lmask; }
    __host__ __device__ __forceinline__ uint32_t lval() const { return o->lval; }
    __host__ __device__ __forceinline__ uint64_t gmask() const { return o->gmask; }
    __host__ __device__ __forceinline__ uint64_t gval() const { return o->gval; }
    __host__ __device__ __forceinline__ uint32_t st_roffsw(int j) const { return u16_at(2 * j); }
    __host__ __device__ __forceinline__ uint32_t st_rpos(int q) const { return u8_at(32 + q); }
};

template <typename C, int B, typename OP>
__host__ __device__ __forceinline__ void twid8_bit(C (&v)[NREG8], const OP& op, const double2* __restrict__ smT, const double2* __restrict__ smE,
                                                   const uint32_t c_lo, const uint32_t c_hi) {
    const uint32_t gsl = op.tw_gslot(B), tsl = op.tw_tslot(B);
    mul_bit_n<C, NREG8, B>(v, cmul(smE[gsl], cmul(smT[tsl + c_lo], smT[tsl + c_hi])));
}

template <typename C, typename R, int NH, bool HAS_W, typename OP>
__host__ __device__ __forceinline__ void fstage8_math(C (&v)[NREG8], const OP& op, const double2* __restrict__ smT,
                                                      const double2* __restrict__ smE, const uint32_t tid) {
    if constexpr (NH >= 1) {
        hadd_n<C, NREG8, 2>(v);
        if constexpr (HAS_W) {
            const double2* w = smT + op.u16_at(36) + 4;
#pragma unroll
            for (int m = 1; m < 4; ++m) {
                const double2 e = w[m];
                v[4 + m] = cmul(v[4 + m], cmake<C, R>((R)e.x, (R)e.y));
            }
        }
    }
    if constexpr (NH >= 2) {
        hadd_n<C, NREG8, 1>(v);
        if constexpr (HAS_W) {
            const double2 e = smT[op.u16_at(38) + 3];
            const C ee = cmake<C, R>((R)e.x, (R)e.y);
            v[3] = cmul(v[3], ee);
            v[7] = cmul(v[7], ee);
        }
    }
    if constexpr (NH >= 3) hadd_n<C, NREG8, 0>(v);
    const uint32_t c_lo = tid & 15u, c_hi = 16u + (tid >> 4);
    const uint32_t fl = op.u8_at(34);
    if (fl & 2u) {  // thread-wide complex factor and / or real scale
        const double scale = op.m(0);
        if (fl & 4u) {
            const uint32_t ts0 = op.tw_tslot(RB), gs0 = op.tw_gslot(RB);
            const double2 g0 = smE[gs0];
            const double2 c0 = cmul(make_double2(scale * g0.x, scale * g0.y), cmul(smT[ts0 + c_lo], smT[ts0 + c_hi]));
            const C cc = cmake<C, R>((R)c0.x, (R)c0.y);
#pragma unroll
            for (int j = 0; j < NREG8; ++j) v[j] = cmul(v[j], cc);
        } else {
            const R sc = (R)scale;
#pragma unroll
            for (int j = 0; j < NREG8; ++j) {
                v[j].x *= sc;
                v[j].y *= sc;
            }
        }
    }
    if (fl & 1u) {
        if (NH >= 1 || op.tw_abit(2) != 0xFFu) twid8_bit<C, 2>(v, op, smT, smE, c_lo, c_hi);
        if (NH >= 2 || op.tw_abit(1) != 0xFFu) twid8_bit<C, 1>(v, op, smT, smE, c_lo, c_hi);
        if (NH >= 3 || op.tw_abit(0) != 0xFFu) twid8_bit<C, 0>(v, op, smT, smE, c_lo, c_hi);
    }
}

// thread id -> local index of its register 0: zero bits inserted at the stage's three register positions (ascending)
__host__ __device__ __forceinline__ uint32_t expand_rb8(uint32_t g, const uint32_t p0, const uint32_t p1, const uint32_t p2) {
    g = ((g >> p0) << (p0 + 1)) | (g & ((1u << p0) - 1u));
    g = ((g >> p1) << (p1 + 1)) | (g & ((1u << p1) - 1u));
    g = ((g >> p2) << (p2 + 1)) | (g & ((1u << p2) - 1u));
    return g;
}

// ... and back: the thread number the planner's thread-phase tables assume for a register-0 index (positions descending)
__host__ __device__ __forceinline__ uint32_t compress_rb8(uint32_t g, const uint32_t p0, const uint32_t p1, const uint32_t p2) {
    g = ((g >> (p2 + 1)) << p2) | (g & ((1u << p2) - 1u));
    g = ((g >> (p1 + 1)) << p1) | (g & ((1u << p1) - 1u));
    g = ((g >> (p0 + 1)) << p0) | (g & ((1u << p0) - 1u));
    return g;
}

// The eight tile slots of a thread from the slot of its register 0 (swizzle and register offsets are XOR-linear).
__host__ __device__ __forceinline__ void slots8(uint32_t (&a)[NREG8], const uint32_t a0, const uint32_t b0, const uint32_t b1, const uint32_t b2) {
    a[0] = a0;
    a[1] = a0 ^ b0;
    a[2] = a0 ^ b1;
    a[3] = a[1] ^ b1;
    a[4] = a0 ^ b2;
    a[5] = a[1] ^ b2;
    a[6] = a[2] ^ b2;
    a[7] = a[3] ^ b2;
}

// One radix-8 stage on a tile in the TMA layout, first half: shared memory -> 8 registers -> fstage8_math.
// `a0` = slot of the thread's register 0 = swz128(expand_rb8(tid, rpos...)) (tile independent: the kernel keeps it parked).
template <typename C, typename R, typename OP>
__host__ __device__ __forceinline__ void fstage8_compute(const OP& op, const C* tile, const double2* __restrict__ smT,
                                                         const double2* __restrict__ smE, const uint32_t tid, const uint32_t a0,
                                                         const uint32_t b0, const uint32_t b1, const uint32_t b2, C (&w)[NREG8]) {
    uint32_t a[NREG8];
    slots8(a, a0, b0, b1, b2);
#pragma unroll
    for (int j = 0; j < NREG8; ++j) w[j] = tile[a[j]];
    switch (op.type() - DC_FSTAGE) {  // shape = 2 * (number of butterflies) + (phase tables present)
        case 0: case 1: fstage8_math<C, R, 0, false>(w, op, smT, smE, tid); break;
        case 2: fstage8_math<C, R, 1, false>(w, op, smT, smE, tid); break;
        case 3: fstage8_math<C, R, 1, true>(w, op, smT, smE, tid); break;
        case 4: fstage8_math<C, R, 2, false>(w, op, smT, smE, tid); break;
        case 5: fstage8_math<C, R, 2, true>(w, op, smT, smE, tid); break;
        case 6: fstage8_math<C, R, 3, false>(w, op, smT, smE, tid); break;
        default: fstage8_math<C, R, 3, true>(w, op, smT, smE, tid); break;
    }
}
// ... second half: the registers go back to shared memory, to the slots they came from (a0, the stage's offsets) or -- last
// stage of a bit-permuting sweep, after a barrier -- to the slots of the OUTPUT tile order (a0 and offsets of that order).
template <typename C>
__host__ __device__ __forceinline__ void fstage8_put(C* tile, const uint32_t a0, const uint32_t b0, const uint32_t b1, const uint32_t b2,
                                                     const C (&w)[NREG8]) {
    uint32_t a[NREG8];
    slots8(a, a0, b0, b1, b2);
#pragma unroll
    for (int j = 0; j < NREG8; ++j) tile[a[j]] = w[j];
}

// ---- generic ops on 8 registers (radix-8 image, launches that are not FFT-shaped) ------------------------------------
// The same op set as generic_op below, with three register bits and every field read where it lies (OpView).
template <typename C, typename R, bool CTRL, bool REAL, int A, typename OP>
__host__ __device__ __forceinline__ void dense8(C (&v)[NREG8], const OP& op, const bool tpred) {
    const uint32_t rm = op.rmask(), rv = op.rval();
    if constexpr (REAL) {
        const R m00 = (R)op.m(0), m01 = (R)op.m(2), m10 = (R)op.m(4), m11 = (R)op.m(6);
#pragma unroll
        for (int j0 = 0; j0 < NREG8; ++j0)
            if (!(j0 & (1 << A))) {
                const int j1 = j0 | (1 << A);
                const C a0 = v[j0], a1 = v[j1];
                v[j0].x = fma(m01, a1.x, m00 * a0.x);
                v[j0].y = fma(m01, a1.y, m00 * a0.y);
                v[j1].x = fma(m11, a1.x, m10 * a0.x);
                v[j1].y = fma(m11, a1.y, m10 * a0.y);
            }
    } else {
        const C m00 = cmake<C, R>((R)op.m(0), (R)op.m(1)), m01 = cmake<C, R>((R)op.m(2), (R)op.m(3));
        const C m10 = cmake<C, R>((R)op.m(4), (R)op.m(5)), m11 = cmake<C, R>((R)op.m(6), (R)op.m(7));
#pragma unroll
        for (int j0 = 0; j0 < NREG8; ++j0)
            if (!(j0 & (1 << A))) {
                const int j1 = j0 | (1 << A);
                if (!CTRL || (((j0 & rm) == rv) && tpred)) {
                    const C a0 = v[j0], a1 = v[j1];
                    v[j0] = cfma(m01, a1, cmul(m00, a0));
                    v[j1] = cfma(m11, a1, cmul(m10, a0));
                }
            }
    }
}
template <typename C, int A, typename OP>
__host__ __device__ __forceinline__ void permx8(C (&v)[NREG8], const OP& op, const bool tpred) {
    const uint32_t rm = op.rmask(), rv = op.rval();
#pragma unroll
    for (int j0 = 0; j0 < NREG8; ++j0)
        if (!(j0 & (1 << A))) {
            const int j1 = j0 | (1 << A);
            if (((j0 & rm) == rv) && tpred) {
                const C t = v[j0];
                v[j0] = v[j1];
                v[j1] = t;
            }
        }
}
template <typename C, typename R, typename OP>
__host__ __device__ __forceinline__ void wtable8(C (&v)[NREG8], const OP& op, const double2* __restrict__ smT) {
    const double2* w = smT + op.u16_at(0);
    const uint32_t wm = op.u16_at(2);
#pragma unroll
    for (int j = 0; j < NREG8; ++j)
        if ((wm >> j) & 1) {
            const double2 ww = w[j];
            v[j] = cmul(v[j], cmake<C, R>((R)ww.x, (R)ww.y));
        }
}
template <typename C, typename R, typename OP>
__host__ __device__ __forceinline__ void twid8_math(C (&v)[NREG8], const OP& op, const double2* __restrict__ smT, const double2* __restrict__ smE,
                                                    const uint32_t tid) {
    const uint32_t c_lo = tid & 15u, c_hi = 16u + (tid >> 4);
    const double scale = op.m(0);
    const uint32_t ts0 = op.tw_tslot(RB), gs0 = op.tw_gslot(RB);
    if (ts0 != NONE16 || gs0 != NONE16) {
        double2 c0 = make_double2(scale, 0.0);
        if (gs0 != NONE16) {
            const double2 g0 = smE[gs0];
            c0 = make_double2(scale * g0.x, scale * g0.y);
        }
        if (ts0 != NONE16) c0 = cmul(c0, cmul(smT[ts0 + c_lo], smT[ts0 + c_hi]));
        const C cc = cmake<C, R>((R)c0.x, (R)c0.y);
#pragma unroll
        for (int j = 0; j < NREG8; ++j) v[j] = cmul(v[j], cc);
    } else if (scale != 1.0) {
        const R sc = (R)scale;
#pragma unroll
        for (int j = 0; j < NREG8; ++j) {
            v[j].x *= sc;
            v[j].y *= sc;
        }
    }
#define TWID8_BIT(B)                                                            \
    if (op.tw_abit(B) != 0xFFu) {                                               \
        double2 e = make_double2(1.0, 0.0);                                     \
        const uint32_t gsl = op.tw_gslot(B), tsl = op.tw_tslot(B);              \
        if (gsl != NONE16) e = smE[gsl];                                        \
        if (tsl != NONE16) e = cmul(e, cmul(smT[tsl + c_lo], smT[tsl + c_hi])); \
        mul_bit_n<C, NREG8, B>(v, e);                                           \
    }
    TWID8_BIT(0)
    TWID8_BIT(1)
    TWID8_BIT(2)
#undef TWID8_BIT
}

template <typename C, typename R, typename OP>
__host__ __device__ __forceinline__ void generic_op8(C (&v)[NREG8], const OP& op, const uint32_t type, const bool tpred,
                                                     const double2* __restrict__ smT, const double2* __restrict__ smE, const uint32_t tid) {
    switch (type) {
        case DC_DENSE_REAL_NC + 0: dense8<C, R, false, true, 0>(v, op, tpred); break;
        case DC_DENSE_REAL_NC + 1: dense8<C, R, false, true, 1>(v, op, tpred); break;
        case DC_DENSE_REAL_NC + 2: dense8<C, R, false, true, 2>(v, op, tpred); break;
        case DC_DENSE_CPLX_NC + 0: dense8<C, R, false, false, 0>(v, op, tpred); break;
        case DC_DENSE_CPLX_NC + 1: dense8<C, R, false, false, 1>(v, op, tpred); break;
        case DC_DENSE_CPLX_NC + 2: dense8<C, R, false, false, 2>(v, op, tpred); break;
        case DC_DENSE_CPLX_C + 0: dense8<C, R, true, false, 0>(v, op, tpred); break;
        case DC_DENSE_CPLX_C + 1: dense8<C, R, true, false, 1>(v, op, tpred); break;
        case DC_DENSE_CPLX_C + 2: dense8<C, R, true, false, 2>(v, op, tpred); break;
        case DC_PERMX + 0: permx8<C, 0>(v, op, tpred); break;
        case DC_PERMX + 1: permx8<C, 1>(v, op, tpred); break;
        case DC_PERMX + 2: permx8<C, 2>(v, op, tpred); break;
        case DC_MPHASE: {
            const C d = cmake<C, R>((R)op.m(0), (R)op.m(1));
            const uint32_t rm = op.rmask(), rv = op.rval();
#pragma unroll
            for (int j = 0; j < NREG8; ++j)
                if (((j & rm) == rv) && tpred) v[j] = cmul(v[j], d);
            break;
        }
        case DC_HADD + 0: hadd_n<C, NREG8, 0>(v); break;
        case DC_HADD + 1: hadd_n<C, NREG8, 1>(v); break;
        case DC_HADD + 2: hadd_n<C, NREG8, 2>(v); break;
        case DC_W: wtable8<C, R>(v, op, smT); break;
        case DC_HW + 0: hadd_n<C, NREG8, 0>(v); wtable8<C, R>(v, op, smT); break;
        case DC_HW + 1: hadd_n<C, NREG8, 1>(v); wtable8<C, R>(v, op, smT); break;
        case DC_HW + 2: hadd_n<C, NREG8, 2>(v); wtable8<C, R>(v, op, smT); break;
        case DC_TWID: twid8_math<C, R>(v, op, smT, smE, tid); break;
        default: break;
    }
}

// One op of the generic interpreter applied to a thread's 16 registers (everything but the stage changes).
template <typename C, typename R, bool LEAN>
__host__ __device__ __forceinline__ void generic_op(C (&v)[NREG], const OpR& op, const uint32_t type, const bool tpred,
                                                    const double2* __restrict__ smT, const double2* __restrict__ smE, const uint32_t tid) {
#define DENSE_CASES(BASE, FN, CTRL)                                   \
    case BASE + 0: FN<C, CTRL, 0>(v, op, tpred); break;               \
    case BASE + 1: FN<C, CTRL, 1>(v, op, tpred); break;               \
    case BASE + 2: FN<C, CTRL, 2>(v, op, tpred); break;               \
    case BASE + 3: FN<C, CTRL, 3>(v, op, tpred); break;
    switch (type) {
        DENSE_CASES(DC_DENSE_REAL_NC, op_dense_real, false)
        DENSE_CASES(DC_DENSE_CPLX_NC, op_dense, false)
        case DC_DENSE_CPLX_C + 0: if constexpr (!LEAN) op_dense<C, true, 0>(v, op, tpred); break;
        case DC_DENSE_CPLX_C + 1: if constexpr (!LEAN) op_dense<C, true, 1>(v, op, tpred); break;
        case DC_DENSE_CPLX_C + 2: if constexpr (!LEAN) op_dense<C, true, 2>(v, op, tpred); break;
        case DC_DENSE_CPLX_C + 3: if constexpr (!LEAN) op_dense<C, true, 3>(v, op, tpred); break;
        case DC_PERMX + 0: if constexpr (!LEAN) op_permx<C, 0>(v, op, tpred); break;
        case DC_PERMX + 1: if constexpr (!LEAN) op_permx<C, 1>(v, op, tpred); break;
        case DC_PERMX + 2: if constexpr (!LEAN) op_permx<C, 2>(v, op, tpred); break;
        case DC_PERMX + 3: if constexpr (!LEAN) op_permx<C, 3>(v, op, tpred); break;
        case DC_MPHASE: {
            if constexpr (!LEAN) {
                const C d = cmake<C, R>((R)op.m(0), (R)op.m(1));
                const uint32_t rm = op.rmask(), rv = op.rval();
#pragma unroll
                for (int j = 0; j < NREG; ++j)
                    if (((j & rm) == rv) && tpred) v[j] = cmul(v[j], d);
            }
            break;
        }
        case DC_HADD + 0: op_hadd<C, 0>(v); break;
        case DC_HADD + 1: op_hadd<C, 1>(v); break;
        case DC_HADD + 2: op_hadd<C, 2>(v); break;
        case DC_HADD + 3: op_hadd<C, 3>(v); break;
#define APPLY_W()                                                              \
    {                                                                          \
const double2* w = smT + op.u16_at(0);                                 \
const uint32_t wm = op.u16_at(2);                                      \
_Pragma("unroll") for (int j = 0; j < NREG; ++j) if ((wm >> j) & 1) {  \
    const double2 ww = w[j];                                           \
    v[j] = cmul(v[j], cmake<C, R>((R)ww.x, (R)ww.y));                  \
}                                                                      \
    }
        case DC_W: APPLY_W() break;
        case DC_HW + 0: op_hadd<C, 0>(v); APPLY_W() break;
        case DC_HW + 1: op_hadd<C, 1>(v); APPLY_W() break;
        case DC_HW + 2: op_hadd<C, 2>(v); APPLY_W() break;
        case DC_HW + 3: op_hadd<C, 3>(v); APPLY_W() break;
#undef APPLY_W
        case DC_TWID: twid_math<C, R>(v, op, smT, smE, tid); break;
        default:
            break;
    }
#undef DENSE_CASES
}

// Three images per (dtype, tile size):
//   IMG_FULL  every op kind;
//   IMG_LEAN  the op set of Hadamard/phase circuits only (stage, uncontrolled dense, unnormalised butterfly, phase tables);
//   IMG_FFT   launches made of OP_FSTAGE ops only: no amplitude is carried in registers from one op to the next, so the
//             op loop has no register-allocation joins (the generic loop spends ~20 % of its instructions on moves there).
//   IMG_FFT8  the same shape with three register bits per stage, 256-thread CTAs, TMA tile loads and stores (qsv_sweep8.cu);
//             chosen per launch when the sweep's targets fit radix-8 stages as well and its tile fits the TMA descriptor.
#define IMG_FULL 0
#define IMG_LEAN 1
#define IMG_FFT 2
#define IMG_FFT8 3

struct PlanStep {
    bool fused;
    size_t gate_index;      // !fused: index into ctx->gates
    SweepParams* params;    // fused (host copy)
    bool out_of_place;
    int img;                // kernel image that runs it: IMG_FFT / IMG_FFT8 (only OP_FSTAGE ops), IMG_LEAN (lean op set) or IMG_FULL
    uint64_t n_gates;       // source gates carried
    Sweep8Extra* tma;       // IMG_FFT8: how the TMA engine sees a tile on load / store (host copy, owned)
    void* launch_cache;     // IMG_FFT8: kernel parameters and tensor maps of the last launch (qsv_sweep8.cu), reused while the buffers stay
};

void sweep8_free_cache(void* cache);  // qsv_sweep8.cu

struct SweepPlan {
    std::vector<PlanStep> steps;
    std::vector<double> aux;  // host copy of tables
    double* d_aux = nullptr;
    int k = 0;
    uint64_t dropped_identities = 0;
    std::vector<cudaEvent_t> ev;  // profiling
    std::vector<float> ms;
    ~SweepPlan() {
        for (auto& s : steps) {
            delete s.params;
            delete s.tma;
            sweep8_free_cache(s.launch_cache);
        }
        if (d_aux) cudaFree(d_aux);
        for (auto e : ev) cudaEventDestroy(e);
    }
};


// the fusion pass (qsv_fuse.cu, host only: builds plans without touching a device)
int sweep_plan_build_host(qsv_ctx* ctx, int max_tile_bits, SweepPlan** out_plan, const int* final_pos = nullptr, bool carry_exchange = false);
int sweep_permute_plan_build(qsv_ctx* ctx, int n_bits, const int* perm, int k, SweepPlan* tmp);
