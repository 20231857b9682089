// Column-slab kernel for streams of two-level (Givens) row operations -- BASELINE config 4.
//
// A two-level UnitaryM (quompiler/construct/unitary.py:20-46) mixes two rows of the [dim][batch] array.
// The SINGLET-level stream the compiler emits for it (Gray-code chain of fully controlled X gates around one
// fully controlled U, quompiler/utils/cnot_decompose.py:84-107) is the same thing written with qubits:
// a gate whose controls cover every other qubit acts on exactly two rows.  The host pass below folds
//   * fully controlled X  -> a relabel of two row names (no data moves, bit-exact by construction),
//   * fully controlled U  -> one row-pair op on the current names of its two rows,
//   * QSV_ROWS2 records   -> one row-pair op,
// and the kernel keeps a slab of C columns x all rows resident in shared memory while it walks the
// whole op list: HBM traffic is one read and one write of the state for the entire program.
//
// Scheduling (round 2).  The first version gave every column to one lane of a single warp and walked the op list in order:
// 8 busy lanes per CTA and a dependent chain of half a million ops through shared memory (55.7 ms for config 4, 106 ns per
// op).  The chain is a property of that schedule, not of the problem: two row-pair ops depend on each other only if they
// share a row.  The host therefore LEVELS the ops -- level(op) = 1 + max(level of the last op on either of its rows) -- and
// sorts them by level (stable).  All ops of a level touch pairwise disjoint rows, so a CTA of up to 1024 threads applies
// them at once, one (op, column) pair per thread, with ONE CTA barrier per level; for the Givens QR stream of a 1024 x 1024
// unitary that is ~2 000 levels of ~256 ops instead of 523 776 sequential ops.  The ops of the next level are staged into
// shared memory with cp.async while the current level computes.
#include <algorithm>

#include "qsv_common.cuh"

struct __align__(16) RowOp {  // 80 bytes
    uint32_t r0, r1;
    uint32_t swap;  // 1: exchange the two rows (kept arithmetic-free)
    uint32_t pad;
    double m[8];
};
static_assert(sizeof(RowOp) == 80, "RowOp layout");

struct RowSlabPlan {
    RowOp* d_ops = nullptr;          // sorted by level
    uint32_t* d_level_start = nullptr;  // [n_levels + 1] offsets into d_ops
    uint64_t n_ops = 0, n_levels = 0;
    uint64_t n_source_gates = 0;
    uint64_t dim = 0, batch = 0;
    int cols = 0;     // columns per CTA (power of two)
    int threads = 0;  // CTA size
    ~RowSlabPlan() {
        if (d_ops) cudaFree(d_ops);
        if (d_level_start) cudaFree(d_level_start);
    }
};

// Ops of one level are contiguous in the sorted stream; the ops of level l+1 are copied into a shared-memory staging buffer
// with cp.async while level l computes (two buffers), so that no pass of a level waits on L2 and no op sits in registers.
#define ROWLEVEL_STAGE_OPS 512   // ops per staging buffer (40 KB); wider levels (never seen for Givens streams) read the rest from L2

// One THREAD per op, all COLS columns of the CTA's slab in its registers: the 80-byte op record is then read once per op
// instead of once per (op, column) -- ncu of the (op, column)-per-thread version showed the shared-memory pipe 80 % busy,
// more than half of it on op records (profiles/r2q_ncu_full_rowlevels_config4.csv) -- and the 2 x COLS butterflies of a
// thread are independent.  Rows are stored with the column index XOR-ed with the row number, so that lanes holding
// different rows read different banks.
template <typename R, int COLS>
__global__ void __launch_bounds__(512, 1) k_rowlevels(typename cplx_of<R>::type* __restrict__ st, uint64_t dim, uint64_t batch,
                                                     const RowOp* __restrict__ ops, const uint32_t* __restrict__ level_start,
                                                     uint32_t n_levels) {
    typedef typename cplx_of<R>::type C;
    constexpr int CL = COLS == 1 ? 0 : COLS == 2 ? 1 : COLS == 4 ? 2 : 3;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RowOp* stage = reinterpret_cast<RowOp*>(smem_raw);                                        // [2][ROWLEVEL_STAGE_OPS]
    C* slab = reinterpret_cast<C*>(smem_raw + 2 * ROWLEVEL_STAGE_OPS * sizeof(RowOp));        // [dim][COLS], column c of row r at c ^ (r % COLS)
    const uint32_t tid = threadIdx.x, T = blockDim.x;
    const uint64_t col0 = (uint64_t)blockIdx.x << CL;
    const uint64_t slab_amps = dim << CL;
    for (uint64_t i = tid; i < slab_amps; i += T) {
        const uint64_t r = i >> CL, c = i & (COLS - 1);
        slab[(r << CL) | (c ^ (r & (COLS - 1)))] = st[r * batch + col0 + c];
    }
    auto stage_level = [&](uint32_t first, uint32_t last, int buf) {  // all threads: 16-byte pieces of the level's op records
        const uint32_t n = min(last - first, (uint32_t)ROWLEVEL_STAGE_OPS);
        const char* src = reinterpret_cast<const char*>(ops + first);
        char* dst = reinterpret_cast<char*>(stage + buf * ROWLEVEL_STAGE_OPS);
        for (uint32_t b = tid * 16u; b < n * (uint32_t)sizeof(RowOp); b += T * 16u) {
            const uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst + b);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(src + b));
        }
        asm volatile("cp.async.commit_group;\n" ::);
    };
    auto apply = [&](const RowOp& o) {
        C* p0 = slab + ((uint64_t)o.r0 << CL);
        C* p1 = slab + ((uint64_t)o.r1 << CL);
        const uint32_t x0 = o.r0 & (COLS - 1), x1 = o.r1 & (COLS - 1);
        C a0[COLS], a1[COLS];
#pragma unroll
        for (int c = 0; c < COLS; ++c) {
            a0[c] = p0[c ^ x0];
            a1[c] = p1[c ^ x1];
        }
        if (o.swap) {
#pragma unroll
            for (int c = 0; c < COLS; ++c) {
                p0[c ^ x0] = a1[c];
                p1[c ^ x1] = a0[c];
            }
        } else {
            const C m00 = cmake<C, R>((R)o.m[0], (R)o.m[1]), m01 = cmake<C, R>((R)o.m[2], (R)o.m[3]);
            const C m10 = cmake<C, R>((R)o.m[4], (R)o.m[5]), m11 = cmake<C, R>((R)o.m[6], (R)o.m[7]);
#pragma unroll
            for (int c = 0; c < COLS; ++c) {
                p0[c ^ x0] = cfma(m01, a1[c], cmul(m00, a0[c]));
                p1[c ^ x1] = cfma(m11, a1[c], cmul(m10, a0[c]));
            }
        }
    };
    uint32_t s = n_levels ? level_start[0] : 0, e = n_levels ? level_start[1] : 0;
    if (n_levels) stage_level(s, e, 0);
    for (uint32_t l = 0; l < n_levels; ++l) {
        const uint32_t s2 = e, e2 = level_start[l + 2];   // (two padding entries follow the last level)
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncthreads();   // level l's ops are staged; level l-1's row updates are visible; the other staging buffer is free
        if (l + 1 < n_levels) stage_level(s2, e2, (l + 1) & 1);
        const RowOp* cur = stage + (l & 1) * ROWLEVEL_STAGE_OPS;
        const uint32_t n = e - s;
        for (uint32_t i = tid; i < n; i += T) apply(i < ROWLEVEL_STAGE_OPS ? cur[i] : ops[s + i]);
        s = s2;
        e = e2;
    }
    __syncthreads();
    for (uint64_t i = tid; i < slab_amps; i += T) {
        const uint64_t r = i >> CL, c = i & (COLS - 1);
        st[r * batch + col0 + c] = slab[(r << CL) | (c ^ (r & (COLS - 1)))];
    }
}

template <typename R>
static cudaError_t rowlevels_dispatch(bool attr_only, int cols, int smem, unsigned grid, int threads, cudaStream_t stream, void* st, uint64_t dim,
                                      uint64_t batch, const RowOp* ops, const uint32_t* ls, uint32_t n_levels) {
    typedef typename cplx_of<R>::type C;
#define RL(COLS)                                                                                                             \
    if (attr_only) return cudaFuncSetAttribute(k_rowlevels<R, COLS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);     \
    k_rowlevels<R, COLS><<<grid, threads, smem, stream>>>((C*)st, dim, batch, ops, ls, n_levels);                            \
    return cudaGetLastError();
    switch (cols) {
        case 1: RL(1)
        case 2: RL(2)
        case 4: RL(4)
        default: RL(8)
    }
#undef RL
}

void rowslab_free(RowSlabPlan* p) { delete p; }

// Returns QSV_OK and a plan if the whole uploaded program consists of row-pair operations, QSV_E_UNSUPPORTED
// otherwise (the caller then uses the general paths).
int rowslab_build(qsv_ctx* ctx, RowSlabPlan** out) {
    *out = nullptr;
    const int b = ctx->prog_batch_bits, n = ctx->prog_bits;
    const int q = n - b;
    if (q < 1 || q > 13 || ctx->gates.empty()) return QSV_E_UNSUPPORTED;
    const uint64_t dim = 1ull << q, batch = 1ull << b;
    const uint64_t qmask = (dim - 1) << b;
    // columns per CTA: as many as fit next to the op buffers in ~200 KB, at most a warp and at most the batch
    const size_t amp = amp_size(ctx->dtype);
    const size_t budget = 224 * 1024 - 2 * ROWLEVEL_STAGE_OPS * sizeof(RowOp);
    uint64_t cols = budget / (dim * amp);
    if (cols < 1) return QSV_E_UNSUPPORTED;
    cols = std::min<uint64_t>(std::min<uint64_t>(cols, 8), batch);   // (a thread holds 2 x cols amplitudes in registers)
    while (cols & (cols - 1)) cols &= cols - 1;  // power of two
    {
        // The work per level is fixed per COLUMN, and a level's latency (barrier + one trip through shared memory) does not
        // shrink with fewer columns per CTA: spread a small batch over the SMs rather than filling a few CTAs (128 columns of
        // config 4 on each of 8 GPUs: 16 CTAs of 8 columns took 3.5 ms, as long as 128 CTAs of 8 on one GPU).
        int sm_count = 148;
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, ctx->device);
        while (cols > 1 && batch / cols < (uint64_t)sm_count / 2) cols >>= 1;
    }
    if (batch % cols) return QSV_E_UNSUPPORTED;
    // Worth it only for long streams on states that are small next to the number of gates
    if (ctx->gates.size() < 4 * (size_t)q) return QSV_E_UNSUPPORTED;
    std::vector<uint32_t> name(dim);  // name[i] = physical row currently holding logical row i
    for (uint64_t i = 0; i < dim; ++i) name[i] = (uint32_t)i;
    std::vector<RowOp> ops;
    ops.reserve(ctx->gates.size() / 4 + 16);
    for (const qsv_gate_t& g : ctx->gates) {
        RowOp op;
        memset(&op, 0, sizeof(op));
        uint64_t i0, i1;
        const double* m = ctx->pool.data() + 2 * g.matrix_offset;
        if (g.kind == QSV_ROWS2) {
            i0 = g.row0;
            i1 = g.row1;
        } else if (g.n_targets == 1 && (g.kind == QSV_DENSE1 || g.kind == QSV_PERMX || g.kind == QSV_DIAG1)) {
            const uint64_t tm = 1ull << g.target_bits[0];
            if ((g.ctrl_mask | tm) != qmask) return QSV_E_UNSUPPORTED;  // not fully controlled
            i0 = g.ctrl_value >> b;
            i1 = (g.ctrl_value | tm) >> b;
            if (g.kind == QSV_PERMX) {  // relabel, no data movement
                std::swap(name[i0], name[i1]);
                continue;
            }
        } else {
            return QSV_E_UNSUPPORTED;
        }
        op.r0 = name[i0];
        op.r1 = name[i1];
        memcpy(op.m, m, sizeof(op.m));
        ops.push_back(op);
    }
    // rows that ended up under a different name: physical exchange (cycle decomposition into transpositions)
    {
        std::vector<uint32_t> at(name);  // at[i] = physical row holding logical i; want at[i] == i
        std::vector<uint32_t> inv(dim);
        for (uint64_t i = 0; i < dim; ++i) inv[at[i]] = (uint32_t)i;  // inv[p] = logical row stored in physical p
        for (uint64_t i = 0; i < dim; ++i) {
            if (at[i] == i) continue;
            const uint32_t p = at[i];  // logical i lives in p; logical inv[i] lives in i
            RowOp op;
            memset(&op, 0, sizeof(op));
            op.r0 = (uint32_t)i;
            op.r1 = p;
            op.swap = 1;
            ops.push_back(op);
            const uint32_t other = inv[i];
            at[other] = p;
            inv[p] = other;
            at[i] = (uint32_t)i;
            inv[i] = (uint32_t)i;
        }
    }
    // level schedule: an op waits only for the last op on either of its rows
    std::vector<uint32_t> level(ops.size()), last(dim, 0);
    uint32_t n_levels = 0;
    for (size_t i = 0; i < ops.size(); ++i) {
        const uint32_t lv = std::max(last[ops[i].r0], last[ops[i].r1]) + 1;
        level[i] = lv - 1;
        last[ops[i].r0] = last[ops[i].r1] = lv;
        n_levels = std::max(n_levels, lv);
    }
    std::vector<uint32_t> start(n_levels + 3, 0);  // (+2 padding entries: the kernel reads two levels ahead)
    for (size_t i = 0; i < ops.size(); ++i) start[level[i] + 1]++;
    for (uint32_t l = 0; l < n_levels; ++l) start[l + 1] += start[l];
    start[n_levels + 1] = start[n_levels + 2] = start[n_levels];
    std::vector<RowOp> sorted(ops.size());
    {
        std::vector<uint32_t> fill(start.begin(), start.begin() + n_levels + 1);
        for (size_t i = 0; i < ops.size(); ++i) sorted[fill[level[i]]++] = ops[i];
    }
    RowSlabPlan* plan = new RowSlabPlan();
    plan->n_ops = ops.size();
    plan->n_levels = n_levels;
    plan->n_source_gates = ctx->gates.size();
    plan->dim = dim;
    plan->batch = batch;
    plan->cols = (int)cols;
    {
        // CTA size: one thread per op of the widest level, a multiple of the warp, at most 512
        uint32_t widest = 1;
        for (uint32_t l = 0; l < n_levels; ++l) widest = std::max(widest, start[l + 1] - start[l]);
        plan->threads = (int)std::min<uint32_t>(512, (widest + 31) / 32 * 32);
    }
    if (!ops.empty()) {
        cudaError_t e = cudaMalloc(&plan->d_ops, sorted.size() * sizeof(RowOp));
        if (e == cudaSuccess) e = cudaMemcpy(plan->d_ops, sorted.data(), sorted.size() * sizeof(RowOp), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMalloc(&plan->d_level_start, start.size() * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaMemcpy(plan->d_level_start, start.data(), start.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            delete plan;
            return qsv_fail(ctx, QSV_E_CUDA, "uploading the row-op stream failed: %s", cudaGetErrorString(e));
        }
    }
    const int smem = (int)(2 * ROWLEVEL_STAGE_OPS * sizeof(RowOp) + dim * cols * amp);
    cudaError_t e = ctx->dtype == QSV_C128
                        ? rowlevels_dispatch<double>(true, (int)cols, smem, 0, 0, nullptr, nullptr, 0, 0, nullptr, nullptr, 0)
                        : rowlevels_dispatch<float>(true, (int)cols, smem, 0, 0, nullptr, nullptr, 0, 0, nullptr, nullptr, 0);
    if (e != cudaSuccess) {
        delete plan;
        return qsv_fail(ctx, QSV_E_CUDA, "cudaFuncSetAttribute(%d B smem) failed: %s", smem, cudaGetErrorString(e));
    }
    *out = plan;
    return QSV_OK;
}

int rowslab_run(qsv_ctx* ctx, RowSlabPlan* plan, void* state, cudaStream_t stream) {
    const size_t amp = amp_size(ctx->dtype);
    const int smem = (int)(2 * ROWLEVEL_STAGE_OPS * sizeof(RowOp) + plan->dim * plan->cols * amp);
    const unsigned grid = (unsigned)(plan->batch / plan->cols);
    if (plan->n_ops) {
        QSV_CUDA(ctx, ctx->dtype == QSV_C128
                          ? rowlevels_dispatch<double>(false, plan->cols, smem, grid, plan->threads, stream, state, plan->dim, plan->batch,
                                                       plan->d_ops, plan->d_level_start, (uint32_t)plan->n_levels)
                          : rowlevels_dispatch<float>(false, plan->cols, smem, grid, plan->threads, stream, state, plan->dim, plan->batch,
                                                      plan->d_ops, plan->d_level_start, (uint32_t)plan->n_levels));
        ctx->stats.kernel_launches++;
    }
    ctx->stats.sweeps++;
    ctx->stats.fused_sweeps++;
    ctx->stats.amp_bytes += 2 * amp * plan->dim * plan->batch;
    ctx->stats.gates_applied += plan->n_source_gates;
    return QSV_OK;
}
