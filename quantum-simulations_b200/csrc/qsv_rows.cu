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
// Columns are independent, so lane t of the (single) warp of a CTA owns column t of the slab and never
// synchronises with other lanes; ops are staged into shared memory in double-buffered cp.async chunks.
// The row shared by consecutive ops (Givens rotations of one column of the QR sweep all touch the pivot
// row) stays in registers.  Throughput is bound by the dependent chain through that pivot row, not by
// HBM: the 16 MiB state of config 4 is read and written once.
#include <algorithm>

#include "qsv_common.cuh"

struct __align__(16) RowOp {  // 80 bytes
    uint32_t r0, r1;
    uint32_t swap;  // 1: exchange the two rows (kept arithmetic-free)
    uint32_t pad;
    double m[8];
};
static_assert(sizeof(RowOp) == 80, "RowOp layout");

#define ROW_CHUNK 64  // ops per staged chunk (5 KB)

struct RowSlabPlan {
    RowOp* d_ops = nullptr;
    uint64_t n_ops = 0;
    uint64_t n_source_gates = 0;
    uint64_t dim = 0, batch = 0;
    int cols = 0;  // columns per CTA
    ~RowSlabPlan() {
        if (d_ops) cudaFree(d_ops);
    }
};

template <typename R>
__global__ void __launch_bounds__(32, 1) k_rowslab(typename cplx_of<R>::type* __restrict__ st, uint64_t dim, uint64_t batch,
                                                   int cols, const RowOp* __restrict__ ops, uint64_t n_ops) {
    typedef typename cplx_of<R>::type C;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RowOp* opbuf = reinterpret_cast<RowOp*>(smem_raw);                        // 2 x ROW_CHUNK ops
    C* slab = reinterpret_cast<C*>(smem_raw + 2 * ROW_CHUNK * sizeof(RowOp));  // [dim][cols]
    const int lane = threadIdx.x;
    const uint64_t col0 = (uint64_t)blockIdx.x * cols;
    // load the slab (each row segment is cols x 16 B contiguous)
    for (uint64_t i = lane; i < dim * cols; i += 32) {
        const uint64_t r = i / cols, c = i % cols;
        slab[i] = st[r * batch + col0 + c];
    }
    auto stage = [&](uint64_t chunk, int buf) {
        const uint64_t first = chunk * ROW_CHUNK;
        const uint64_t n = min((uint64_t)ROW_CHUNK, n_ops - first);
        const char* src = reinterpret_cast<const char*>(ops + first);
        char* dst = reinterpret_cast<char*>(opbuf + buf * ROW_CHUNK);
        for (uint64_t b = (uint64_t)lane * 16; b < n * sizeof(RowOp); b += 32 * 16) {
            const uint32_t s = (uint32_t)__cvta_generic_to_shared(dst + b);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(src + b));
        }
        asm volatile("cp.async.commit_group;\n" ::);
    };
    const uint64_t n_chunks = (n_ops + ROW_CHUNK - 1) / ROW_CHUNK;
    if (n_chunks) stage(0, 0);
    __syncwarp();
    const bool active = lane < cols;
    uint32_t cached = 0xFFFFFFFFu;  // row whose value for this lane's column is held in `cv`
    C cv = cmake<C, R>((R)0, (R)0);
    for (uint64_t ch = 0; ch < n_chunks; ++ch) {
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncwarp();
        if (ch + 1 < n_chunks) stage(ch + 1, (int)((ch + 1) & 1));
        const RowOp* cur = opbuf + (ch & 1) * ROW_CHUNK;
        const int n = (int)min((uint64_t)ROW_CHUNK, n_ops - ch * ROW_CHUNK);
        if (active) {
            for (int i = 0; i < n; ++i) {
                const RowOp& op = cur[i];
                const uint32_t r0 = op.r0, r1 = op.r1;
                if (cached != r0) {
                    if (cached != 0xFFFFFFFFu) slab[(uint64_t)cached * cols + lane] = cv;
                    cv = slab[(uint64_t)r0 * cols + lane];
                    cached = r0;
                }
                // (r1 == cached cannot happen: r1 != r0 and cached == r0 here)
                const C a1 = slab[(uint64_t)r1 * cols + lane];
                C b1;
                if (op.swap) {
                    b1 = cv;
                    cv = a1;
                } else {
                    const C m00 = cmake<C, R>((R)op.m[0], (R)op.m[1]), m01 = cmake<C, R>((R)op.m[2], (R)op.m[3]);
                    const C m10 = cmake<C, R>((R)op.m[4], (R)op.m[5]), m11 = cmake<C, R>((R)op.m[6], (R)op.m[7]);
                    const C a0 = cv;
                    cv = cfma(m01, a1, cmul(m00, a0));
                    b1 = cfma(m11, a1, cmul(m10, a0));
                }
                slab[(uint64_t)r1 * cols + lane] = b1;
            }
        }
        __syncwarp();
    }
    if (active && cached != 0xFFFFFFFFu) slab[(uint64_t)cached * cols + lane] = cv;
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncwarp();
    for (uint64_t i = lane; i < dim * cols; i += 32) {
        const uint64_t r = i / cols, c = i % cols;
        st[r * batch + col0 + c] = slab[i];
    }
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
    const size_t budget = 200 * 1024 - 2 * ROW_CHUNK * sizeof(RowOp);
    uint64_t cols = budget / (dim * amp);
    if (cols < 1) return QSV_E_UNSUPPORTED;
    cols = std::min<uint64_t>(std::min<uint64_t>(cols, 32), batch);
    while (cols & (cols - 1)) cols &= cols - 1;  // power of two
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
    RowSlabPlan* plan = new RowSlabPlan();
    plan->n_ops = ops.size();
    plan->n_source_gates = ctx->gates.size();
    plan->dim = dim;
    plan->batch = batch;
    plan->cols = (int)cols;
    if (!ops.empty()) {
        cudaError_t e = cudaMalloc(&plan->d_ops, ops.size() * sizeof(RowOp));
        if (e == cudaSuccess) e = cudaMemcpy(plan->d_ops, ops.data(), ops.size() * sizeof(RowOp), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            delete plan;
            return qsv_fail(ctx, QSV_E_CUDA, "uploading the row-op stream failed: %s", cudaGetErrorString(e));
        }
    }
    const int smem = (int)(2 * ROW_CHUNK * sizeof(RowOp) + dim * cols * amp);
    cudaError_t e = ctx->dtype == QSV_C128 ? cudaFuncSetAttribute(k_rowslab<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                                           : cudaFuncSetAttribute(k_rowslab<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) {
        delete plan;
        return qsv_fail(ctx, QSV_E_CUDA, "cudaFuncSetAttribute(%d B smem) failed: %s", smem, cudaGetErrorString(e));
    }
    *out = plan;
    return QSV_OK;
}

int rowslab_run(qsv_ctx* ctx, RowSlabPlan* plan, void* state, cudaStream_t stream) {
    const size_t amp = amp_size(ctx->dtype);
    const size_t smem = 2 * ROW_CHUNK * sizeof(RowOp) + plan->dim * plan->cols * amp;
    const unsigned grid = (unsigned)(plan->batch / plan->cols);
    if (ctx->dtype == QSV_C128)
        k_rowslab<double><<<grid, 32, smem, stream>>>((double2*)state, plan->dim, plan->batch, plan->cols, plan->d_ops, plan->n_ops);
    else
        k_rowslab<float><<<grid, 32, smem, stream>>>((float2*)state, plan->dim, plan->batch, plan->cols, plan->d_ops, plan->n_ops);
    QSV_CUDA(ctx, cudaGetLastError());
    ctx->stats.kernel_launches++;
    ctx->stats.sweeps++;
    ctx->stats.fused_sweeps++;
    ctx->stats.amp_bytes += 2 * amp * plan->dim * plan->batch;
    ctx->stats.gates_applied += plan->n_source_gates;
    return QSV_OK;
}
