// Fused tile sweep ("F" in SURVEY.md 2.1): gate-fusion pass + shared-memory-staged sweep kernel.
//
// One launch = one pass over the state ("sweep").  The state index is split into k LOCAL bits (the
// tile, 2^k amplitudes staged in shared memory) and n-k GLOBAL bits (the tile id).  Every gate whose
// target bit is local can be applied inside the tile; controls may sit anywhere (local controls are
// tested per amplitude, global controls per tile); diagonal gates can always be applied because
// they only need the amplitude's own index.  HBM traffic of a sweep is one read + one write of the
// state no matter how many gates it carries.
//
// Inside a tile the work proceeds in REGISTER STAGES: each thread holds the 16 amplitudes that
// differ in 4 chosen local bits (the stage's register bits), applies every op of the stage to them
// in registers, and the tile is re-distributed through shared memory between stages (like the
// passes of a radix-16 FFT).  Shared memory is XOR-swizzled at 16-byte granularity so that the
// 128-bit accesses of a quarter-warp hit 8 distinct bank groups for the usual register-bit choices.
//
// Runs of consecutive diagonal gates with at most one control are merged on the host into a phase
// polynomial  phi(x) = t0 + sum_a A_a x_a + sum_{a<b} B_ab x_a x_b  and applied as ONE op:
//   register-register terms -> a 16-entry table W[j]                              (shared memory)
//   register-thread terms   -> two 16-entry tables over 4-bit chunks of the thread id (shared memory,
//                              copied once per CTA), multiplied together
//   register-global terms   -> per-tile scalars (one sincos per slot per tile, kept in shared memory)
// so a QFT sweep with 6 Hadamards and ~150 controlled phases costs 6 butterflies + a few table lookups.
//
// A sweep may store the tile to a bit-permuted location of a second buffer (out of place): SWAP
// gates then cost no arithmetic and no extra traffic.
//
// Launches whose every stage is "FFT shaped" (butterflies, each followed by phases that couple its bit
// only to bits whose butterfly comes later, then the deferred phases) are rewritten by the host into
// ONE op per stage (OP_FSTAGE) and run in a kernel image of their own (IMG_FFT) in which no amplitude
// is carried in registers from one op to the next -- see fstage_math / to_fstages.  Everything else
// runs in the generic interpreter images (IMG_LEAN, IMG_FULL).
//
// The per-op arithmetic (generic_op, fstage_op and below) is __host__ __device__: the same source is
// (qsv_sweep_ops.cuh) is interpreted on the host by qsv_hostcheck_fused (qsv_hostcheck.cu, test infrastructure)
// so that the fusion pass can be checked against the oracle without a GPU.
#include <stdlib.h>

#include <algorithm>
#include <cmath>
#include <complex>
#include <map>

#include "qsv_common.cuh"
#include "qsv_sweep_ops.cuh"

// gridDim.x re-read at every use: held in a register across a tile it was spilled to local memory
__device__ __forceinline__ uint32_t grid_x() {
    uint32_t gx;
    asm volatile("mov.u32 %0, %%nctaid.x;" : "=r"(gx));
    return gx;
}

template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gsrc) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
    if (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gsrc));
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

// Two instantiations: <256, 1> for 64 KB tiles (k = 12, one CTA per SM) and <128, 3> for smaller tiles
// (register cap 168 so that three double-buffered 32 KB-tile CTAs share an SM).
template <typename R, int MAXT, int MINB, int IMG>
// (src and dst are NOT __restrict__: every in-place sweep is launched with dst == src)
__global__ void __launch_bounds__(MAXT, MINB) k_sweep(const typename cplx_of<R>::type* src,
                                                  typename cplx_of<R>::type* dst,
                                                  const double* __restrict__ aux,
                                                  const __grid_constant__ SweepParams P) {
    typedef typename cplx_of<R>::type C;
    constexpr bool LEAN = IMG != IMG_FULL;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const SweepDesc& D = P.d;
    const size_t tile_bytes = (size_t)sizeof(C) << D.k;
    auto buf_ptr = [&](int b) { return reinterpret_cast<C*>(smem_raw + (size_t)b * tile_bytes); };
    double2* smE0 = reinterpret_cast<double2*>(smem_raw + 2 * tile_bytes);  // [2][MAX_GSLOT]: per-tile scalars, by tile parity
    double2* smT = smE0 + 2 * MAX_GSLOT;
    uint64_t* park = reinterpret_cast<uint64_t*>(smT + MAX_TAB);  // [PARK_WORDS][blockDim]
    const double2* aux2 = reinterpret_cast<const double2*>(aux);

    const uint32_t tid = threadIdx.x;
    const uint32_t tile_amps = 1u << D.k;
    const uint32_t n_groups = tile_amps >> RB;
    const bool active = tid < n_groups;
    const uint32_t n_tiles = 1u << (D.n_bits - D.k);  // (at most 2^31: checked by sweep_launch)
    // per-thread parts of the tile-copy addresses (tile independent)
    const bool dstore = D.direct_store;
    // (the two 64-bit ones live in shared memory and are re-read where needed -- one LDS each per tile: kept in
    // registers they were spilled to local memory, whose reloads were the kernel's largest single stall)
    const uint32_t t_st_l = dstore ? expand_rb(tid, D.last_rpos) : seg_apply(tid, D.st_enum);
    volatile uint64_t* park_ld = park + tid;
    volatile uint64_t* park_st = park + blockDim.x + tid;
    volatile uint32_t* park_w = reinterpret_cast<volatile uint32_t*>(park + 2 * blockDim.x + tid);  // [0] store slot, [1] tile id
    *park_ld = seg_apply64(tid, D.in_loc);
    *park_st = seg_apply64(t_st_l, D.out_loc);
    park_w[0] = swz(t_st_l);
#define t_ld_g (*park_ld)
#define t_st_g (*park_st)
#define t_st_s (park_w[0])
    // amplitudes each thread copies (<= NREG); tiles smaller than a warp use only the first tile_amps threads
    // (kept as a CTA-uniform count and a per-thread predicate rather than one per-thread integer: registers are scarce)
    const int n_slots_u = (int)max(1u, tile_amps >> D.tb);
    const bool copies = tid < tile_amps;
#define n_slots (copies ? n_slots_u : 0)

    // phase tables of this sweep: global -> shared, once per CTA
    for (uint32_t i = tid; i < D.tab_len; i += blockDim.x) smT[i] = aux2[D.tab_off + i];

    // tile prefetch: 128-bit cp.async straight into the swizzled tile buffer (no registers involved)
    auto prefetch = [&](uint64_t t, C* buf) {
        const uint64_t gb = seg_apply64x(t, D.in_glob) | t_ld_g;
        const uint32_t t_ld_s = swz(tid);
#pragma unroll
        for (int i = 0; i < NREG; ++i)
            if (i < n_slots) cp_async<sizeof(C)>(&buf[t_ld_s ^ D.ld_soff[i]], &src[gb | D.ld_goff[i]]);
    };

    int cur = 0;
    if (blockIdx.x < n_tiles) prefetch(blockIdx.x, buf_ptr(0));
    cp_async_commit();

    for (uint32_t t = blockIdx.x; t < n_tiles; t += grid_x()) {
        C* tile = buf_ptr(cur);
        double2* smE = smE0 + cur * MAX_GSLOT;
        // ---- per-tile phase scalars: product of 16-entry chunk tables over the tile id ---------
        for (uint32_t gs = tid; gs < (uint32_t)D.n_gslots; gs += blockDim.x) {
            const double2* gt = aux2 + D.gtab_off + (size_t)gs * D.g_chunks * 16;
            double2 e = gt[t & 15ull];
            for (int c = 1; c < D.g_chunks; ++c) e = cmul(e, gt[c * 16 + ((t >> (4 * c)) & 15ull)]);
            smE[gs] = e;
        }
        cp_async_wait_all();
        __syncthreads();  // tile t has landed; per-tile scalars visible; previous tile fully retired
        {
            const uint32_t tn = t + grid_x();
            if (tn > t && tn < n_tiles) prefetch(tn, buf_ptr(cur ^ 1));
            cp_async_commit();
        }

        if constexpr (IMG == IMG_FFT) {
            // every op is a whole stage in one straight line: shared memory -> registers -> butterflies + phases ->
            // shared memory (or, for the launch's last stage, straight to HBM)
            bool stored = false;
            const uint64_t gbs = seg_apply64x(t, D.out_glob) | t_st_g;
            for (int i = 0; i < D.n_ops; ++i) {
                const OpR op = fetch_op(P.ops, i);
                const bool to_hbm = dstore && i == D.n_ops - 1;
#define FS(NH, HAS_W) stored = fstage_op<C, R, NH, HAS_W>(op, tile, smT, smE, tid, active, to_hbm, dst, gbs, D.st_goff); break;
                switch (op.type() - DC_FSTAGE) {  // shape = 2 * (number of butterflies) + (phase tables present)
                    case 0: case 1: FS(0, false)
                    case 2: FS(1, false)
                    case 3: FS(1, true)
                    case 4: FS(2, false)
                    case 5: FS(2, true)
                    case 6: FS(3, false)
                    case 7: FS(3, true)
                    case 8: FS(4, false)
                    default: FS(4, true)
                }
#undef FS
            }
            if (!stored) {
                const uint64_t gb = gbs;
#pragma unroll
                for (int i = 0; i < NREG; ++i)
                    if (i < n_slots) {
                        const uint64_t y = gb | D.st_goff[i];
                        if (D.xg) {  // the store IS the exchange: top xg bits of the output index -> destination rank
                            const int sh = D.n_bits - D.xg;
                            C* pd = reinterpret_cast<C*>(D.peer[y >> sh]);
                            pd[((uint64_t)D.xrank << sh) | (y & ((1ull << sh) - 1ull))] = tile[t_st_s ^ D.st_soff[i]];
                        } else {
                            dst[y] = tile[t_st_s ^ D.st_soff[i]];
                        }
                    }
            }
        } else {
        C v[NREG];
        uint32_t base_l = 0, swbase = 0;
        bool have_stage = false;
        uint64_t stq[4] = {0, 0, 0, 0};  // roffsw table of the open stage
        auto roff_of = [&](int j) { return (uint32_t)(stq[j >> 2] >> (16 * (j & 3))) & 0xFFFFu; };
        // each op is read from the constant bank into registers in one go (fetching one op ahead was measured: no gain,
        // the copy costs as much as the latency it hides)
        for (int i = 0; i < D.n_ops; ++i) {
            const OpR op = fetch_op(P.ops, i);
            const uint32_t type = op.type();
            if (type == DC_STAGE) {
                if (have_stage) {
                    if (active) {
#pragma unroll
                        for (int j = 0; j < NREG; ++j) tile[swbase ^ roff_of(j)] = v[j];
                    }
                    __syncthreads();
                }
                have_stage = true;
                stq[0] = op.p[0]; stq[1] = op.p[1]; stq[2] = op.p[2]; stq[3] = op.p[3];
                base_l = tid;
#pragma unroll
                for (int q = 0; q < RB; ++q) {
                    const uint32_t pz = op.rpos(q);
                    base_l = ((base_l >> pz) << (pz + 1)) | (base_l & ((1u << pz) - 1u));
                }
                swbase = swz(base_l);
                if (active) {
#pragma unroll
                    for (int j = 0; j < NREG; ++j) v[j] = tile[swbase ^ roff_of(j)];
                }
                continue;
            }
            // (threads beyond the tile's groups -- tiles smaller than a warp's worth -- run the math on garbage and never store it)
            bool tpred = true;
            if constexpr (!LEAN) {
                if (op.gmask && (t & op.gmask) != op.gval) continue;  // control on a tile-id bit not satisfied for this tile
                tpred = (base_l & op.lmask()) == op.lval();
            }
            generic_op<C, R, LEAN>(v, op, type, tpred, smT, smE, tid);
        }
        // ---- store the tile (possibly to a bit-permuted location of the other buffer) ----------
        const uint64_t gb = seg_apply64x(t, D.out_glob) | t_st_g;
        if (dstore && have_stage) {
            if (active) {
#pragma unroll
                for (int j = 0; j < NREG; ++j) dst[gb | D.st_goff[j]] = v[j];
            }
        } else {
            if (have_stage) {
                if (active) {
#pragma unroll
                    for (int j = 0; j < NREG; ++j) tile[swbase ^ roff_of(j)] = v[j];
                }
                __syncthreads();
            }
#pragma unroll
            for (int i = 0; i < NREG; ++i)
                if (i < n_slots) v[i] = tile[t_st_s ^ D.st_soff[i]];
#pragma unroll
            for (int i = 0; i < NREG; ++i)
                if (i < n_slots) {
                    const uint64_t y = gb | D.st_goff[i];
                    if (D.xg) {
                        const int sh = D.n_bits - D.xg;
                        C* pd = reinterpret_cast<C*>(D.peer[y >> sh]);
                        pd[((uint64_t)D.xrank << sh) | (y & ((1ull << sh) - 1ull))] = v[i];
                    } else {
                        dst[y] = v[i];
                    }
                }
        }
        }  // generic images
        // No barrier here: the tile buffer and the scalars of this parity are next written two tiles from now, i.e. after
        // every thread has passed the barrier at the start of the next tile -- and so has finished this one.
        cur ^= 1;
    }
    cp_async_wait_all();
}
#undef t_ld_g
#undef t_st_g
#undef t_st_s
#undef n_slots

// ================================================================================================
// host: plan upload, launch, run (the fusion pass itself is in qsv_fuse.cu)
// ================================================================================================
void sweep_plan_free(SweepPlan* p) { delete p; }
void sweep_plan_count(const SweepPlan* plan, uint64_t* sweeps, uint64_t* out_of_place) {
    uint64_t s = 0, o = 0;
    for (auto& st : plan->steps) {
        s += st.fused;
        o += st.fused && st.out_of_place;
    }
    *sweeps = s;
    *out_of_place = o;
}


static int sweep_set_smem_attr(qsv_ctx* ctx, int k);
// qsv_sweep8.cu: the radix-8 / TMA image
int sweep8_prepare(qsv_ctx* ctx, int k);
int sweep8_launch(qsv_ctx* ctx, int n, int k, const void* src, void* dst, const double* d_aux, const PlanStep& st, cudaStream_t stream, int xg,
                  int xrank, const uint64_t* peers);

int sweep_plan_build(qsv_ctx* ctx, int max_tile_bits, uint64_t* out_sweeps, const int* final_pos, bool carry_exchange) {
    SweepPlan* plan = nullptr;
    {
        int rc = sweep_plan_build_host(ctx, max_tile_bits, &plan, final_pos, carry_exchange);
        if (rc) return rc;
    }
    const int k = plan->k;
    cudaError_t e = cudaMalloc(&plan->d_aux, plan->aux.size() * sizeof(double));
    if (e == cudaSuccess) e = cudaMemcpy(plan->d_aux, plan->aux.data(), plan->aux.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        delete plan;
        return qsv_fail(ctx, QSV_E_CUDA, "uploading the sweep tables failed: %s", cudaGetErrorString(e));
    }
    // opt in to > 48 KB of dynamic shared memory
    {
        int rc2 = sweep_set_smem_attr(ctx, k);
        if (rc2) {
            delete plan;
            return rc2;
        }
    }
    for (auto& st : plan->steps)
        if (st.fused && st.img == IMG_FFT8) {
            int rc3 = sweep8_prepare(ctx, k);
            if (rc3) {
                delete plan;
                return rc3;
            }
            break;
        }
    ctx->plan = plan;
    if (out_sweeps) {
        uint64_t s = 0;
        for (auto& st : plan->steps) s += st.fused;
        *out_sweeps = s;
    }
    return QSV_OK;
}

// defined in qsv.cu
int qsv_apply_planned_single(qsv_ctx* ctx, void* state, size_t gate_index, cudaStream_t s);

template <typename R, int MAXT, int MINB, int IMG>
static cudaError_t sweep_image_launch(bool set_attr_only, int* occ_out, unsigned grid, int threads, size_t smem, cudaStream_t stream,
                                      const void* src, void* dst, const double* d_aux, const SweepParams* params) {
    typedef typename cplx_of<R>::type C;
    auto kern = k_sweep<R, MAXT, MINB, IMG>;
    if (set_attr_only) return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (occ_out) return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ_out, kern, threads, smem);
    kern<<<grid, threads, smem, stream>>>((const C*)src, (C*)dst, d_aux, *params);
    return cudaGetLastError();
}

// picks the kernel image for (dtype, tile size, image kind) and either sets its smem attribute, queries occupancy or launches it
static cudaError_t sweep_image(qsv_ctx* ctx, int k, int img, bool set_attr_only, int* occ_out, unsigned grid, int threads, size_t smem,
                               cudaStream_t stream, const void* src, void* dst, const double* d_aux, const SweepParams* params) {
#define LAUNCH(R, MAXT, MINB, I) sweep_image_launch<R, MAXT, MINB, I>(set_attr_only, occ_out, grid, threads, smem, stream, src, dst, d_aux, params)
#define BY_IMG(R, MAXT, MINB) \
    (img == IMG_FFT ? LAUNCH(R, MAXT, MINB, IMG_FFT) : img == IMG_LEAN ? LAUNCH(R, MAXT, MINB, IMG_LEAN) : LAUNCH(R, MAXT, MINB, IMG_FULL))
    const bool dbl = ctx->dtype == QSV_C128;
    if (k > 11) return dbl ? BY_IMG(double, 256, 1) : BY_IMG(float, 256, 1);
    return dbl ? BY_IMG(double, 128, 3) : BY_IMG(float, 128, 3);
#undef BY_IMG
#undef LAUNCH
}

static int sweep_set_smem_attr(qsv_ctx* ctx, int k) {
    const size_t smem = SWEEP_SMEM(amp_size(ctx->dtype), k, std::max(32, 1 << (k - RB)));
    for (int img = 0; img < 3; ++img) {
        cudaError_t e = sweep_image(ctx, k, img, true, nullptr, 0, 0, smem, nullptr, nullptr, nullptr, nullptr, nullptr);
        if (e != cudaSuccess) return qsv_fail(ctx, QSV_E_CUDA, "cudaFuncSetAttribute(%zu B smem) failed: %s", smem, cudaGetErrorString(e));
    }
    return QSV_OK;
}

static int sweep_launch(qsv_ctx* ctx, int n, int k, const void* src, void* dst, const double* d_aux, const SweepParams& params,
                        cudaStream_t stream, int img = IMG_FULL) {
    int sm_count = 148;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, ctx->device);
    const int threads = std::max(32, 1 << (k - RB));
    const size_t smem = SWEEP_SMEM(amp_size(ctx->dtype), k, threads);
    if (n - k > 31) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "unsupported: %d tile-id bits (max 31)", n - k);
    const uint64_t n_tiles = 1ull << (n - k);
    int occ = 1;
    sweep_image(ctx, k, img, false, &occ, 0, threads, smem, stream, nullptr, nullptr, nullptr, nullptr);
    occ = std::max(occ, 1);
    const unsigned grid = (unsigned)std::min<uint64_t>(n_tiles, (uint64_t)sm_count * occ);
    QSV_CUDA(ctx, sweep_image(ctx, k, img, false, nullptr, grid, threads, smem, stream, src, dst, d_aux, &params));
    return QSV_OK;
}

// Bit permutation as ONE gate-free tile sweep with a permuting store: out[y] = in[x], bit perm[i] of y = bit i of x.
// Returns QSV_E_UNSUPPORTED when the state is too small for the tile machinery (caller falls back to k_permute_bits).
int sweep_permute_run(qsv_ctx* ctx, const void* in, void* out, int n_bits, const int* perm, cudaStream_t stream, int xg, int xrank,
                      const uint64_t* peers) {
    if (n_bits < 9) return QSV_E_UNSUPPORTED;
    const int k = std::min(n_bits, 11);
    SweepPlan tmp;
    int rc = sweep_permute_plan_build(ctx, n_bits, perm, k, &tmp);
    if (rc) return rc;
    rc = sweep_set_smem_attr(ctx, k);
    if (rc) return rc;
    if (xg > 0) {
        SweepDesc& D = tmp.steps[0].params->d;
        D.xg = xg;
        D.xrank = xrank;
        for (int i = 0; i < (1 << xg); ++i) D.peer[i] = peers[i];
    }
    rc = sweep_launch(ctx, n_bits, k, in, out, nullptr, *tmp.steps[0].params, stream);
    if (rc) return rc;
    ctx->stats.fused_sweeps++;
    return QSV_OK;
}

int sweep_plan_run(qsv_ctx* ctx, void* state, void* spare, int* result_in_spare, cudaStream_t stream, int xg, int xrank,
                   const uint64_t* peers) {
    SweepPlan* plan = ctx->plan;
    const int n = ctx->prog_bits, k = plan->k;
    void* cur = state;
    void* other = spare;
    if (xg > 0) {
        if (plan->steps.empty() || !plan->steps.back().fused || !plan->steps.back().out_of_place)
            return qsv_fail(ctx, QSV_E_INVALID, "this plan was not built to carry an exchange (qsv_fuse_to_layout with carry_exchange)");
    }
    for (auto& st : plan->steps) {
        const bool carries = xg > 0 && &st == &plan->steps.back();
        SweepParams* patched = nullptr;
        if (carries && st.img != IMG_FFT8) {
            // the last sweep's store goes to the peers: same launch, destination table filled in
            patched = new SweepParams(*st.params);
            patched->d.xg = xg;
            patched->d.xrank = xrank;
            for (int i = 0; i < (1 << xg); ++i) patched->d.peer[i] = peers[i];
        }
        if (!st.fused) {
            int rc = qsv_apply_planned_single(ctx, cur, st.gate_index, stream);
            if (rc) return rc;
            continue;
        }
        void* dstp = cur;
        if (st.out_of_place) {
            if (!other) return qsv_fail(ctx, QSV_E_INVALID, "this plan has bit-permuting sweeps: pass a spare buffer to qsv_run");
            dstp = other;
        }
        cudaEvent_t ev0 = nullptr, ev1 = nullptr;
        if (ctx->profile) {
            QSV_CUDA(ctx, cudaEventCreate(&ev0));
            QSV_CUDA(ctx, cudaEventCreate(&ev1));
            QSV_CUDA(ctx, cudaEventRecord(ev0, stream));
        }
        {
            int rc = st.img == IMG_FFT8 ? sweep8_launch(ctx, n, k, cur, dstp, plan->d_aux, st, stream, carries ? xg : 0, xrank, peers)
                                        : sweep_launch(ctx, n, k, cur, dstp, plan->d_aux, patched ? *patched : *st.params, stream, st.img);
            delete patched;
            if (rc) return rc;
        }
        static const bool sync_each = getenv("QSV_SYNC") != nullptr;  // debugging aid: surface kernel faults at the launch site
        if (sync_each) QSV_CUDA(ctx, cudaStreamSynchronize(stream));
        if (ctx->profile) {
            QSV_CUDA(ctx, cudaEventRecord(ev1, stream));
            plan->ev.push_back(ev0);
            plan->ev.push_back(ev1);
        }
        ctx->stats.kernel_launches++;
        ctx->stats.sweeps++;
        ctx->stats.fused_sweeps++;
        ctx->stats.amp_bytes += 2 * (amp_size(ctx->dtype) << n);
        ctx->stats.gates_applied += st.n_gates;
        if (st.out_of_place) std::swap(cur, other);
    }
    ctx->stats.gates_applied += plan->dropped_identities;
    if (result_in_spare) *result_in_spare = (cur != state);
    else if (cur != state) return qsv_fail(ctx, QSV_E_INVALID, "result landed in the spare buffer but result_in_spare is NULL");
    return QSV_OK;
}

// Collects the CUDA-event durations of the fused sweeps launched since the last call (profiling mode).
int sweep_plan_collect_profile(qsv_ctx* ctx, float* ms, int cap, int* n_out) {
    SweepPlan* plan = ctx->plan;
    int n = 0;
    if (plan) {
        for (size_t i = 0; i + 1 < plan->ev.size(); i += 2) {
            QSV_CUDA(ctx, cudaEventSynchronize(plan->ev[i + 1]));
            float t = 0;
            QSV_CUDA(ctx, cudaEventElapsedTime(&t, plan->ev[i], plan->ev[i + 1]));
            if (n < cap && ms) ms[n] = t;
            ++n;
            cudaEventDestroy(plan->ev[i]);
            cudaEventDestroy(plan->ev[i + 1]);
        }
        plan->ev.clear();
    }
    if (n_out) *n_out = n;
    return QSV_OK;
}
