// Host check of the fusion pass: the fused plan of a gate program interpreted on the host (include/qsv_hostcheck.h).
// TEST INFRASTRUCTURE -- not a compute path, never called by the package.
#include <stdlib.h>

#include <algorithm>

#include "qsv_sweep_ops.cuh"

// ================================================================================================
// Host check of the fusion pass (TEST INFRASTRUCTURE, see include/qsv_hostcheck.h)
// ================================================================================================
// Interprets a fused plan on the host, thread by thread, with the SAME op functions the kernel is compiled from
// (generic_op / fstage_op are __host__ __device__) and the same address tables (swizzled tile slots, per-slot
// global offsets, per-tile scalars).  It exists so that the planner -- stage selection, phase polynomials,
// relabelling, budgets, FFT-shape conversion -- can be fuzzed against the oracle on a machine without a GPU.
// It is not a compute path: nothing in the package calls it, and it refuses states of more than 2^22 amplitudes.
// The radix-8 / TMA image (qsv_sweep8.cu), restated: the tile copies follow the TmaTile descriptions the way the engine
// does (box coordinates from the tile id, 128-byte rows, SWIZZLE_128B slots), the stages run fstage8_compute /
// fstage8_put thread by thread, and a permuting launch re-orders the tile before the store like the kernel's last stage.
template <typename R>
static int sweep8_emulate_host(qsv_ctx* ctx, const SweepParams& P, const Sweep8Extra& X, const std::vector<double>& auxv,
                               const typename cplx_of<R>::type* src, typename cplx_of<R>::type* dst, bool thread_store) {
    typedef typename cplx_of<R>::type C;
    const SweepDesc& D = P.d;
    const int k = D.k;
    if (sizeof(C) != 16) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: the radix-8 image is complex128 only");
    if (D.n_ops > MAX_OPS8 || D.n_gslots > MAX_GSLOT8) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: %d ops / %d scalars in a radix-8 launch", D.n_ops, D.n_gslots);
    const uint32_t tile_amps = 1u << k, threads = 1u << (k - RB8);
    const uint64_t n_tiles = 1ull << (D.n_bits - k);
    const double2* aux2 = reinterpret_cast<const double2*>(auxv.data());
    std::vector<C> tile_v(tile_amps);
    C* tile = tile_v.data();
    std::vector<double2> smE(MAX_GSLOT8), smT(MAX_TAB);
    for (uint32_t i = 0; i < D.tab_len; ++i) smT[i] = aux2[D.tab_off + i];
    auto engine_copy = [&](const TmaTile& T, uint64_t phys_tile, bool load, const C* g_in, C* g_out) {
        int box_bits[5] = {3, 0, 0, 0, 0};
        for (int d = 1; d < T.rank; ++d) box_bits[d] = __builtin_ctz(T.box[d]);
        for (int e = 0; e < T.n_inst; ++e) {
            const uint64_t p = phys_tile | T.inst_goff[e];
            uint64_t c[5] = {0, 0, 0, 0, 0};
            for (int d = 1; d < T.rank; ++d) c[d] = (p >> T.shift[d]) & ((1ull << T.bits[d]) - 1ull);
            for (uint32_t l = 0; l < T.inst_amps; ++l) {
                // box element l (dimension 0 fastest) -> byte address in the tensor = sum coordinate x stride
                uint64_t byte = (uint64_t)(l & 7u) * 16;
                uint32_t rest = l >> 3;
                for (int d = 1; d < T.rank; ++d) {
                    const uint64_t x = c[d] + (rest & (T.box[d] - 1u));
                    if (x >= T.size[d]) return false;
                    byte += x * T.stride_bytes[d];
                    rest >>= box_bits[d];
                }
                const uint32_t slot = swz128(e * T.inst_amps + l);
                if (load) tile[slot] = g_in[byte / 16];
                else g_out[byte / 16] = tile[slot];
            }
        }
        return true;
    };
    std::vector<C> regs((size_t)threads * NREG8);
    for (uint64_t t = 0; t < n_tiles; ++t) {
        for (uint32_t gs = 0; gs < (uint32_t)D.n_gslots; ++gs) {
            const double2* gt = aux2 + D.gtab_off + (size_t)gs * D.g_chunks * 16;
            double2 e = gt[t & 15ull];
            for (int c = 1; c < D.g_chunks; ++c) e = cmul(e, gt[c * 16 + ((t >> (4 * c)) & 15ull)]);
            smE[gs] = e;
        }
        if (!engine_copy(X.in, seg_apply64x(t, D.in_glob), true, src, nullptr)) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: TMA load box leaves the tensor");
        if (X.mid_layout) {
            // two-stage permuting launch through the intermediate swz() layout: each stage computes everything, then writes
            if (D.n_ops != 2) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: mid layout with %d ops", D.n_ops);
            const OpView op0{&P.ops[0]}, op1{&P.ops[1]};
            for (uint32_t tid = 0; tid < threads; ++tid) {
                const uint32_t bl = expand_rb8(tid, op0.rpos(0), op0.rpos(1), op0.rpos(2));
                C w[NREG8];
                fstage8_compute<C, R>(op0, tile, smT.data(), smE.data(), tid, swz128(bl), op0.boff(0), op0.boff(1), op0.boff(2), w);
                for (int j = 0; j < NREG8; ++j) regs[(size_t)tid * NREG8 + j] = w[j];
            }
            for (uint32_t tid = 0; tid < threads; ++tid) {
                const uint32_t bl = expand_rb8(tid, op0.rpos(0), op0.rpos(1), op0.rpos(2));
                C w[NREG8];
                for (int j = 0; j < NREG8; ++j) w[j] = regs[(size_t)tid * NREG8 + j];
                fstage8_put<C>(tile, swz(bl), X.boff_mid_w[0], X.boff_mid_w[1], X.boff_mid_w[2], w);
            }
            for (uint32_t tid = 0; tid < threads; ++tid) {
                const uint32_t bl1 = seg_apply(tid, X.tmap_last);
                C w[NREG8];
                const uint32_t tid_std = compress_rb8(bl1, op1.rpos(0), op1.rpos(1), op1.rpos(2));
                fstage8_compute<C, R>(op1, tile, smT.data(), smE.data(), tid_std, swz(bl1), X.boff_last_r[0], X.boff_last_r[1], X.boff_last_r[2], w);
                for (int j = 0; j < NREG8; ++j) regs[(size_t)tid * NREG8 + j] = w[j];
            }
            for (uint32_t tid = 0; tid < threads; ++tid) {
                const uint32_t bl1 = seg_apply(tid, X.tmap_last);
                C w[NREG8];
                for (int j = 0; j < NREG8; ++j) w[j] = regs[(size_t)tid * NREG8 + j];
                fstage8_put<C>(tile, swz128(seg_apply(bl1, X.loc_perm)), X.boff_out[0], X.boff_out[1], X.boff_out[2], w);
            }
        }
        const bool generic = D.n_ops > 0 && P.ops[0].type == DC_STAGE;
        for (int i = 0; generic && i < D.n_ops;) {
            // a launch that is not FFT-shaped: DC_STAGE, then the ops it holds in registers (threads in sequence: in place)
            const OpView sop{&P.ops[i]};
            if (sop.type() != DC_STAGE) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: op %d of a generic radix-8 launch should open a stage", i);
            int j = i + 1;
            while (j < D.n_ops && P.ops[j].type != DC_STAGE) ++j;
            for (uint32_t tid = 0; tid < threads; ++tid) {
                const uint32_t bl = expand_rb8(tid, sop.st_rpos(0), sop.st_rpos(1), sop.st_rpos(2));
                const uint32_t a0 = swz128(bl), b0 = sop.st_roffsw(1), b1 = sop.st_roffsw(2), b2 = sop.st_roffsw(4);
                uint32_t a[NREG8];
                slots8(a, a0, b0, b1, b2);
                C w[NREG8];
                for (int r = 0; r < NREG8; ++r) w[r] = tile[a[r]];
                for (int q = i + 1; q < j; ++q) {
                    const OpView op{&P.ops[q]};
                    if (op.type() >= DC_FSTAGE && op.type() < DC_FIRST_FULL) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: OP_FSTAGE in a generic launch");
                    if (op.gmask() && (t & op.gmask()) != op.gval()) continue;
                    generic_op8<C, R>(w, op, op.type(), (bl & op.lmask()) == op.lval(), smT.data(), smE.data(), tid);
                }
                fstage8_put<C>(tile, a0, b0, b1, b2, w);
            }
            i = j;
        }
        for (int i = 0; i < D.n_ops && !X.mid_layout && !generic; ++i) {
            const OpView op{&P.ops[i]};
            if (op.type() < DC_FSTAGE || op.type() >= DC_FSTAGE + 8) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: op %d of a radix-8 launch is not a radix-8 OP_FSTAGE", i);
            const bool reorder = i == D.n_ops - 1 && X.permuting;
            // (a stage reads all its slots before it writes them, and threads own disjoint slots: threads in sequence are exact;
            // the re-ordering stage writes other threads' slots, so it computes everything first -- the kernel's extra barrier)
            for (uint32_t tid = 0; tid < threads; ++tid) {
                const uint32_t bl = expand_rb8(tid, op.rpos(0), op.rpos(1), op.rpos(2));
                C w[NREG8];
                fstage8_compute<C, R>(op, tile, smT.data(), smE.data(), tid, swz128(bl), op.boff(0), op.boff(1), op.boff(2), w);
                if (!reorder) fstage8_put<C>(tile, swz128(bl), op.boff(0), op.boff(1), op.boff(2), w);
                else for (int j = 0; j < NREG8; ++j) regs[(size_t)tid * NREG8 + j] = w[j];
            }
            if (reorder)
                for (uint32_t tid = 0; tid < threads; ++tid) {
                    const uint32_t bl = expand_rb8(tid, op.rpos(0), op.rpos(1), op.rpos(2));
                    C w[NREG8];
                    for (int j = 0; j < NREG8; ++j) w[j] = regs[(size_t)tid * NREG8 + j];
                    fstage8_put<C>(tile, swz128(seg_apply(bl, X.loc_perm)), X.boff_out[0], X.boff_out[1], X.boff_out[2], w);
                }
        }
        if (thread_store) {
            // an exchange-carrying launch: the threads store the output-ordered tile themselves (k_sweep8, D.xg branch) --
            // slot tid + (j << tb) goes to the tile's base plus the output offset of every set bit of the slot number
            const int tb = k - RB8;
            const uint64_t gb = seg_apply64x(t, D.out_glob);
            for (uint32_t tid = 0; tid < threads; ++tid) {
                uint64_t off = 0;
                for (int u = 0; u < tb; ++u)
                    if ((tid >> u) & 1u) off |= 1ull << X.out_pos[u];
                for (int j = 0; j < NREG8; ++j) {
                    uint64_t y = gb | off;
                    for (int q = 0; q < RB8; ++q)
                        if ((j >> q) & 1) y |= 1ull << X.out_pos[tb + q];
                    if (y >> D.n_bits) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: exchange store leaves the shard");
                    dst[y] = tile[swz128(tid) | ((uint32_t)j << tb)];
                }
            }
        } else if (!engine_copy(X.out, seg_apply64x(t, D.out_glob), false, nullptr, dst)) {
            return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: TMA store box leaves the tensor");
        }
    }
    return QSV_OK;
}

template <typename R>
static int sweep_emulate_host(qsv_ctx* ctx, const SweepParams& P, const std::vector<double>& auxv, int img,
                              const typename cplx_of<R>::type* src, typename cplx_of<R>::type* dst) {
    typedef typename cplx_of<R>::type C;
    const SweepDesc& D = P.d;
    if (D.xg) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "hostcheck: exchange sweeps are not emulated");
    const int k = D.k;
    const uint32_t tile_amps = 1u << k;
    const uint32_t threads = (uint32_t)std::max(32, 1 << (k - RB));  // as sweep_launch
    const uint32_t n_groups = tile_amps >> RB;
    const uint64_t n_tiles = 1ull << (D.n_bits - k);
    const double2* aux2 = reinterpret_cast<const double2*>(auxv.data());
    const bool dstore = D.direct_store;
    // (inactive threads of tiny tiles read beyond the tile like the device does beyond its buffer: give them room)
    std::vector<C> tile_v(std::max<size_t>(tile_amps, 1024));
    C* tile = tile_v.data();
    std::vector<double2> smE(MAX_GSLOT), smT(MAX_TAB);
    for (uint32_t i = 0; i < D.tab_len; ++i) smT[i] = aux2[D.tab_off + i];
    std::vector<C> regs((size_t)threads * NREG);
    std::vector<uint32_t> base_l(threads), swbase(threads);
    for (uint64_t t = 0; t < n_tiles; ++t) {
        for (uint32_t gs = 0; gs < (uint32_t)D.n_gslots; ++gs) {
            const double2* gt = aux2 + D.gtab_off + (size_t)gs * D.g_chunks * 16;
            double2 e = gt[t & 15ull];
            for (int c = 1; c < D.g_chunks; ++c) e = cmul(e, gt[c * 16 + ((t >> (4 * c)) & 15ull)]);
            smE[gs] = e;
        }
        const uint64_t gin = seg_apply64x(t, D.in_glob), gout = seg_apply64x(t, D.out_glob);
        auto n_slots_of = [&](uint32_t tid) { return (tid < tile_amps) ? (int)std::max(1u, tile_amps >> D.tb) : 0; };
        for (uint32_t tid = 0; tid < threads; ++tid) {
            const uint64_t gb = gin | seg_apply64(tid, D.in_loc);
            for (int i = 0; i < n_slots_of(tid); ++i) tile[swz(tid) ^ D.ld_soff[i]] = src[gb | D.ld_goff[i]];
        }
        auto st_parts = [&](uint32_t tid, uint64_t* t_st_g, uint32_t* t_st_s) {
            const uint32_t t_st_l = dstore ? expand_rb(tid, D.last_rpos) : seg_apply(tid, D.st_enum);
            *t_st_g = seg_apply64(t_st_l, D.out_loc);
            *t_st_s = swz(t_st_l);
        };
        bool stored = false, have_stage = false;
        uint64_t stq[4] = {0, 0, 0, 0};
        auto roff_of = [&](int j) { return (uint32_t)(stq[j >> 2] >> (16 * (j & 3))) & 0xFFFFu; };
        for (int i = 0; i < D.n_ops; ++i) {
            const OpR op = fetch_op(P.ops, i);
            const uint32_t type = op.type();
            if (img == IMG_FFT) {
                if (type < DC_FSTAGE || type >= DC_FIRST_FULL) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: op %d of an FFT launch is not OP_FSTAGE", i);
                const bool to_hbm = dstore && i == D.n_ops - 1;
                for (uint32_t tid = 0; tid < threads; ++tid) {
                    const bool active = tid < n_groups;
                    uint64_t t_st_g;
                    uint32_t t_st_s;
                    st_parts(tid, &t_st_g, &t_st_s);
                    const uint64_t gbs = gout | t_st_g;
#define FS(NH, HAS_W) fstage_op<C, R, NH, HAS_W>(op, tile, smT.data(), smE.data(), tid, active, to_hbm, dst, gbs, D.st_goff); break;
                    switch (type - DC_FSTAGE) {
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
                stored = to_hbm;
                continue;
            }
            if (type >= DC_FSTAGE && type < DC_FIRST_FULL) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: OP_FSTAGE in a generic launch");
            if (img == IMG_LEAN && type >= DC_FIRST_FULL) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: full-image op %u in a lean launch", type);
            if (type == DC_STAGE) {
                if (have_stage)
                    for (uint32_t tid = 0; tid < n_groups && tid < threads; ++tid)
                        for (int j = 0; j < NREG; ++j) tile[swbase[tid] ^ roff_of(j)] = regs[(size_t)tid * NREG + j];
                have_stage = true;
                stq[0] = op.p[0]; stq[1] = op.p[1]; stq[2] = op.p[2]; stq[3] = op.p[3];
                for (uint32_t tid = 0; tid < threads; ++tid) {
                    uint32_t b = tid;
                    for (int q = 0; q < RB; ++q) {
                        const uint32_t pz = op.rpos(q);
                        b = ((b >> pz) << (pz + 1)) | (b & ((1u << pz) - 1u));
                    }
                    base_l[tid] = b;
                    swbase[tid] = swz(b);
                    if (tid < n_groups)
                        for (int j = 0; j < NREG; ++j) regs[(size_t)tid * NREG + j] = tile[swbase[tid] ^ roff_of(j)];
                }
                continue;
            }
            if (!have_stage) return qsv_fail(ctx, QSV_E_INVALID, "hostcheck: op %d before any stage", i);
            if (op.gmask && (t & op.gmask) != op.gval) continue;
            for (uint32_t tid = 0; tid < n_groups && tid < threads; ++tid) {
                const bool tpred = (base_l[tid] & op.lmask()) == op.lval();
                C v[NREG];
                for (int j = 0; j < NREG; ++j) v[j] = regs[(size_t)tid * NREG + j];
                if (img == IMG_LEAN) generic_op<C, R, true>(v, op, type, tpred, smT.data(), smE.data(), tid);
                else generic_op<C, R, false>(v, op, type, tpred, smT.data(), smE.data(), tid);
                for (int j = 0; j < NREG; ++j) regs[(size_t)tid * NREG + j] = v[j];
            }
        }
        if (stored) continue;
        if (dstore && have_stage) {
            for (uint32_t tid = 0; tid < n_groups && tid < threads; ++tid) {
                uint64_t t_st_g;
                uint32_t t_st_s;
                st_parts(tid, &t_st_g, &t_st_s);
                for (int j = 0; j < NREG; ++j) dst[gout | t_st_g | D.st_goff[j]] = regs[(size_t)tid * NREG + j];
            }
            continue;
        }
        if (have_stage)
            for (uint32_t tid = 0; tid < n_groups && tid < threads; ++tid)
                for (int j = 0; j < NREG; ++j) tile[swbase[tid] ^ roff_of(j)] = regs[(size_t)tid * NREG + j];
        // (in place: every tile reads before it writes and tiles are disjoint, exactly as on the device)
        for (uint32_t tid = 0; tid < threads; ++tid) {
            uint64_t t_st_g;
            uint32_t t_st_s;
            st_parts(tid, &t_st_g, &t_st_s);
            for (int i = 0; i < n_slots_of(tid); ++i) dst[gout | t_st_g | D.st_goff[i]] = tile[t_st_s ^ D.st_soff[i]];
        }
    }
    return QSV_OK;
}

template <typename R>
static int hostcheck_run(qsv_ctx* ctx, SweepPlan* plan, void* state, uint64_t* out_sweeps, uint64_t* out_fft_sweeps, bool carry) {
    typedef typename cplx_of<R>::type C;
    const size_t amps = (size_t)1 << ctx->prog_bits;
    std::vector<C> spare_v(amps);
    C* cur = static_cast<C*>(state);
    C* other = spare_v.data();
    uint64_t sweeps = 0, fft = 0;
    for (auto& st : plan->steps) {
        if (!st.fused) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "hostcheck: the plan holds a gate that is not fused (dense-k / row pair)");
        C* d = st.out_of_place ? other : cur;
        int rc = st.img == IMG_FFT8 ? sweep8_emulate_host<R>(ctx, *st.params, *st.tma, plan->aux, cur, d, carry && &st == &plan->steps.back())
                                    : sweep_emulate_host<R>(ctx, *st.params, plan->aux, st.img, cur, d);
        if (rc) return rc;
        if (st.out_of_place) std::swap(cur, other);
        ++sweeps;
        fft += st.img == IMG_FFT || st.img == IMG_FFT8;
    }
    if (cur != static_cast<C*>(state)) memcpy(state, cur, amps * sizeof(C));
    if (out_sweeps) *out_sweeps = sweeps;
    if (out_fft_sweeps) *out_fft_sweeps = fft;
    return QSV_OK;
}

extern "C" int qsv_hostcheck_fused_layout(int dtype, int n_bits, const qsv_gate_t* gates, uint64_t n_gates, const double* pool,
                                          uint64_t pool_elems, int max_tile_bits, const int* final_pos, int carry_exchange, void* state,
                                          uint64_t* out_sweeps, uint64_t* out_fft_sweeps, char* err, uint64_t err_len);

extern "C" int qsv_hostcheck_fused(int dtype, int n_bits, const qsv_gate_t* gates, uint64_t n_gates, const double* pool, uint64_t pool_elems,
                                   int max_tile_bits, void* state, uint64_t* out_sweeps, uint64_t* out_fft_sweeps, char* err,
                                   uint64_t err_len) {
    return qsv_hostcheck_fused_layout(dtype, n_bits, gates, n_gates, pool, pool_elems, max_tile_bits, nullptr, 0, state, out_sweeps,
                                      out_fft_sweeps, err, err_len);
}

extern "C" int qsv_hostcheck_fused_layout(int dtype, int n_bits, const qsv_gate_t* gates, uint64_t n_gates, const double* pool,
                                          uint64_t pool_elems, int max_tile_bits, const int* final_pos, int carry_exchange, void* state,
                                          uint64_t* out_sweeps, uint64_t* out_fft_sweeps, char* err, uint64_t err_len) {
    qsv_ctx ctx;  // never touches a device
    ctx.dtype = dtype;
    int rc = QSV_OK;
    // state == NULL: plan only (any register the planner accepts) -- the shape of a full-size plan can be asserted on a CPU
    if ((dtype != QSV_C128 && dtype != QSV_C64) || n_bits < RB || n_bits > (state ? 22 : 40) || (n_gates && !gates)) {
        rc = qsv_fail(&ctx, QSV_E_INVALID, "hostcheck: bad arguments (n_bits must be %d..22, up to 40 without a state)", RB);
    } else {
        for (uint64_t i = 0; i < n_gates && !rc; ++i) {
            const qsv_gate_t& g = gates[i];
            if (g.kind > QSV_DIAG1 || g.n_targets != 1 || g.target_bits[0] >= n_bits || (g.ctrl_mask >> n_bits) ||
                (g.ctrl_mask & (1ull << g.target_bits[0])) || (g.ctrl_value & ~g.ctrl_mask) ||
                (g.kind != QSV_PERMX && g.matrix_offset + 4 > pool_elems))
                rc = qsv_fail(&ctx, QSV_E_INVALID, "hostcheck: gate %llu is not a valid one-target gate", (unsigned long long)i);
        }
    }
    SweepPlan* plan = nullptr;
    if (!rc) {
        ctx.gates.assign(gates, gates + n_gates);
        ctx.pool.assign(pool, pool + 2 * pool_elems);
        ctx.prog_bits = n_bits;
        ctx.has_program = true;
        rc = sweep_plan_build_host(&ctx, max_tile_bits, &plan, final_pos, carry_exchange != 0);
        if (!rc && carry_exchange && (plan->steps.empty() || !plan->steps.back().out_of_place))
            rc = qsv_fail(&ctx, QSV_E_INVALID, "hostcheck: the plan's last sweep cannot carry an exchange");
    }
    if (!rc && !state) {
        uint64_t sweeps = 0, fft = 0;
        for (auto& st : plan->steps) {
            sweeps += st.fused;
            fft += st.fused && (st.img == IMG_FFT || st.img == IMG_FFT8);
        }
        if (out_sweeps) *out_sweeps = sweeps;
        if (out_fft_sweeps) *out_fft_sweeps = fft;
    } else if (!rc) {
        rc = dtype == QSV_C128 ? hostcheck_run<double>(&ctx, plan, state, out_sweeps, out_fft_sweeps, carry_exchange != 0)
                               : hostcheck_run<float>(&ctx, plan, state, out_sweeps, out_fft_sweeps, carry_exchange != 0);
    }
    if (plan) sweep_plan_free(plan);
    if (err && err_len) {
        strncpy(err, ctx.err, err_len - 1);
        err[err_len - 1] = 0;
    }
    return rc;
}
