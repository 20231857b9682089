// Gate-fusion pass (host): packs the uploaded gate stream into tile sweeps for k_sweep (qsv_sweep.cu) -- tile and stage
// selection, phase polynomials and their tables, SWAP relabelling, launch budgets, FFT-shape conversion, balanced
// sweeps.  No CUDA call in this file: qsv_hostcheck.cu runs the same pass on a machine without a GPU.
#include <stdlib.h>

#include <algorithm>
#include <cmath>
#include <complex>
#include <map>

#include "qsv_sweep_ops.cuh"

// ================================================================================================
// host: fusion pass
// ================================================================================================

static SegMap make_segmap(const std::vector<int>& dst_of_src) {
    // dst_of_src[i] = destination bit of source bit i; merge consecutive runs
    SegMap m;
    memset(&m, 0, sizeof(m));
    int i = 0, n = (int)dst_of_src.size();
    while (i < n) {
        int j = i + 1;
        while (j < n && dst_of_src[j] == dst_of_src[j - 1] + 1) ++j;
        m.src[m.n] = (uint8_t)i;
        m.len[m.n] = (uint8_t)(j - i);
        m.dst[m.n] = (uint8_t)dst_of_src[i];
        ++m.n;
        i = j;
    }
    return m;
}

static SegMap64 make_segmap64(const std::vector<int>& dst_of_src) {
    SegMap64 m;
    memset(&m, 0, sizeof(m));
    int i = 0, n = (int)dst_of_src.size();
    while (i < n) {
        int j = i + 1;
        while (j < n && dst_of_src[j] == dst_of_src[j - 1] + 1) ++j;
        m.src[m.n] = (uint8_t)i;
        m.len[m.n] = (uint8_t)(j - i);
        m.dst[m.n] = (uint8_t)dst_of_src[i];
        ++m.n;
        i = j;
    }
    return m;
}

namespace {

struct Poly {  // phase polynomial over physical bits
    double t0 = 0;
    std::map<int, double> a;
    std::map<std::pair<int, int>, double> b;
    bool empty() const { return t0 == 0 && a.empty() && b.empty(); }
};

struct Builder {
    qsv_ctx* ctx;
    SweepPlan* plan;
    int n, k, c_low;
    // current sweep
    std::vector<size_t> cur;  // indices into pg
    uint64_t cur_targets = 0;
    // SWAP gates (CNOT triples) are pure relabels: pos[b] = physical position that currently holds logical bit b.
    // Gates are rewritten to physical positions (pg) when they are scheduled.
    std::vector<qsv_gate_t> pg;
    int pos[64];
    struct SwapEvent { size_t first; int a, b; };
    std::vector<SwapEvent> swaps;      // in program order
    std::vector<char> consumed;        // gates that belong to a recognised SWAP
    size_t next_gate = 0;              // index of the next unscheduled source gate (for look-ahead)
    uint64_t relabelled_gates = 0;
    // Register bits per stage of the launch being emitted: RB (4) for the cp.async images, RB8 (3) when flush() tries the
    // radix-8 TMA image first.  Arrays stay sized for RB; with rb = 3 register bit 3 simply never occurs.
    static constexpr int RETRY_RB4 = -1000;  // internal: the radix-8 attempt of flush() cannot hold this launch
    int n_targets_cur = 0;                   // butterfly targets of the sweep being flushed
    int rb = RB;
    bool tma_swizzle = false;          // tile slots follow the TMA engine's 128-byte swizzle instead of swz()
    bool allow8 = true;                // (sweep_plan_build_host: complex128, QSV_NO_FFT8 unset)
    // Layout the plan must END in: logical bit b at physical position final_pos[b] (identity = canonical).  A sharded run
    // asks for the layout its next exchange step needs, so that the correcting permutation rides on the last compute
    // sweep's store instead of costing a sweep of its own.  carry_exchange: the plan's last sweep must store out of place
    // (its store becomes the exchange: it writes into the peers' buffers) and must run in an image that can do that.
    int final_pos[64];
    bool carry_exchange = false;
    bool last_flush = false;
    // With a requested end layout the permutation is folded into the LAST sweep only: the exchange that follows wants it
    // there, and an early out-of-place sweep would make the segment ping-pong through the buffer the peers store into.
    bool defer_fold = false;
    uint32_t sw(uint32_t l) const { return tma_swizzle ? swz128(l) : swz(l); }

    Builder(qsv_ctx* c, SweepPlan* p, int n_, int k_) : ctx(c), plan(p), n(n_), k(k_) {
        c_low = (k < n) ? std::min(4, k - 1) : 0;
        for (int b = 0; b < 64; ++b) pos[b] = final_pos[b] = b;
        find_swaps();
    }

    static bool is_cnot(const qsv_gate_t& g, int* c, int* t) {
        if (g.kind != QSV_PERMX || g.n_targets != 1) return false;
        if (__builtin_popcountll(g.ctrl_mask) != 1 || g.ctrl_value != g.ctrl_mask) return false;
        *c = __builtin_ctzll(g.ctrl_mask);
        *t = g.target_bits[0];
        return true;
    }

    void find_swaps() {
        const auto& G = ctx->gates;
        consumed.assign(G.size(), 0);
        static const bool off = getenv("QSV_NO_RELABEL") != nullptr;  // debugging aid
        if (off) return;
        for (size_t i = 0; i + 2 < G.size(); ++i) {
            int c0, t0, c1, t1, c2, t2;
            if (is_cnot(G[i], &c0, &t0) && is_cnot(G[i + 1], &c1, &t1) && is_cnot(G[i + 2], &c2, &t2) && c1 == t0 && t1 == c0 &&
                c2 == c0 && t2 == t0) {
                swaps.push_back({i, c0, t0});
                consumed[i] = consumed[i + 1] = consumed[i + 2] = 1;
                i += 2;
            }
        }
    }

    // rewrite a source gate to the physical positions its bits occupy right now
    qsv_gate_t to_physical(const qsv_gate_t& g) const {
        qsv_gate_t r = g;
        r.ctrl_mask = r.ctrl_value = 0;
        for (int b = 0; b < n; ++b)
            if (g.ctrl_mask & (1ull << b)) {
                r.ctrl_mask |= 1ull << pos[b];
                if (g.ctrl_value & (1ull << b)) r.ctrl_value |= 1ull << pos[b];
            }
        for (uint32_t j = 0; j < g.n_targets; ++j) r.target_bits[j] = (uint8_t)pos[g.target_bits[j]];
        return r;
    }

    const double* mat(const qsv_gate_t& g) const { return ctx->pool.data() + 2 * g.matrix_offset; }
    static bool is_identity_diag(const double* m) { return m[0] == 1.0 && m[1] == 0.0 && m[6] == 1.0 && m[7] == 0.0; }

    uint64_t forced_mask() const { return (1ull << c_low) - 1ull; }

    int target_cap = 64;  // max butterfly targets per sweep (see sweep_plan_build_host: balanced sweeps)
    // Mixed caps: the first n_small sweeps take at most six targets (the radix-8 / TMA image), the others base_cap.
    int base_cap = 64, n_small = 0, flushes = 0;
    bool fits(uint64_t targets) const {
        return __builtin_popcountll(targets | forced_mask()) <= k && __builtin_popcountll(targets) <= target_cap;
    }

    int add_gate(size_t gi) {
        next_gate = gi;
        const qsv_gate_t& src = ctx->gates[gi];
        if (consumed[gi]) {
            // first gate of a SWAP triple performs the relabel; the other two are absorbed
            for (auto& ev : swaps)
                if (ev.first == gi) std::swap(pos[ev.a], pos[ev.b]);
            relabelled_gates++;
            return QSV_OK;
        }
        if (src.kind == QSV_ROWS2 || src.kind == QSV_DENSEK) {
            // not fusable: run as its own kernel on the canonical layout
            next_gate = gi;
            int rc = flush();
            if (rc) return rc;
            rc = restore_layout();
            if (rc) return rc;
            PlanStep s{};
            s.fused = false;
            s.gate_index = gi;
            s.params = nullptr;
            s.n_gates = 1;
            plan->steps.push_back(s);
            return QSV_OK;
        }
        if (src.kind == QSV_DIAG1) {
            if (is_identity_diag(mat(src))) {
                plan->dropped_identities++;
                return QSV_OK;
            }
            pg.push_back(to_physical(src));
            cur.push_back(pg.size() - 1);
            return QSV_OK;
        }
        uint64_t tb = 1ull << pos[src.target_bits[0]];
        if (!fits(cur_targets | tb)) {
            int rc = flush();   // may permute the layout
            if (rc) return rc;
            tb = 1ull << pos[src.target_bits[0]];
        }
        cur_targets |= tb;
        pg.push_back(to_physical(src));
        cur.push_back(pg.size() - 1);
        return QSV_OK;
    }

    // Data permutation (physical -> physical) that, applied now, makes the layout canonical once all
    // SWAP relabels from source gate `from` onwards have happened.  Returns false if none is needed.
    bool pending_permutation(size_t from, std::vector<int>& P) const {
        int tau[64];
        for (int b = 0; b < n; ++b) tau[b] = b;
        for (auto& ev : swaps)
            if (ev.first >= from) std::swap(tau[ev.a], tau[ev.b]);
        // at the end logical b would live at pos[tau[b]]; we want it at final_pos[b] (canonical: b)
        P.assign(n, 0);
        bool any = false;
        for (int b = 0; b < n; ++b) {
            P[pos[tau[b]]] = final_pos[b];
            any |= pos[tau[b]] != final_pos[b];
        }
        return any;
    }

    void apply_permutation(const std::vector<int>& P) {
        for (int b = 0; b < n; ++b) pos[b] = P[pos[b]];
    }

    // stand-alone permuting sweep (no gates) that restores the canonical layout
    int restore_layout(bool even_if_identity = false) {
        std::vector<int> P;
        if (!pending_permutation((size_t)-1, P) && !even_if_identity) return QSV_OK;   // only relabels already made count
        uint64_t need = forced_mask();
        for (int p = 0; p < n; ++p)
            if (P[p] < c_low) need |= 1ull << p;
        cur_targets = need;  // (2 * c_low <= k always holds: c_low <= 4, k >= 9 whenever c_low == 4)
        if (__builtin_popcountll(need) > k) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "unsupported: tile too small for the layout restore");
        choose_tile();
        Emit e;
        int rc = push_sweep(e, true, &P);
        apply_permutation(P);
        cur_targets = 0;
        return rc;
    }

    // ---- emit one sweep from `cur` ----------------------------------------------------------
    struct StageCtx {
        std::vector<int> rloc;     // local bit positions of register bits (ascending), size RB
        std::vector<int> tloc;     // thread local bits (ascending)
    };

    std::vector<int> tile_pos;     // physical position of each local bit
    std::vector<int> glob_pos;     // physical position of each tile-id bit
    int loc_of[64], glob_of[64];

    void choose_tile() {
        uint64_t m = cur_targets | forced_mask();
        // fill with the lowest free bits (longer contiguous runs -> better coalescing)
        for (int b = 0; b < n && __builtin_popcountll(m) < k; ++b) m |= 1ull << b;
        tile_pos.clear();
        glob_pos.clear();
        for (int b = 0; b < n; ++b) {
            loc_of[b] = glob_of[b] = -1;
            if (m & (1ull << b)) {
                loc_of[b] = (int)tile_pos.size();
                tile_pos.push_back(b);
            } else {
                glob_of[b] = (int)glob_pos.size();
                glob_pos.push_back(b);
            }
        }
    }

    // control decomposition of a (mask,value) over physical bits for the current stage
    void split_ctrl(uint64_t mask, uint64_t value, const StageCtx& sc, SwOp& op) const {
        op.rmask = op.rval = 0;
        op.lmask = op.lval = 0;
        op.gmask = op.gval = 0;
        for (int b = 0; b < n; ++b) {
            if (!(mask & (1ull << b))) continue;
            const bool one = value & (1ull << b);
            if (loc_of[b] >= 0) {
                const int l = loc_of[b];
                auto it = std::find(sc.rloc.begin(), sc.rloc.end(), l);
                if (it != sc.rloc.end()) {
                    const int ri = (int)(it - sc.rloc.begin());
                    op.rmask |= 1u << ri;
                    if (one) op.rval |= 1u << ri;
                } else {
                    op.lmask |= (uint16_t)(1u << l);
                    if (one) op.lval |= (uint16_t)(1u << l);
                }
            } else {
                op.gmask |= 1ull << glob_of[b];
                if (one) op.gval |= 1ull << glob_of[b];
            }
        }
    }

    struct Emit {
        std::vector<SwOp> ops;
        std::vector<std::vector<double>> gslots;  // per slot: [const, coef per global bit]
        std::vector<double> tab;                  // shared phase tables of this launch (re,im pairs)
        std::vector<int> first_rloc, last_rloc;   // register bits of the first / last stage
        uint64_t n_gates = 0;
    };

    int new_gslot(Emit& e, double cst, const std::vector<double>& coef) {
        std::vector<double> s(1 + glob_pos.size(), 0.0);
        s[0] = cst;
        for (size_t i = 0; i < coef.size(); ++i) s[1 + i] = coef[i];
        e.gslots.push_back(s);
        return (int)e.gslots.size() - 1;
    }

    void emit_stage_op(Emit& e, const StageCtx& sc) {
        SwOp op;
        memset(&op, 0, sizeof(op));
        op.type = OP_STAGE;
        for (int q = 0; q < rb; ++q) op.st.rpos[q] = (uint8_t)sc.rloc[q];
        for (int j = 0; j < (1 << rb); ++j) {
            uint32_t off = 0;
            for (int q = 0; q < rb; ++q)
                if ((j >> q) & 1) off |= 1u << sc.rloc[q];
            op.st.roffsw[j] = (uint16_t)sw(off);
        }
        if (e.first_rloc.empty()) e.first_rloc = sc.rloc;
        e.last_rloc = sc.rloc;
        e.ops.push_back(op);
    }

    // close a run of diagonal gates: register-register phases -> OP_W now, everything else -> deferred (Pending)
    int emit_diag_run(Emit& e, const StageCtx& sc, const std::vector<size_t>& run) {
        if (run.empty()) return QSV_OK;
        Poly P;
        std::vector<SwOp> fallbacks;
        auto in_reg = [&](int b) { return loc_of[b] >= 0 && std::find(sc.rloc.begin(), sc.rloc.end(), loc_of[b]) != sc.rloc.end(); };
        for (size_t gi : run) {
            const qsv_gate_t& g = pg[gi];
            const double* m = mat(g);
            const int t = g.target_bits[0];
            const double a0 = std::atan2(m[1], m[0]), a1 = std::atan2(m[7], m[6]);
            const int nc = __builtin_popcountll(g.ctrl_mask);
            bool poly_ok = nc <= 1;
            int cb = -1;
            if (nc == 1) {
                cb = __builtin_ctzll(g.ctrl_mask);
                if (!in_reg(t) && !in_reg(cb)) poly_ok = false;  // a pair term needs one register bit
            }
            if (poly_ok) {
                const double d = a1 - a0;
                if (nc == 0) {
                    P.t0 += a0;
                    if (d != 0) P.a[t] += d;
                } else if (g.ctrl_value & g.ctrl_mask) {  // control on 1
                    if (a0 != 0) P.a[cb] += a0;
                    if (d != 0) P.b[{std::min(cb, t), std::max(cb, t)}] += d;
                } else {  // control on 0: (1 - x_c)(a0 + d x_t)
                    P.t0 += a0;
                    if (d != 0) P.a[t] += d;
                    if (a0 != 0) P.a[cb] -= a0;
                    if (d != 0) P.b[{std::min(cb, t), std::max(cb, t)}] -= d;
                }
            } else {
                for (int side = 0; side < 2; ++side) {
                    const double re = m[side ? 6 : 0], im = m[side ? 7 : 1];
                    if (re == 1.0 && im == 0.0) continue;
                    SwOp op;
                    memset(&op, 0, sizeof(op));
                    op.type = OP_MPHASE;
                    const uint64_t tm = 1ull << t;
                    split_ctrl(g.ctrl_mask | tm, g.ctrl_value | (side ? tm : 0), sc, op);
                    op.m[0] = re;
                    op.m[1] = im;
                    fallbacks.push_back(op);
                }
            }
            e.n_gates++;
        }
        if (!P.empty()) {
            // classify the polynomial's terms for this stage
            const int n_thr = (int)sc.tloc.size();
            auto reg_index = [&](int b) -> int {
                if (loc_of[b] < 0) return -1;
                auto it = std::find(sc.rloc.begin(), sc.rloc.end(), loc_of[b]);
                return it == sc.rloc.end() ? -1 : (int)(it - sc.rloc.begin());
            };
            auto thr_index = [&](int b) -> int {
                if (loc_of[b] < 0) return -1;
                auto it = std::find(sc.tloc.begin(), sc.tloc.end(), loc_of[b]);
                return it == sc.tloc.end() ? -1 : (int)(it - sc.tloc.begin());
            };
            pend.ensure_size(n_thr, (int)glob_pos.size());
            pend.c0.cst += P.t0;
            pend.c0.any |= P.t0 != 0;
            double w[RB][RB] = {{0}};
            for (auto& kv : P.a) {
                const int b = kv.first;
                const int ri = reg_index(b);
                Pending::Term& tm = ri >= 0 ? pend.bit[ri] : pend.c0;
                if (ri >= 0) tm.cst += kv.second;
                else if (thr_index(b) >= 0) tm.thr[thr_index(b)] += kv.second;
                else tm.glb[glob_of[b]] += kv.second;
                tm.any = true;
            }
            bool has_w = false;
            for (auto& kv : P.b) {
                const int b1 = kv.first.first, b2 = kv.first.second;
                const int r1 = reg_index(b1), r2 = reg_index(b2);
                if (r1 >= 0 && r2 >= 0) {
                    w[std::min(r1, r2)][std::max(r1, r2)] += kv.second;
                    has_w = true;
                } else {
                    const int ri = r1 >= 0 ? r1 : r2;
                    const int ob = r1 >= 0 ? b2 : b1;
                    if (thr_index(ob) >= 0) pend.bit[ri].thr[thr_index(ob)] += kv.second;
                    else pend.bit[ri].glb[glob_of[ob]] += kv.second;
                    pend.bit[ri].any = true;
                }
            }
            if (has_w) {
                // register-register terms do not commute with later butterflies on those bits: apply them here
                int rc = ensure_budget(e, sc, 1, 0, NREG);
                if (rc) return rc;
                const size_t off = e.tab.size() / 2;
                e.tab.resize(e.tab.size() + 2 * NREG);
                uint16_t wm = 0;
                for (int j = 0; j < NREG; ++j) {
                    double ang = 0;
                    for (int p_ = 0; p_ < RB; ++p_)
                        for (int q = p_ + 1; q < RB; ++q)
                            if (((j >> p_) & 1) && ((j >> q) & 1)) ang += w[p_][q];
                    e.tab[2 * (off + j)] = std::cos(ang);
                    e.tab[2 * (off + j) + 1] = std::sin(ang);
                    if (ang != 0) wm |= (uint16_t)(1u << j);
                }
                if (!e.ops.empty() && e.ops.back().type == OP_HADD) {
                    // the butterfly just emitted is directly followed by this table: one op instead of two
                    e.ops.back().type = OP_HW;
                    e.ops.back().w.w_off = (uint16_t)off;
                    e.ops.back().w.wmask = wm;
                } else {
                    SwOp op;
                    memset(&op, 0, sizeof(op));
                    op.type = OP_W;
                    op.w.w_off = (uint16_t)off;
                    op.w.wmask = wm;
                    e.ops.push_back(op);
                }
            }
        }
        if (!fallbacks.empty()) {
            int rc = ensure_budget(e, sc, fallbacks.size(), 0, 0);
            if (rc) return rc;
            for (auto& f : fallbacks) e.ops.push_back(f);
        }
        return QSV_OK;
    }

    // Deferred phases of the open stage.  A factor exp(i x_a (c + sum thread/tile terms)) on register bit a
    // commutes with everything that follows in the stage except a butterfly/swap on bit a itself, so it is
    // accumulated here (angles add) and applied once, by OP_TWID, at the latest possible point.
    struct Pending {
        struct Term {
            double cst = 0;
            std::vector<double> thr, glb;
            bool any = false;
            void clear() {
                cst = 0;
                std::fill(thr.begin(), thr.end(), 0.0);
                std::fill(glb.begin(), glb.end(), 0.0);
                any = false;
            }
        };
        Term bit[RB], c0;
        void ensure_size(int n_thr, int n_glob) {
            for (int i = 0; i <= RB; ++i) {
                Term& t = i < RB ? bit[i] : c0;
                if ((int)t.thr.size() != n_thr) t.thr.assign(n_thr, 0.0);
                if ((int)t.glb.size() != n_glob) t.glb.assign(n_glob, 0.0);
            }
        }
    } pend;
    double sweep_scale = 1.0;  // product of the 1/sqrt2 factors owed by unnormalised butterflies

    static bool nz(const std::vector<double>& x) { return std::any_of(x.begin(), x.end(), [](double d) { return d != 0; }); }

    // make room in the current launch; if it is full, close it and reopen the same stage in a new one
    int ensure_budget(Emit& e, const StageCtx& sc, size_t ops_needed, size_t slots_needed, size_t tab_needed) {
        if (rb == RB8 && (e.ops.size() + ops_needed > MAX_OPS8 || e.gslots.size() + slots_needed > MAX_GSLOT8 ||
                          e.tab.size() / 2 + tab_needed > MAX_TAB))
            return RETRY_RB4;  // (the radix-8 image does not split a sweep: the four-bit attempt decides)
        if (e.ops.size() + ops_needed > MAX_OPS || e.gslots.size() + slots_needed > MAX_GSLOT || e.tab.size() / 2 + tab_needed > MAX_TAB) {
            int rc = push_sweep(e, false, nullptr);
            if (rc) return rc;
            emit_stage_op(e, sc);
        }
        return QSV_OK;
    }

    uint16_t make_thread_table(Emit& e, const std::vector<double>& coef) {
        // two 16-entry tables over the low / high 4-bit chunk of the thread id
        const int n_thr = (int)coef.size();
        const size_t off = e.tab.size() / 2;
        e.tab.resize(e.tab.size() + 2 * TSLOT_LEN);
        for (int chunk = 0; chunk < 2; ++chunk)
            for (int vv = 0; vv < 16; ++vv) {
                double ang = 0;
                for (int q = 0; q < 4; ++q) {
                    const int tb_i = chunk * 4 + q;
                    if (tb_i < n_thr && ((vv >> q) & 1)) ang += coef[tb_i];
                }
                e.tab[2 * (off + chunk * 16 + vv)] = std::cos(ang);
                e.tab[2 * (off + chunk * 16 + vv) + 1] = std::sin(ang);
            }
        return (uint16_t)off;
    }

    // emit OP_TWID for the register bits in `mask` (and, if with_c0, the thread-wide factor and the owed scale)
    int flush_pending(Emit& e, const StageCtx& sc, unsigned mask, bool with_c0) {
        size_t slots = 0, tabs = 0;
        int na = 0;
        for (int ri = 0; ri < RB; ++ri)
            if (((mask >> ri) & 1) && pend.bit[ri].any) {
                ++na;
                if (nz(pend.bit[ri].thr)) tabs += TSLOT_LEN;
                if (pend.bit[ri].cst != 0 || nz(pend.bit[ri].glb)) ++slots;
            }
        const bool c0_on = with_c0 && pend.c0.any;
        const bool scale_on = with_c0 && sweep_scale != 1.0;
        if (c0_on) {
            if (nz(pend.c0.thr)) tabs += TSLOT_LEN;
            if (pend.c0.cst != 0 || nz(pend.c0.glb)) ++slots;
        }
        if (na == 0 && !c0_on && !scale_on) return QSV_OK;
        int rc = ensure_budget(e, sc, 1, slots, tabs);
        if (rc) return rc;
        SwOp op;
        memset(&op, 0, sizeof(op));
        op.type = OP_TWID;
        op.tw.scale = scale_on ? sweep_scale : 1.0;
        if (scale_on) sweep_scale = 1.0;
        for (int s_ = 0; s_ <= RB; ++s_) op.tw.tslot[s_] = op.tw.gslot[s_] = NONE16;
        int k_ = 0;
        for (int ri = 0; ri < RB; ++ri) {
            op.tw.abit[ri] = 0xFF;
            if (((mask >> ri) & 1) && pend.bit[ri].any) {
                Pending::Term& t = pend.bit[ri];
                op.tw.abit[ri] = (uint8_t)ri;
                if (nz(t.thr)) op.tw.tslot[ri] = make_thread_table(e, t.thr);
                if (t.cst != 0 || nz(t.glb)) op.tw.gslot[ri] = (uint16_t)new_gslot(e, t.cst, t.glb);
                t.clear();
                ++k_;
            }
        }
        op.tw.n_active = (uint8_t)k_;
        if (c0_on) {
            Pending::Term& t = pend.c0;
            if (nz(t.thr)) op.tw.tslot[RB] = make_thread_table(e, t.thr);
            if (t.cst != 0 || nz(t.glb)) op.tw.gslot[RB] = (uint16_t)new_gslot(e, t.cst, t.glb);
            t.clear();
        }
        e.ops.push_back(op);
        return QSV_OK;
    }

    void finish_stage_regs(StageCtx& sc, const std::vector<int>& want) {
        // register bits: the wanted local bits, padded with the highest free local bits
        std::vector<int> r = want;
        for (int l = k - 1; l >= 0 && (int)r.size() < rb; --l)
            if (std::find(r.begin(), r.end(), l) == r.end()) r.push_back(l);
        std::sort(r.begin(), r.end());
        sc.rloc = r;
        sc.tloc.clear();
        for (int l = 0; l < k; ++l)
            if (std::find(r.begin(), r.end(), l) == r.end()) sc.tloc.push_back(l);
    }

    // Does the 16-entry table at `off` have the shape OP_FSTAGE applies after the butterfly on register bit A: 1 for
    // registers without bit A, and for the others a value that depends only on the register bits below A?
    static bool fft_table_shape(const std::vector<double>& tab, size_t off, int A) {
        for (int j = 0; j < NREG; ++j) {
            const double re = tab[2 * (off + j)], im = tab[2 * (off + j) + 1];
            if (!((j >> A) & 1)) {
                if (re != 1.0 || im != 0.0) return false;
                continue;
            }
            const int m = j & ((1 << A) - 1);
            if (m == 0) {
                if (re != 1.0 || im != 0.0) return false;
                continue;
            }
            const size_t r = off + (size_t)((1 << A) | m);
            if (re != tab[2 * r] || im != tab[2 * r + 1]) return false;
        }
        return true;
    }

    // If every stage of the launch is "FFT shaped" -- butterflies on distinct register bits, each optionally followed by a
    // phase table that couples it only to bits whose butterfly comes later (or never), and at most one deferred-phase op at
    // the end -- rewrite each stage (OP_STAGE + its ops) as ONE OP_FSTAGE.  Register bits are relabelled so that the
    // butterflies run on bits 3, 2, 1, 0 in that order (tables and per-bit slots are permuted to match).  All-or-nothing per
    // launch: the IMG_FFT kernel image runs only such ops and carries nothing in registers between them.
    bool to_fstages(Emit& e) {
        static const bool off = getenv("QSV_NO_FSTAGE") != nullptr;  // debugging / A-B aid
        static const bool dbg = getenv("QSV_FSTAGE_DEBUG") != nullptr;
#define FS_NO(why)                                                                  \
    do {                                                                            \
        if (dbg) fprintf(stderr, "[qsv] launch stays generic: %s\n", why);          \
        return false;                                                               \
    } while (0)
        if (off || e.ops.empty()) return false;
        std::vector<double> tab = e.tab;  // tables are permuted on a copy: a failed conversion must leave the launch intact
        std::vector<std::vector<double>> gslots = e.gslots;
        int one_g = -1, one_t = -1, one_w = -1;
        auto ones_gslot = [&]() {
            if (one_g < 0) {
                gslots.push_back(std::vector<double>(1 + glob_pos.size(), 0.0));
                one_g = (int)gslots.size() - 1;
            }
            return (uint16_t)one_g;
        };
        auto ones_table = [&](int& slot, size_t len) {
            if (slot < 0) {
                slot = (int)(tab.size() / 2);
                for (size_t i = 0; i < len; ++i) {
                    tab.push_back(1.0);
                    tab.push_back(0.0);
                }
            }
            return (uint16_t)slot;
        };
        std::vector<SwOp> out;
        std::vector<int> last_order;
        size_t i = 0;
        while (i < e.ops.size()) {
            if (e.ops[i].type != OP_STAGE) FS_NO("does not start with a stage");
            const SwOp st = e.ops[i++];
            // ---- pass 1: the stage's ops and the butterfly order
            std::vector<const SwOp*> body;
            while (i < e.ops.size() && e.ops[i].type != OP_STAGE) body.push_back(&e.ops[i++]);
            std::vector<int> order;  // old register bits in butterfly order
            const SwOp* tw = nullptr;
            for (const SwOp* o : body) {
                if (tw) FS_NO("an op follows the deferred phases");
                if (o->type == OP_HADD || o->type == OP_HW) {
                    if (std::find(order.begin(), order.end(), (int)o->tgt) != order.end()) FS_NO("two butterflies on one register bit");
                    order.push_back(o->tgt);
                } else if (o->type == OP_TWID) {
                    tw = o;
                } else {
                    if (dbg) fprintf(stderr, "[qsv] op kind %d\n", (int)o->type);
                    FS_NO("an op kind outside the FFT shape");
                }
            }
            const int nh = (int)order.size();
            for (int q = rb - 1; q >= 0; --q)  // the other register bits keep their relative order below the butterfly bits
                if (std::find(order.begin(), order.end(), q) == order.end()) order.push_back(q);
            int newbit[RB];  // old register bit -> new register bit
            for (int r = 0; r < RB; ++r) newbit[r] = r;   // (rb = 3: bit 3 does not exist and maps to itself)
            for (int r = 0; r < rb; ++r) newbit[order[r]] = rb - 1 - r;
            auto map_j = [&](int j) {
                int r = 0;
                for (int q = 0; q < RB; ++q)
                    if ((j >> q) & 1) r |= 1 << newbit[q];
                return r;
            };
            SwOp f;
            memset(&f, 0, sizeof(f));
            f.type = OP_FSTAGE;
            f.fs.scale = 1.0;
            f.fs.hmask = (uint8_t)nh;
            for (int s_ = 0; s_ <= RB; ++s_) f.fs.tslot[s_] = f.fs.gslot[s_] = NONE16;
            for (int q = 0; q < RB; ++q) f.fs.abit[q] = 0xFF;
            for (int q = 0; q < rb; ++q) {
                f.fs.rpos[q] = st.st.rpos[q];
                f.fs.boff[newbit[q]] = (uint16_t)sw(1u << st.st.rpos[q]);
            }
            for (int q = 0; q < 3; ++q) f.fs.w_off[q] = NONE16;
            // ---- pass 2: tables
            bool any_w = false;
            for (const SwOp* o : body) {
                if (o->type != OP_HW) continue;
                const int A = newbit[o->tgt];
                double perm[2 * NREG];
                for (int j = 0; j < NREG; ++j) {
                    perm[2 * map_j(j)] = tab[2 * (o->w.w_off + j)];
                    perm[2 * map_j(j) + 1] = tab[2 * (o->w.w_off + j) + 1];
                }
                memcpy(&tab[2 * o->w.w_off], perm, sizeof(perm));
                if (!fft_table_shape(tab, o->w.w_off, A)) FS_NO("a phase table couples bits outside the FFT shape");
                if (A > 0 && o->w.wmask) {
                    f.fs.w_off[rb - 1 - A] = o->w.w_off;
                    any_w = true;
                }
            }
            if (any_w)
                for (int q = 0; q < rb - 1; ++q)
                    if (f.fs.w_off[q] == NONE16) f.fs.w_off[q] = ones_table(one_w, NREG);
            // ---- deferred phases
            if (tw) {
                f.fs.scale = tw->tw.scale;
                bool any_bit = false;
                for (int q = 0; q < rb; ++q)
                    if (tw->tw.abit[q] != 0xFF) {
                        const int nq = newbit[q];
                        f.fs.abit[nq] = (uint8_t)nq;
                        f.fs.tslot[nq] = tw->tw.tslot[q];
                        f.fs.gslot[nq] = tw->tw.gslot[q];
                        any_bit = true;
                    }
                f.fs.n_active = tw->tw.n_active;
                f.fs.tslot[RB] = tw->tw.tslot[RB];
                f.fs.gslot[RB] = tw->tw.gslot[RB];
                const bool c0 = f.fs.tslot[RB] != NONE16 || f.fs.gslot[RB] != NONE16;
                if (any_bit) f.fs.flags |= 1;
                if (c0 || f.fs.scale != 1.0) f.fs.flags |= 2;
                if (c0) f.fs.flags |= 4;
                // every factor the kernel reads unconditionally must point somewhere: absent ones at all-ones slots
                for (int nq = 0; nq <= RB; ++nq) {
                    if (nq >= rb && nq < RB) continue;
                    const bool used = nq == RB ? c0 : (any_bit && (f.fs.abit[nq] != 0xFF || nq >= rb - nh));
                    if (!used) continue;
                    if (f.fs.tslot[nq] == NONE16) f.fs.tslot[nq] = ones_table(one_t, TSLOT_LEN);
                    if (f.fs.gslot[nq] == NONE16) f.fs.gslot[nq] = ones_gslot();
                }
                for (int nq = rb - nh; nq < rb; ++nq)
                    if (any_bit) f.fs.abit[nq] = (uint8_t)nq;
            }
            f.tgt = (uint8_t)(2 * nh + (any_w ? 1 : 0));  // shape code (added to DC_FSTAGE at conversion)
            out.push_back(f);
            last_order = order;
        }
        if (gslots.size() > MAX_GSLOT || tab.size() / 2 > MAX_TAB) FS_NO("no room for the all-ones slots");
#undef FS_NO
        e.ops.swap(out);
        e.tab.swap(tab);
        e.gslots.swap(gslots);
        // the direct store enumerates the last stage's register bits by (new) register index
        if (!e.last_rloc.empty()) {
            std::vector<int> r(rb);
            for (int rnk = 0; rnk < rb; ++rnk) r[rb - 1 - rnk] = e.last_rloc[last_order[rnk]];
            e.last_rloc = r;
        }
        return true;
    }

    // How the TMA engine sees a tile whose bits sit at the index positions `bits_at` (ascending): false when the bit runs
    // cannot be described (needs index bits 0..2 in the tile and at most TMA_MAX_INST instructions of five dimensions).
    bool build_tma(const std::vector<int>& bits_at, TmaTile& T) const {
        memset(&T, 0, sizeof(T));
        if (k < 4 || k > 11 || (int)bits_at.size() != k || bits_at[0] != 0 || bits_at[1] != 1 || bits_at[2] != 2) return false;
        std::vector<std::pair<int, int>> runs;  // (first index bit, length), ascending, at most 8 bits each (box extent <= 256)
        for (int l = 3; l < k; ++l) {
            const int b = bits_at[l];
            if (!runs.empty() && runs.back().first + runs.back().second == b && runs.back().second < 8) runs.back().second++;
            else runs.push_back({b, 1});
        }
        std::vector<int> inst_bits;  // index bits walked by separate instructions (the top tile bits)
        while (runs.size() > 4) {
            auto& r = runs.back();
            inst_bits.insert(inst_bits.begin(), r.first + r.second - 1);
            if (--r.second == 0) runs.pop_back();
        }
        if (runs.empty() || runs[0].first != 3 || (1 << inst_bits.size()) > TMA_MAX_INST) return false;
        T.rank = 1 + (int)runs.size();
        T.box[0] = 16;
        T.size[0] = 16;
        T.shift[0] = 0;
        T.bits[0] = 3;
        int in_dims = 3;
        for (size_t d = 0; d < runs.size(); ++d) {
            const int s0 = runs[d].first, s1 = d + 1 < runs.size() ? runs[d + 1].first : n;
            if (s1 - s0 > 31) return false;  // (coordinates are signed 32-bit)
            T.shift[d + 1] = (uint8_t)s0;
            T.bits[d + 1] = (uint8_t)(s1 - s0);
            T.box[d + 1] = 1u << runs[d].second;
            T.size[d + 1] = 1ull << (s1 - s0);
            T.stride_bytes[d + 1] = 16ull << s0;
            in_dims += runs[d].second;
        }
        T.n_inst = 1 << inst_bits.size();
        T.inst_amps = 1u << in_dims;
        for (int e_ = 0; e_ < T.n_inst; ++e_)
            for (size_t i = 0; i < inst_bits.size(); ++i)
                if ((e_ >> i) & 1) T.inst_goff[e_] |= 1ull << inst_bits[i];
        return true;
    }

    // Everything the radix-8 image needs for the launch in `e` (after to_fstages): load / store views of the tile, and for a
    // permuting store the re-ordering of the tile by the last stage.
    bool build_sweep8(const Emit& e, const std::vector<int>* out_pos_of_phys, Sweep8Extra& X) const {
        memset(&X, 0, sizeof(X));
        if (!build_tma(tile_pos, X.in)) return false;
        X.out = X.in;
        for (int l = 0; l < k && l < KMAX; ++l) X.out_pos[l] = (uint8_t)tile_pos[l];
        if (!out_pos_of_phys) return true;
        std::vector<int> out_at(k), order(k), rank_of(k);
        for (int l = 0; l < k; ++l) {
            out_at[l] = (*out_pos_of_phys)[tile_pos[l]];
            order[l] = l;
        }
        std::sort(order.begin(), order.end(), [&](int x, int y) { return out_at[x] < out_at[y]; });
        std::vector<int> sorted_out(k);
        bool identity = true;
        for (int u = 0; u < k; ++u) {
            sorted_out[u] = out_at[order[u]];
            rank_of[order[u]] = u;
            identity &= order[u] == u;
        }
        if (!build_tma(sorted_out, X.out)) return false;
        for (int u = 0; u < k && u < KMAX; ++u) X.out_pos[u] = (uint8_t)sorted_out[u];
        if (identity) return true;  // the tile keeps its order in shared memory; only the addresses differ
        X.permuting = 1;
        X.loc_perm = make_segmap(rank_of);
        // the last stage's register bit q sits at tile bit p_q (recovered from its slot offset: the swizzle only adds lower bits)
        const SwOp& last = e.ops.back();
        int p_last[RB8];
        for (int q = 0; q < RB8; ++q) {
            p_last[q] = 31 - __builtin_clz((unsigned)last.fs.boff[q]);
            X.boff_out[q] = (uint16_t)swz128(1u << rank_of[p_last[q]]);
        }
        static const bool no_mid = getenv("QSV_NO_MIDLAYOUT") != nullptr;  // A/B aid
        if (e.ops.size() == 2 && !no_mid) {
            // intermediate layout: see Sweep8Extra
            X.mid_layout = 1;
            const SwOp& first = e.ops[0];
            for (int q = 0; q < RB8; ++q) {
                const int p_first = 31 - __builtin_clz((unsigned)first.fs.boff[q]);
                X.boff_mid_w[q] = (uint16_t)swz(1u << p_first);
                X.boff_last_r[q] = (uint16_t)swz(1u << p_last[q]);
            }
            std::vector<int> rest;  // the last stage's thread bits: its non-register tile bits, lowest OUTPUT position first
            for (int l = 0; l < k; ++l)
                if (l != p_last[0] && l != p_last[1] && l != p_last[2]) rest.push_back(l);
            std::sort(rest.begin(), rest.end(), [&](int x, int y) { return rank_of[x] < rank_of[y]; });
            X.tmap_last = make_segmap(rest);
        }
        return true;
    }

    int push_sweep(Emit& e, bool out_of_place, const std::vector<int>* out_pos_of_phys) {
        const bool has_work = std::any_of(e.ops.begin(), e.ops.end(), [](const SwOp& o) { return o.type != OP_STAGE; });
        if (!has_work && !out_of_place) {
            const uint64_t carried = e.n_gates;
            e = Emit();
            e.n_gates = carried;   // (only identities could get here)
            return QSV_OK;
        }
        if (e.ops.size() > MAX_OPS || e.gslots.size() > MAX_GSLOT || e.tab.size() / 2 > MAX_TAB)
            return qsv_fail(ctx, QSV_E_TOOBIG, "internal: sweep budget exceeded (%zu ops, %zu slots, %zu table entries)", e.ops.size(),
                            e.gslots.size(), e.tab.size() / 2);
        const bool all_fft = has_work && to_fstages(e);
        Sweep8Extra extra8;
        if (rb == RB8) {
            // the radix-8 image runs FFT-shaped launches only, a handful of stages and scalars, on tiles the TMA engine can
            // move; nothing has been committed to the plan yet, so flush() can start over with four register bits
            // (launches that are not FFT-shaped run its generic stage loop, in place only)
            static const bool no_gen8 = getenv("QSV_NO_GEN8") != nullptr;  // A/B aid
            if (!has_work || (!all_fft && (out_of_place || no_gen8)) || (all_fft && n_targets_cur > 6) || e.ops.size() > MAX_OPS8 ||
                e.gslots.size() > MAX_GSLOT8 ||
                e.tab.size() / 2 > MAX_TAB || !build_sweep8(e, all_fft ? out_pos_of_phys : nullptr, extra8))
                return RETRY_RB4;
        }
        SweepParams* sp = new SweepParams();
        memset(sp, 0, sizeof(*sp));
        SweepDesc& D = sp->d;
        D.n_bits = n;
        D.k = k;
        D.n_ops = (int)e.ops.size();
        D.n_gslots = (int)e.gslots.size();
        D.n_glob = (int)glob_pos.size();
        // per-tile scalars exp(i (c + sum_b coef_b t_b)) as a product of 16-entry tables over 4-bit chunks of the tile id
        D.g_chunks = std::max(1, (D.n_glob + 3) / 4);
        D.gtab_off = (uint32_t)(plan->aux.size() / 2);
        for (auto& gsl : e.gslots)
            for (int c = 0; c < D.g_chunks; ++c)
                for (int vv = 0; vv < 16; ++vv) {
                    double ang = c == 0 ? gsl[0] : 0.0;
                    for (int q = 0; q < 4; ++q) {
                        const int b = 4 * c + q;
                        if (b < D.n_glob && ((vv >> q) & 1)) ang += gsl[1 + b];
                    }
                    plan->aux.push_back(std::cos(ang));
                    plan->aux.push_back(std::sin(ang));
                }
        D.tab_off = (uint32_t)(plan->aux.size() / 2);
        D.tab_len = (uint32_t)(e.tab.size() / 2);
        plan->aux.insert(plan->aux.end(), e.tab.begin(), e.tab.end());
        std::vector<int> in_dst(k), out_dst(k), enum_dst(k);
        for (int l = 0; l < k; ++l) {
            in_dst[l] = tile_pos[l];
            out_dst[l] = out_pos_of_phys ? (*out_pos_of_phys)[tile_pos[l]] : tile_pos[l];
        }
        // store enumeration: counter bits walk the local bits in ascending OUTPUT position
        std::vector<int> order(k);
        for (int l = 0; l < k; ++l) order[l] = l;
        std::sort(order.begin(), order.end(), [&](int x, int y) { return out_dst[x] < out_dst[y]; });
        for (int u = 0; u < k; ++u) enum_dst[u] = order[u];
        D.in_loc = make_segmap(in_dst);
        D.out_loc = make_segmap(out_dst);
        D.st_enum = make_segmap(enum_dst);
        D.tb = std::max(5, k - rb);
        // direct global <-> register paths: need full tiles (k >= 9) and a stage whose register bits
        // leave the lowest local bits to the lanes, so that the accesses stay coalesced
        auto high_only = [](const std::vector<int>& r) { return !r.empty() && *std::min_element(r.begin(), r.end()) >= 4; };
        static const bool no_direct = getenv("QSV_NO_DIRECT") != nullptr;  // debugging aid
        D.direct_load = 0;  // (loads always go through the cp.async double buffer)
        D.direct_store = !no_direct && has_work && k >= 9 && !out_of_place && high_only(e.last_rloc);
        {
            std::vector<int> lr = e.last_rloc;  // (register order of an OP_FSTAGE stage is not position order)
            std::sort(lr.begin(), lr.end());
            for (int q = 0; q < RB; ++q) {
                D.first_rpos[q] = (uint8_t)(e.first_rloc.empty() || q >= rb ? q : e.first_rloc[q]);
                D.last_rpos[q] = (uint8_t)(lr.empty() || q >= rb ? q : lr[q]);
            }
        }
        auto dep = [&](uint32_t l, const std::vector<int>& dstpos) {
            uint64_t g = 0;
            for (int b = 0; b < k; ++b)
                if ((l >> b) & 1) g |= 1ull << dstpos[b];
            return g;
        };
        for (int i = 0; i < NREG; ++i) {
            // load side
            uint32_t l_ld;
            if (D.direct_load) {
                l_ld = 0;
                for (int q = 0; q < rb; ++q)
                    if ((i >> q) & 1) l_ld |= 1u << e.first_rloc[q];
            } else {
                l_ld = (uint32_t)i << D.tb;
            }
            if (l_ld < (1u << k)) {
                D.ld_goff[i] = dep(l_ld, in_dst);
                D.ld_soff[i] = (uint16_t)sw(l_ld);
            }
            // store side
            uint32_t l_st;
            if (D.direct_store) {
                if (i >= (1 << rb)) continue;
                l_st = 0;
                for (int q = 0; q < rb; ++q)
                    if ((i >> q) & 1) l_st |= 1u << e.last_rloc[q];
            } else {
                const uint32_t u = (uint32_t)i << D.tb;
                if (u >= (1u << k)) continue;
                l_st = 0;
                for (int b = 0; b < k; ++b)
                    if ((u >> b) & 1) l_st |= 1u << enum_dst[b];
            }
            D.st_goff[i] = dep(l_st, out_dst);
            D.st_soff[i] = (uint16_t)sw(l_st);
        }
        {
            std::vector<int> gin(glob_pos.size()), gout(glob_pos.size());
            for (size_t b = 0; b < glob_pos.size(); ++b) {
                gin[b] = glob_pos[b];
                gout[b] = out_pos_of_phys ? (*out_pos_of_phys)[glob_pos[b]] : glob_pos[b];
            }
            D.in_glob = make_segmap64(gin);
            D.out_glob = make_segmap64(gout);
        }
        bool all_lean = true;
        for (size_t i = 0; i < e.ops.size(); ++i) {
            SwOp o = e.ops[i];
            const bool ctrl = o.rmask || o.lmask || o.gmask;
            switch (o.type) {  // builder-level kind -> device dispatch code
                case OP_STAGE: o.type = DC_STAGE; break;
                case OP_DENSE: {
                    const bool real = o.m[1] == 0.0 && o.m[3] == 0.0 && o.m[5] == 0.0 && o.m[7] == 0.0;
                    o.type = (uint8_t)((ctrl ? DC_DENSE_CPLX_C : (real ? DC_DENSE_REAL_NC : DC_DENSE_CPLX_NC)) + o.tgt);
                    break;
                }
                case OP_PERMX: o.type = (uint8_t)(DC_PERMX + o.tgt); break;
                case OP_MPHASE: o.type = DC_MPHASE; break;
                case OP_HADD: o.type = (uint8_t)(DC_HADD + o.tgt); break;
                case OP_HW: o.type = (uint8_t)(DC_HW + o.tgt); break;
                case OP_W: o.type = DC_W; break;
                case OP_FSTAGE: o.type = (uint8_t)(DC_FSTAGE + o.tgt); break;
                default: o.type = DC_TWID; break;
            }
            sp->ops[i] = o;
            const uint8_t c_ = o.type;
            if (c_ >= DC_FIRST_FULL) all_lean = false;
        }
        PlanStep s{};
        s.fused = true;
        s.params = sp;
        s.out_of_place = out_of_place;
        s.img = rb == RB8 ? IMG_FFT8 : all_fft ? IMG_FFT : all_lean ? IMG_LEAN : IMG_FULL;
        s.n_gates = e.n_gates;
        if (s.img == IMG_FFT8) s.tma = new Sweep8Extra(extra8);
        plan->steps.push_back(s);
        e = Emit();
        return QSV_OK;
    }

    int flush() {
        const bool had_gates = !cur.empty();
        const int rc = flush_sweep();
        if (!rc && had_gates) {
            ++flushes;
            target_cap = flushes < n_small ? std::min(6, base_cap) : base_cap;
        }
        return rc;
    }

    int flush_sweep() {
        if (cur.empty()) {
            cur_targets = 0;
            return QSV_OK;
        }
        // Radix-8 TMA image first when the sweep's butterfly targets fill one or two three-bit stages as well as they
        // would fill four-bit ones (1-3 or 5-6 targets); anything that does not fit it starts over with four register bits.
        int n_t = 0;
        {
            uint64_t seen = 0;
            for (size_t gi : cur)
                if (pg[gi].kind != QSV_DIAG1) seen |= 1ull << pg[gi].target_bits[0];
            n_t = __builtin_popcountll(seen);
        }
        // (a sweep whose store carries an exchange fans out over several peers.  The cp.async images do that by default; the
        // radix-8 image can too -- its threads then store the output-ordered tile themselves instead of the engine -- but
        // such a sweep is bound by the wire, not by its stages: measured on 8 B200 the radix-8 version is no faster on QFT-30
        // (2.93 against 2.90 ms) and 4 % slower on QFT-34 (46.7 against 44.8 ms), profiles/r2_ab_carry8.txt.  QSV_CARRY8=1
        // selects it.)
        // (FFT-shaped launches: up to six targets = two stages; generic ones may take a third stage, see push_sweep)
        // A sweep of swaps and phases only (multi-controlled X streams) is cheap per stage: measured faster in three radix-8
        // stages than in two or three radix-16 ones (cyclic shift, 2^30: 25.2 against 28.7 ms), so it may carry up to 9 targets.
        bool dense = false;
        for (size_t gi : cur) dense |= pg[gi].kind == QSV_DENSE1;
        static const int gen8_env = getenv("QSV_GEN8_MAX_TARGETS") ? atoi(getenv("QSV_GEN8_MAX_TARGETS")) : 0;
        const int gen8_max = gen8_env ? gen8_env : (dense ? 6 : 9);
        n_targets_cur = n_t;
        static const bool no_carry8 = getenv("QSV_CARRY8") == nullptr;
        const bool try8 = allow8 && !(last_flush && carry_exchange && no_carry8) && k < n && k >= 9 && k <= 11 && n_t >= 1 && n_t <= std::max(6, gen8_max) &&
                          n_t != 4;
        if (try8) {
            const size_t steps0 = plan->steps.size(), aux0 = plan->aux.size();
            const uint64_t targets0 = cur_targets;
            const double scale0 = sweep_scale;
            rb = RB8;
            tma_swizzle = true;
            const int rc = flush_emit();
            rb = RB;
            tma_swizzle = false;
            if (rc != RETRY_RB4) return rc;
            while (plan->steps.size() > steps0) {
                delete plan->steps.back().params;
                delete plan->steps.back().tma;
                plan->steps.pop_back();
            }
            plan->aux.resize(aux0);
            cur_targets = targets0;
            sweep_scale = scale0;
            pend = Pending();
        }
        return flush_emit();
    }

    int flush_emit() {
        // Can the permutation that the remaining SWAP relabels call for ride on this sweep's store?  It can
        // when the tile also holds the bits that become the low (coalescing) bits of the output.
        std::vector<int> P;
        bool fold = false;
        if ((!defer_fold || last_flush) && (pending_permutation(next_gate, P) || (last_flush && carry_exchange))) {
            uint64_t need = cur_targets | forced_mask();
            for (int p = 0; p < n; ++p)
                if (P[p] < c_low) need |= 1ull << p;
            if (__builtin_popcountll(need) <= k) {
                fold = true;
                cur_targets = need;
            }
            static const bool dbg = getenv("QSV_PLAN_DEBUG") != nullptr;
            if (dbg)
                fprintf(stderr, "[qsv] fold? targets %#llx need %#llx (%d bits, tile %d) -> %s\n", (unsigned long long)cur_targets,
                        (unsigned long long)need, __builtin_popcountll(need), k, fold ? "yes" : "no");
        }
        choose_tile();
        Emit e;
        StageCtx sc;
        std::vector<int> want;            // local bits the open stage needs in registers
        // Two-phase per stage: first collect the gates of a stage (greedy), then emit with the final register set.
        size_t i = 0;
        while (i < cur.size()) {
            // ---- collect one stage ----
            want.clear();
            size_t j = i;
            for (; j < cur.size(); ++j) {
                const qsv_gate_t& g = pg[cur[j]];
                if (g.kind == QSV_DIAG1) continue;
                const int l = loc_of[g.target_bits[0]];
                if (std::find(want.begin(), want.end(), l) != want.end()) continue;
                if ((int)want.size() == rb) break;
                want.push_back(l);
            }
            // diagonal gates prefer a register bit of their own: if the stage has room, give pair gates one
            if ((int)want.size() < rb) {
                for (size_t q = i; q < j && (int)want.size() < rb; ++q) {
                    const qsv_gate_t& g = pg[cur[q]];
                    if (g.kind != QSV_DIAG1 || __builtin_popcountll(g.ctrl_mask) != 1) continue;
                    const int t = g.target_bits[0], cb = __builtin_ctzll(g.ctrl_mask);
                    auto has = [&](int b) { return loc_of[b] >= 0 && std::find(want.begin(), want.end(), loc_of[b]) != want.end(); };
                    if (has(t) || has(cb)) continue;
                    if (loc_of[t] >= 0) want.push_back(loc_of[t]);
                    else if (loc_of[cb] >= 0) want.push_back(loc_of[cb]);
                }
            }
            finish_stage_regs(sc, want);
            // budgets: a launch carries at most MAX_OPS ops, MAX_GSLOT per-tile scalars and MAX_TAB table entries;
            // ensure_budget() closes the launch and reopens the same stage in the next one when any would overflow
            if (rb == RB8 && e.ops.size() + 16 > MAX_OPS8) return RETRY_RB4;
            if (e.ops.size() + 16 > MAX_OPS) {
                int rc = push_sweep(e, false, nullptr);
                if (rc) return rc;
            }
            emit_stage_op(e, sc);
            // unnormalised butterflies pay off when the stage has at least two plain Hadamards (the owed
            // 1/sqrt2 factors are settled by ONE scaling per sweep)
            auto is_plain_h = [&](const qsv_gate_t& g) {
                if (g.kind != QSV_DENSE1 || g.ctrl_mask) return false;
                const double* m = mat(g);
                const double h = m[0];
                return std::fabs(h - 0.70710678118654752) < 1e-15 && m[1] == 0 && m[2] == h && m[3] == 0 && m[4] == h && m[5] == 0 &&
                       m[6] == -h && m[7] == 0;
            };
            int n_plain_h = 0;
            for (size_t q = i; q < j; ++q) n_plain_h += is_plain_h(pg[cur[q]]);
            const bool hadd_mode = n_plain_h >= 1;  // (a single one costs the same either way, and keeps the stage FFT-shaped)
            std::vector<size_t> run;
            const size_t RUN_CHUNK = 64;
            for (size_t q = i; q < j; ++q) {
                const qsv_gate_t& g = pg[cur[q]];
                if (g.kind == QSV_DIAG1) {
                    run.push_back(cur[q]);
                    if (run.size() >= RUN_CHUNK) {
                        int rc = emit_diag_run(e, sc, run);
                        if (rc) return rc;
                        run.clear();
                    }
                    continue;
                }
                int rc = emit_diag_run(e, sc, run);
                if (rc) return rc;
                run.clear();
                const int l = loc_of[g.target_bits[0]];
                const uint8_t tg = (uint8_t)(std::find(sc.rloc.begin(), sc.rloc.end(), l) - sc.rloc.begin());
                // a butterfly / swap on register bit tg does not commute with phases deferred on that bit
                if (pend.bit[tg].any) {
                    rc = flush_pending(e, sc, 1u << tg, false);
                    if (rc) return rc;
                }
                rc = ensure_budget(e, sc, 1, 0, 0);
                if (rc) return rc;
                SwOp op;
                memset(&op, 0, sizeof(op));
                if (hadd_mode && is_plain_h(g)) {
                    op.type = OP_HADD;
                    op.tgt = tg;
                    sweep_scale *= mat(g)[0];
                } else {
                    op.type = g.kind == QSV_PERMX ? OP_PERMX : OP_DENSE;
                    split_ctrl(g.ctrl_mask, g.ctrl_value, sc, op);
                    op.tgt = tg;
                    if (g.kind != QSV_PERMX) memcpy(op.m, mat(g), 8 * sizeof(double));
                }
                e.ops.push_back(op);
                e.n_gates++;
            }
            {
                int rc = emit_diag_run(e, sc, run);
                if (rc) return rc;
                run.clear();
                // stage end: everything deferred must land before the register bits change; the owed scale
                // rides along if this is the sweep's last stage
                const bool last_stage = j >= cur.size();
                const double keep = sweep_scale;
                if (!last_stage) sweep_scale = 1.0;  // (the thread-wide factor is applied per stage, the scale once per sweep)
                rc = flush_pending(e, sc, (1u << RB) - 1u, true);
                if (!last_stage) sweep_scale = keep;
                if (rc) return rc;
            }
            i = j;
        }
        int rc = push_sweep(e, fold, fold ? &P : nullptr);
        if (rc) return rc;   // (RETRY_RB4 included: nothing below may happen before the launch is really in the plan)
        if (fold) apply_permutation(P);
        cur.clear();
        cur_targets = 0;
        return rc;
    }
};

}  // namespace

// Commuting-gate scheduler.  The builder below packs gates into a sweep in program order and closes the sweep at the first
// gate whose target does not fit the tile any more -- although later gates that act on untouched qubits could still ride
// along.  Gates with disjoint supports (targets and controls) commute, and diagonal gates fit any sweep, so the stream is
// re-ordered first: per sweep, walk the remaining gates in order; take a gate if it shares no qubit with any gate deferred
// so far and its target (at the physical position the SWAP relabels give it) still fits; else defer it and block its qubits.
// The relative order of gates that share a qubit never changes, so the operator is the same (to rounding; exactly for
// permutation streams).  Measured on random controlled-gate streams: 11 -> 9, 42 -> 36, 207 -> 167 sweeps.
static void schedule_commuting_gates(qsv_ctx* ctx, int k) {
    auto& G = ctx->gates;
    const int n = ctx->prog_bits;
    const size_t m = G.size();
    if (m < 3 || k >= n) return;
    for (const qsv_gate_t& g : G)
        if (g.kind == QSV_ROWS2 || g.kind == QSV_DENSEK) return;  // (streams with unfusable gates keep their order)
    const int c_low = std::min(4, k - 1);
    const uint64_t forced = (1ull << c_low) - 1ull, all = n >= 64 ? ~0ull : (1ull << n) - 1ull;
    // units: single gates, or a SWAP written as three CNOTs (kept together: the builder turns it into a relabel)
    struct Unit { size_t first; int len; uint64_t support; int target; bool free; int sa, sb; };
    std::vector<Unit> units;
    auto cnot = [&](const qsv_gate_t& g, int* c, int* t) {
        if (g.kind != QSV_PERMX || g.n_targets != 1 || __builtin_popcountll(g.ctrl_mask) != 1 || g.ctrl_value != g.ctrl_mask) return false;
        *c = __builtin_ctzll(g.ctrl_mask);
        *t = g.target_bits[0];
        return true;
    };
    static const bool no_relabel = getenv("QSV_NO_RELABEL") != nullptr;
    for (size_t i = 0; i < m;) {
        int c0, t0, c1, t1, c2, t2;
        if (!no_relabel && i + 2 < m && cnot(G[i], &c0, &t0) && cnot(G[i + 1], &c1, &t1) && cnot(G[i + 2], &c2, &t2) && c1 == t0 && t1 == c0 &&
            c2 == c0 && t2 == t0) {
            units.push_back({i, 3, (1ull << c0) | (1ull << t0), -1, true, c0, t0});
            i += 3;
        } else {
            const qsv_gate_t& g = G[i];
            const int t = g.target_bits[0];
            units.push_back({i, 1, g.ctrl_mask | (1ull << t), t, g.kind == QSV_DIAG1, -1, -1});
            ++i;
        }
    }
    std::vector<char> done(units.size(), 0);
    std::vector<size_t> order;
    order.reserve(units.size());
    int pos[64];
    for (int b = 0; b < 64; ++b) pos[b] = b;
    size_t start = 0;
    while (start < units.size()) {
        uint64_t targets = 0, blocked = 0;
        for (size_t u = start; u < units.size() && blocked != all; ++u) {
            if (done[u]) continue;
            const Unit& un = units[u];
            bool take = !(un.support & blocked);
            if (take && !un.free) {
                const uint64_t nt = targets | (1ull << pos[un.target]);
                if (__builtin_popcountll(nt | forced) <= k) targets = nt;
                else take = false;
            }
            if (take) {
                if (un.sa >= 0) std::swap(pos[un.sa], pos[un.sb]);
                done[u] = 1;
                order.push_back(u);
            } else {
                blocked |= un.support;
            }
        }
        while (start < units.size() && done[start]) ++start;
    }
    std::vector<qsv_gate_t> out;
    out.reserve(m);
    for (size_t u : order)
        for (int j = 0; j < units[u].len; ++j) out.push_back(G[units[u].first + j]);
    G.swap(out);
}

int sweep_plan_build_host(qsv_ctx* ctx, int max_tile_bits, SweepPlan** out_plan, const int* final_pos, bool carry_exchange) {
    const int n = ctx->prog_bits;
    if (n < RB) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "unsupported: fused sweeps need at least %d index bits (have %d)", RB, n);
    // default tile: the whole state when it fits, else 2^11 amplitudes (32 KB): measured best on B200 -- three
    // double-buffered CTAs per SM hide more latency than one 64 KB-tile CTA
    int k = max_tile_bits > 0 ? max_tile_bits : (n <= KMAX ? KMAX : 11);
    if (max_tile_bits <= 0 && getenv("QSV_TILE_BITS")) k = atoi(getenv("QSV_TILE_BITS"));  // tuning / debugging aid
    k = std::min(std::min(k, KMAX), n);
    if (k < RB) return qsv_fail(ctx, QSV_E_INVALID, "max_tile_bits must be >= %d", RB);
    if (ctx->plan) {
        sweep_plan_free(ctx->plan);
        ctx->plan = nullptr;
    }
    static const bool no_sched = getenv("QSV_NO_SCHEDULE") != nullptr;  // A/B aid: keep the program order
    if (!no_sched) schedule_commuting_gates(ctx, k);
    auto build = [&](int cap, SweepPlan** out, int n_small = 0) -> int {
        SweepPlan* plan = new SweepPlan();
        plan->k = k;
        plan->aux.resize(2, 0.0);  // offset 0 reserved
        Builder b(ctx, plan, n, k);
        b.base_cap = cap;
        b.n_small = n_small;
        b.target_cap = n_small > 0 ? std::min(6, cap) : cap;
        static const bool no8 = getenv("QSV_NO_FFT8") != nullptr;  // A/B aid: keep every launch in the cp.async images
        b.allow8 = ctx->dtype == QSV_C128 && !no8;
        if (final_pos)
            for (int q = 0; q < n; ++q) b.final_pos[q] = final_pos[q];
        b.carry_exchange = carry_exchange;
        b.defer_fold = final_pos != nullptr;
        int rc = QSV_OK;
        for (size_t i = 0; i < ctx->gates.size() && !rc; ++i) rc = b.add_gate(i);
        if (!rc) {
            b.next_gate = ctx->gates.size();
            b.last_flush = true;
            rc = b.flush();
        }
        if (!rc) {
            // the exchange rides on an out-of-place store: if the last sweep is not one, a gate-free permuting sweep is added
            const bool last_ok = !plan->steps.empty() && plan->steps.back().fused && plan->steps.back().out_of_place;
            rc = b.restore_layout(carry_exchange && !last_ok);
        }
        if (rc) {
            delete plan;
            return rc;
        }
        plan->dropped_identities += b.relabelled_gates;  // counted as applied source gates
        *out = plan;
        return QSV_OK;
    };
    SweepPlan* best = nullptr;
    int best_cap = 64;
    {
        int rc = build(64, &best);
        if (rc) return rc;
    }
    // Balanced sweeps: a sweep's cost grows with the butterflies it carries, so among the plans with the FEWEST launches
    // take the one with the smallest per-sweep target count (QFT-30 at k = 11: 6+6+6+6+6 instead of 7+7+7+7+2; measured
    // 33.6 vs 34.1 ms).  Planning is host work of O(gates); the search is skipped for very long programs.
    static const char* cap_env = getenv("QSV_SWEEP_TARGETS");  // tuning aid: fixed cap, no search
    if (cap_env) {
        SweepPlan* p = nullptr;
        const char* colon = strchr(cap_env, ':');  // "cap" or "cap:m" (first m sweeps held to six targets)
        if (build(atoi(cap_env), &p, colon ? atoi(colon + 1) : 0) == QSV_OK) {
            delete best;
            best = p;
        }
    } else if (ctx->gates.size() <= 50000 && k < n) {
        for (int cap = k - 2; cap >= 3; --cap) {
            SweepPlan* p = nullptr;
            if (build(cap, &p) != QSV_OK) break;
            if (p->steps.size() > best->steps.size()) {
                delete p;
                break;
            }
            delete best;
            best = p;
            best_cap = cap;
        }
        // Mixed caps.  When the balanced plan needs seven or more targets per sweep, none of its sweeps can take the radix-8 /
        // TMA image; but often the FIRST m sweeps can be held to six targets without adding a launch (31 targets in 5 sweeps:
        // 6+6+6+6+7 instead of 7+7+7+7+3).  A six-target radix-8 sweep is faster than a seven-target radix-16 one (11.9
        // against 13.5 ms over 2^31 amplitudes), and the radix-16 image pays ~25 % extra on a plan's first sweep (tile rows
        // 128 MB apart), which the TMA image does not.  Largest m that keeps the launch count wins.  Measured: QFT-31 on one
        // B200 59.6 against 65.7 ms per step (profiles/r2_ab_mixed_caps.txt).
        static const bool no_mixed = getenv("QSV_NO_MIXED_CAPS") != nullptr;  // A/B aid
        auto n_fft8 = [](const SweepPlan* p) {
            int c = 0;
            for (auto& st : p->steps) c += st.fused && st.img == IMG_FFT8;
            return c;
        };
        if (!no_mixed && best_cap >= 7 && best_cap < 64 && ctx->dtype == QSV_C128 && k >= 9 && k <= 11) {
            for (int m = (int)best->steps.size() - 1; m >= 1; --m) {
                SweepPlan* p = nullptr;
                if (build(best_cap, &p, m) != QSV_OK) break;
                if (getenv("QSV_PLAN_DEBUG"))
                    fprintf(stderr, "[qsv] mixed caps: first %d sweeps <= 6 targets, others <= %d: %zu launches (%d radix-8) against %zu (%d)\n", m,
                            best_cap, p->steps.size(), n_fft8(p), best->steps.size(), n_fft8(best));
                if (p->steps.size() <= best->steps.size() && n_fft8(p) > n_fft8(best)) {
                    delete best;
                    best = p;
                    break;
                }
                delete p;
            }
        }
        // Tried and measured worse (profiles/r2_ab_generic8.txt), kept behind QSV_COST_MODEL=1: trading launches for cheaper
        // ones.  A sweep of at most six targets is eligible for the radix-8 / TMA image; for 96 random controlled dense gates
        // the six-target plan has 14 sweeps of 5.8-7.8 ms instead of 11 of 7.0-9.1 ms -- 95.4 against 86.2 ms per step.
        auto cost = [](const SweepPlan* p) {
            double c = 0;
            for (auto& st : p->steps) c += !st.fused ? 1.0 : st.img == IMG_FFT8 ? 1.0 : st.img == IMG_FFT ? 1.1 : 1.4;
            return c;
        };
        static const bool use_cost = getenv("QSV_COST_MODEL") != nullptr;  // A/B aid
        if (use_cost && ctx->dtype == QSV_C128 && k >= 9 && k <= 11) {
            SweepPlan* p = nullptr;
            if (build(6, &p) == QSV_OK) {
                if (cost(p) < cost(best)) {
                    delete best;
                    best = p;
                } else {
                    delete p;
                }
            }
        }
    }
    if (getenv("QSV_PLAN_DEBUG")) {
        static const char* names[] = {"full", "lean", "fft16", "fft8-tma"};
        for (auto& st : best->steps)
            if (st.fused)
                fprintf(stderr, "[qsv] sweep: image %s%s, %d ops, %d scalars, %llu gates%s%s\n", names[st.img],
                        st.img == IMG_FFT8 && st.params->d.n_ops && st.params->ops[0].type == DC_STAGE ? " (generic stages)" : "", st.params->d.n_ops,
                        st.params->d.n_gslots,
                        (unsigned long long)st.n_gates, st.out_of_place ? ", permuting store" : "",
                        st.img == IMG_FFT8 ? (st.tma->permuting ? ", tile re-ordered for the TMA store" : ", TMA store")
                                           : (st.params->d.direct_store ? ", direct store" : ""));
    }
    *out_plan = best;
    return QSV_OK;
}

// One gate-free sweep whose out-of-place store applies the bit permutation `perm` (bit perm[i] of the output index =
// bit i of the input index); QSV_E_UNSUPPORTED when the tile cannot hold the bits the store needs.
int sweep_permute_plan_build(qsv_ctx* ctx, int n_bits, const int* perm, int k, SweepPlan* tmp) {
    tmp->k = k;
    tmp->aux.resize(2, 0.0);
    const int saved_bits = ctx->prog_bits;
    std::vector<qsv_gate_t> saved_gates;
    saved_gates.swap(ctx->gates);   // the builder scans ctx->gates for SWAPs; give it an empty stream
    ctx->prog_bits = n_bits;
    int rc;
    {
        Builder b(ctx, tmp, n_bits, k);
        std::vector<int> P(perm, perm + n_bits);
        uint64_t need = b.forced_mask();
        for (int p = 0; p < n_bits; ++p)
            if (P[p] < b.c_low) need |= 1ull << p;
        if (__builtin_popcountll(need) > k) {
            rc = QSV_E_UNSUPPORTED;
        } else {
            b.cur_targets = need;
            b.choose_tile();
            Builder::Emit e;
            rc = b.push_sweep(e, true, &P);
        }
    }
    ctx->gates.swap(saved_gates);
    ctx->prog_bits = saved_bits;
    return rc;
}
