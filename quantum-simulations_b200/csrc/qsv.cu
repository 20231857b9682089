// libqsv: C ABI + one-kernel-per-gate statevector kernels for sm_100a.
//
// Kernels in this file (names follow SURVEY.md section 2.1):
//   K1 k_gate1<DENSE>   controlled 2x2 butterfly, control mask folded into the index math
//   K2 k_gate1<SWAP>    controlled X as a pure amplitude swap (bit-exact, no arithmetic)
//   K3 k_gate1<DIAG> / k_phase1   controlled diagonal; touches only the selected amplitudes
//   K4 k_densek<T>      controlled dense 2^T x 2^T, T <= 5
//   K5 k_rows           m x m unitary on m rows of a [dim][batch] array (two-level / Givens)
//   K6 k_permute_bits   bit permutation of the index (qubit remap, local part)
//   K7 k_project        gather rows with ancilla bits = 0
//   K8 k_compare(+final) fused max|a-b| and allclose violation count
//   k_basis             identity columns with ancillas in |0>
// All are HBM-bound streaming kernels: one 128-bit load + one 128-bit store per touched
// complex128 amplitude, 4 independent pairs per thread in flight.
#include <stdarg.h>
#include <stdlib.h>

#include <algorithm>
#include <cmath>

#include "qsv_common.cuh"

static char g_create_err[512] = {0};

int qsv_fail(qsv_ctx* ctx, int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(ctx ? ctx->err : g_create_err, 512, fmt, ap);
    va_end(ap);
    return code;
}

// ================================================================================================
// K1/K2/K3: one target bit
// ================================================================================================
enum { MODE_DENSE = 0, MODE_SWAP = 1, MODE_DIAG = 2 };

template <typename R, int MODE, int U>
__global__ void __launch_bounds__(256) k_gate1(typename cplx_of<R>::type* __restrict__ st, uint64_t npairs,
                                               const __grid_constant__ BitInsert ins, uint64_t tmask, Mat2 mat) {
    typedef typename cplx_of<R>::type C;
    const uint64_t base = (uint64_t)blockIdx.x * (256 * U) + threadIdx.x;
    uint64_t i0[U];
    C a0[U], a1[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const uint64_t p = base + (uint64_t)u * 256;
        i0[u] = (p < npairs) ? bit_expand(p, ins) : ~0ull;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        if (i0[u] != ~0ull) {
            a0[u] = st[i0[u]];
            a1[u] = st[i0[u] | tmask];
        }
    }
    C m00, m01, m10, m11;
    if (MODE != MODE_SWAP) {
        m00 = cmake<C, R>((R)mat.m[0], (R)mat.m[1]);
        m01 = cmake<C, R>((R)mat.m[2], (R)mat.m[3]);
        m10 = cmake<C, R>((R)mat.m[4], (R)mat.m[5]);
        m11 = cmake<C, R>((R)mat.m[6], (R)mat.m[7]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        if (i0[u] != ~0ull) {
            C b0, b1;
            if (MODE == MODE_SWAP) {
                b0 = a1[u];
                b1 = a0[u];
            } else if (MODE == MODE_DIAG) {
                b0 = cmul(m00, a0[u]);
                b1 = cmul(m11, a1[u]);
            } else {
                b0 = cfma(m01, a1[u], cmul(m00, a0[u]));
                b1 = cfma(m11, a1[u], cmul(m10, a0[u]));
            }
            st[i0[u]] = b0;
            st[i0[u] | tmask] = b1;
        }
    }
}

// K1 for a LOW target bit (below the warp width), uncontrolled: every lane loads ONE amplitude (a warp reads 512 contiguous
// bytes), gets its partner from lane ^ 2^t with warp shuffles, and writes its own output -- the "warp-shuffle exchange for
// low-index target qubits".  k_gate1 handles the same gates by giving each thread both amplitudes of a pair (32 contiguous
// bytes per thread for t = 0); measured 1-3 % slower (profiles/r2_ab_shuffle.txt), so this kernel is the default for these
// gates and QSV_NO_SHFL=1 selects the pair kernel.
template <typename R, int U>
__global__ void __launch_bounds__(256) k_gate1_shfl(typename cplx_of<R>::type* __restrict__ st, uint64_t namps, int t, Mat2 mat) {
    typedef typename cplx_of<R>::type C;
    const uint64_t base = (uint64_t)blockIdx.x * (256 * U) + threadIdx.x;
    const bool hi = (threadIdx.x >> t) & 1;   // (the index bits below 8 are the thread's own: stride 256 between its amplitudes)
    // row of the matrix this lane computes: out = m_self * own + m_other * partner
    const C m_self = hi ? cmake<C, R>((R)mat.m[6], (R)mat.m[7]) : cmake<C, R>((R)mat.m[0], (R)mat.m[1]);
    const C m_other = hi ? cmake<C, R>((R)mat.m[4], (R)mat.m[5]) : cmake<C, R>((R)mat.m[2], (R)mat.m[3]);
    C a[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const uint64_t p = base + (uint64_t)u * 256;
        a[u] = p < namps ? st[p] : cmake<C, R>((R)0, (R)0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        C b;
        b.x = __shfl_xor_sync(0xffffffffu, a[u].x, 1 << t);
        b.y = __shfl_xor_sync(0xffffffffu, a[u].y, 1 << t);
        const uint64_t p = base + (uint64_t)u * 256;
        if (p < namps) st[p] = cfma(m_other, b, cmul(m_self, a[u]));
    }
}

// K3 special case diag(1, d): only amplitudes with the target bit set are touched
template <typename R, int U>
__global__ void __launch_bounds__(256) k_phase1(typename cplx_of<R>::type* __restrict__ st, uint64_t namps,
                                                const __grid_constant__ BitInsert ins, double dre, double dim) {
    typedef typename cplx_of<R>::type C;
    const uint64_t base = (uint64_t)blockIdx.x * (256 * U) + threadIdx.x;
    const C d = cmake<C, R>((R)dre, (R)dim);
    uint64_t ix[U];
    C a[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const uint64_t p = base + (uint64_t)u * 256;
        ix[u] = (p < namps) ? bit_expand(p, ins) : ~0ull;
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
        if (ix[u] != ~0ull) a[u] = st[ix[u]];
#pragma unroll
    for (int u = 0; u < U; ++u)
        if (ix[u] != ~0ull) st[ix[u]] = cmul(d, a[u]);
}

// ================================================================================================
// K4: dense 2^T x 2^T on T target bits (matrix in global memory, staged in shared)
// ================================================================================================
struct TargetSet { int t; uint64_t mask[QSV_MAX_TARGETS]; };  // mask[j] = bit of matrix-index bit (t-1-j)

template <typename R, int T>
__global__ void __launch_bounds__(128) k_densek(typename cplx_of<R>::type* __restrict__ st, uint64_t ngroups,
                                                const __grid_constant__ BitInsert ins, TargetSet ts, const double* __restrict__ mat) {
    typedef typename cplx_of<R>::type C;
    constexpr int D = 1 << T;
    __shared__ double2 sm[D * D];
    for (int i = threadIdx.x; i < D * D; i += blockDim.x) sm[i] = make_double2(mat[2 * i], mat[2 * i + 1]);
    __syncthreads();
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ngroups) return;
    const uint64_t base = bit_expand(g, ins);
    uint64_t off[D];
#pragma unroll
    for (int v = 0; v < D; ++v) {
        uint64_t o = 0;
#pragma unroll
        for (int j = 0; j < T; ++j)
            if ((v >> (T - 1 - j)) & 1) o |= ts.mask[j];
        off[v] = o;
    }
    C in[D];
#pragma unroll
    for (int v = 0; v < D; ++v) in[v] = st[base | off[v]];
#pragma unroll 1
    for (int r = 0; r < D; ++r) {
        C acc = cmake<C, R>((R)0, (R)0);
#pragma unroll
        for (int c = 0; c < D; ++c) {
            const double2 m = sm[r * D + c];
            acc = cfma(cmake<C, R>((R)m.x, (R)m.y), in[c], acc);
        }
        st[base | off[r]] = acc;
    }
}

// ================================================================================================
// K5: m x m unitary on rows of a [dim][batch] array
// ================================================================================================
struct RowSet { int m; uint64_t row[16]; };

template <typename R>
__global__ void __launch_bounds__(256) k_rows(typename cplx_of<R>::type* __restrict__ st, uint64_t batch,
                                              RowSet rs, const double* __restrict__ mat) {
    typedef typename cplx_of<R>::type C;
    __shared__ double2 sm[256];
    for (int i = threadIdx.x; i < rs.m * rs.m; i += blockDim.x) sm[i] = make_double2(mat[2 * i], mat[2 * i + 1]);
    __syncthreads();
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= batch) return;
    C in[16];
#pragma unroll
    for (int c = 0; c < 16; ++c)
        if (c < rs.m) in[c] = st[rs.row[c] * batch + j];
    for (int r = 0; r < rs.m; ++r) {
        C acc = cmake<C, R>((R)0, (R)0);
#pragma unroll
        for (int c = 0; c < 16; ++c)
            if (c < rs.m) {
                const double2 m = sm[r * rs.m + c];
                acc = cfma(cmake<C, R>((R)m.x, (R)m.y), in[c], acc);
            }
        st[rs.row[r] * batch + j] = acc;
    }
}

// two-level fast path: 2x2 by value, swap variant kept arithmetic-free
template <typename R, int MODE>
__global__ void __launch_bounds__(256) k_rows2(typename cplx_of<R>::type* __restrict__ st, uint64_t batch,
                                               uint64_t r0, uint64_t r1, Mat2 mat) {
    typedef typename cplx_of<R>::type C;
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= batch) return;
    const C a0 = st[r0 * batch + j], a1 = st[r1 * batch + j];
    C b0, b1;
    if (MODE == MODE_SWAP) {
        b0 = a1;
        b1 = a0;
    } else {
        const C m00 = cmake<C, R>((R)mat.m[0], (R)mat.m[1]), m01 = cmake<C, R>((R)mat.m[2], (R)mat.m[3]);
        const C m10 = cmake<C, R>((R)mat.m[4], (R)mat.m[5]), m11 = cmake<C, R>((R)mat.m[6], (R)mat.m[7]);
        b0 = cfma(m01, a1, cmul(m00, a0));
        b1 = cfma(m11, a1, cmul(m10, a0));
    }
    st[r0 * batch + j] = b0;
    st[r1 * batch + j] = b1;
}

// ================================================================================================
// state preparation / read-out
// ================================================================================================
template <typename R>
__global__ void k_basis(typename cplx_of<R>::type* __restrict__ st, uint64_t ncols, int batch_bits, uint64_t col0,
                        const __grid_constant__ BitInsert anc) {
    typedef typename cplx_of<R>::type C;
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ncols) return;
    const uint64_t row = bit_expand(col0 + j, anc);
    st[(row << batch_bits) | j] = cmake<C, R>((R)1, (R)0);
}

template <typename R>
__global__ void __launch_bounds__(256) k_project(const typename cplx_of<R>::type* __restrict__ in,
                                                 typename cplx_of<R>::type* __restrict__ out, uint64_t nout,
                                                 const __grid_constant__ BitInsert anc) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nout; i += stride)
        out[i] = in[bit_expand(i, anc)];
}

// Ancillas in a pure state psi instead of |0>: thread (j, v) writes column j's amplitude for ancilla assignment v
struct Psi { double a_re, a_im, b_re, b_im; };

template <typename R>
__global__ void k_basis_psi(typename cplx_of<R>::type* __restrict__ st, uint64_t ncols, int batch_bits, uint64_t col0,
                            const __grid_constant__ BitInsert anc, Psi psi) {
    typedef typename cplx_of<R>::type C;
    const uint64_t id = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t j = id % ncols, v = id / ncols;
    if (v >> anc.n) return;
    uint64_t row = bit_expand(col0 + j, anc);
    double2 w = make_double2(1.0, 0.0);
    for (int i = 0; i < anc.n; ++i) {
        const bool one = (v >> i) & 1ull;
        if (one) row |= 1ull << anc.pos[i];
        w = cmul(w, one ? make_double2(psi.b_re, psi.b_im) : make_double2(psi.a_re, psi.a_im));
    }
    st[(row << batch_bits) | j] = cmake<C, R>((R)w.x, (R)w.y);
}

// out[i] = sum_v conj(prod_i psi_{v_i}) in[expand(i) | ancillas = v]   (<psi| on every ancilla)
template <typename R>
__global__ void __launch_bounds__(256) k_project_psi(const typename cplx_of<R>::type* __restrict__ in,
                                                     typename cplx_of<R>::type* __restrict__ out, uint64_t nout,
                                                     const __grid_constant__ BitInsert anc, Psi psi) {
    typedef typename cplx_of<R>::type C;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nout; i += stride) {
        const uint64_t base = bit_expand(i, anc);
        double2 acc = make_double2(0.0, 0.0);
        for (uint64_t v = 0; !(v >> anc.n); ++v) {
            uint64_t x = base;
            double2 w = make_double2(1.0, 0.0);
            for (int q = 0; q < anc.n; ++q) {
                const bool one = (v >> q) & 1ull;
                if (one) x |= 1ull << anc.pos[q];
                w = cmul(w, one ? make_double2(psi.b_re, -psi.b_im) : make_double2(psi.a_re, -psi.a_im));
            }
            const C a = in[x];
            acc = cfma(w, make_double2((double)a.x, (double)a.y), acc);
        }
        out[i] = cmake<C, R>((R)acc.x, (R)acc.y);
    }
}

struct BitPerm { int n; uint8_t to[64]; };

template <typename R>
__global__ void __launch_bounds__(256) k_permute_bits(const typename cplx_of<R>::type* __restrict__ in,
                                                      typename cplx_of<R>::type* __restrict__ out, uint64_t n,
                                                      const __grid_constant__ BitPerm bp, int fixed_low) {
    // threads walk the OUTPUT index so that stores coalesce; bits below fixed_low map to themselves
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t y = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; y < n; y += stride) {
        uint64_t x = y & ((1ull << fixed_low) - 1ull);
        for (int i = fixed_low; i < bp.n; ++i) x |= ((y >> bp.to[i]) & 1ull) << i;
        out[y] = in[x];
    }
}

// ================================================================================================
// K8: fused compare
// ================================================================================================
#define CMP_BLOCKS 1184  // 148 SMs x 8
template <typename R>
__global__ void __launch_bounds__(256) k_compare(const typename cplx_of<R>::type* __restrict__ a,
                                                 const typename cplx_of<R>::type* __restrict__ b, uint64_t n,
                                                 double rtol, double atol, double* __restrict__ partial) {
    typedef typename cplx_of<R>::type C;
    double mx = 0.0;
    double bad = 0.0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const C x = a[i], y = b[i];
        const double dx = (double)x.x - (double)y.x, dy = (double)x.y - (double)y.y;
        const double d = hypot(dx, dy);
        const double ref = hypot((double)y.x, (double)y.y);
        // NaN anywhere counts as a violation, like np.allclose without equal_nan
        if (!(d <= atol + rtol * ref)) bad += 1.0;
        if (!(d <= mx)) mx = d;  // propagates NaN into the max
    }
    __shared__ double smx[256], sbad[256];
    smx[threadIdx.x] = mx;
    sbad[threadIdx.x] = bad;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            const double o = smx[threadIdx.x + s];
            if (!(o <= smx[threadIdx.x])) smx[threadIdx.x] = o;
            sbad[threadIdx.x] += sbad[threadIdx.x + s];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        partial[2 * blockIdx.x] = smx[0];
        partial[2 * blockIdx.x + 1] = sbad[0];
    }
}

__global__ void __launch_bounds__(256) k_compare_final(const double* __restrict__ partial, int nblocks,
                                                       double* __restrict__ result) {
    __shared__ double smx[256], sbad[256];
    double mx = 0.0, bad = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 256) {
        const double o = partial[2 * i];
        if (!(o <= mx)) mx = o;
        bad += partial[2 * i + 1];
    }
    smx[threadIdx.x] = mx;
    sbad[threadIdx.x] = bad;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            const double o = smx[threadIdx.x + s];
            if (!(o <= smx[threadIdx.x])) smx[threadIdx.x] = o;
            sbad[threadIdx.x] += sbad[threadIdx.x + s];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        result[0] = smx[0];
        result[1] = sbad[0];
    }
}

// first index whose entry is not ~0 (|b| > atol), and compare against a scaled second operand: the two pieces of
// the reference's allprop (quompiler/utils/mfun.py:34-52)
template <typename R>
__global__ void __launch_bounds__(256) k_first_nonzero(const typename cplx_of<R>::type* __restrict__ b, uint64_t n, double atol,
                                                       unsigned long long* __restrict__ out) {
    typedef typename cplx_of<R>::type C;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    unsigned long long best = ~0ull;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const C y = b[i];
        if (!(hypot((double)y.x, (double)y.y) <= atol)) {
            best = i;
            break;  // indexes only grow along this thread's stride
        }
    }
    if (best != ~0ull) atomicMin(out, best);
}

template <typename R>
__global__ void __launch_bounds__(256) k_compare_scaled(const typename cplx_of<R>::type* __restrict__ a,
                                                        const typename cplx_of<R>::type* __restrict__ b, uint64_t n, double lre,
                                                        double lim, double rtol, double atol, double* __restrict__ partial) {
    typedef typename cplx_of<R>::type C;
    double mx = 0.0, bad = 0.0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const C x = a[i], y = b[i];
        const double sx = lre * (double)y.x - lim * (double)y.y, sy = lre * (double)y.y + lim * (double)y.x;
        const double d = hypot((double)x.x - sx, (double)x.y - sy);
        if (!(d <= atol + rtol * hypot(sx, sy))) bad += 1.0;
        if (!(d <= mx)) mx = d;
    }
    __shared__ double smx[256], sbad[256];
    smx[threadIdx.x] = mx;
    sbad[threadIdx.x] = bad;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            const double o = smx[threadIdx.x + s];
            if (!(o <= smx[threadIdx.x])) smx[threadIdx.x] = o;
            sbad[threadIdx.x] += sbad[threadIdx.x + s];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        partial[2 * blockIdx.x] = smx[0];
        partial[2 * blockIdx.x + 1] = sbad[0];
    }
}

// ================================================================================================
// host side
// ================================================================================================
static int build_insert(qsv_ctx* ctx, const qsv_gate_t* g, int n_bits, BitInsert* ins, bool include_targets) {
    uint64_t m = g->ctrl_mask;
    if (include_targets)
        for (uint32_t j = 0; j < g->n_targets; ++j) {
            if (g->target_bits[j] >= n_bits) return qsv_fail(ctx, QSV_E_INVALID, "target bit %d >= n_bits %d", g->target_bits[j], n_bits);
            if (m & (1ull << g->target_bits[j])) return qsv_fail(ctx, QSV_E_INVALID, "target bit %d is also a control or listed twice", g->target_bits[j]);
            m |= 1ull << g->target_bits[j];
        }
    if (n_bits < 64 && (m >> n_bits)) return qsv_fail(ctx, QSV_E_INVALID, "control mask exceeds n_bits");
    if (g->ctrl_value & ~g->ctrl_mask) return qsv_fail(ctx, QSV_E_INVALID, "ctrl_value has bits outside ctrl_mask");
    ins->n = 0;
    for (int b = 0; b < 64; ++b)
        if (m & (1ull << b)) {
            if (ins->n >= QSV_MAX_INSERT) return qsv_fail(ctx, QSV_E_TOOBIG, "more than %d control+target bits", QSV_MAX_INSERT);
            ins->pos[ins->n++] = (uint8_t)b;
        }
    ins->value = g->ctrl_value;
    return QSV_OK;
}

static inline bool is_one(const double* z) { return z[0] == 1.0 && z[1] == 0.0; }
static inline bool is_zero(const double* z) { return z[0] == 0.0 && z[1] == 0.0; }

template <typename R>
static int launch_gate1(qsv_ctx* ctx, void* state, int n_bits, const qsv_gate_t* g, const double* mat, cudaStream_t s) {
    typedef typename cplx_of<R>::type C;
    BitInsert ins;
    int rc;
    const uint64_t tmask = 1ull << g->target_bits[0];
    Mat2 m2;
    memset(&m2, 0, sizeof(m2));
    uint32_t kind = g->kind;
    if (kind != QSV_PERMX) {
        if (!mat) return qsv_fail(ctx, QSV_E_INVALID, "matrix required");
        memcpy(m2.m, mat, sizeof(m2.m));
    }
    constexpr int U = 4;
    if (kind == QSV_DIAG1 && is_one(mat)) {
        // diag(1, d): touch only target=1 amplitudes -> fold the target bit into the selection
        qsv_gate_t sel = *g;
        sel.ctrl_mask |= tmask;
        sel.ctrl_value |= tmask;
        sel.n_targets = 0;
        if ((rc = build_insert(ctx, &sel, n_bits, &ins, false))) return rc;
        const uint64_t namps = 1ull << (n_bits - ins.n);
        const uint64_t blocks = (namps + 256 * U - 1) / (256 * U);
        k_phase1<R, U><<<(unsigned)blocks, 256, 0, s>>>((C*)state, namps, ins, mat[6], mat[7]);
        ctx->stats.amp_bytes += 2 * namps * sizeof(C);
    } else {
        static const bool use_shfl = getenv("QSV_NO_SHFL") == nullptr;  // measured 1-3 % faster than the pair kernel (profiles/r2_ab_shuffle.txt)
        if (use_shfl && kind == QSV_DENSE1 && g->ctrl_mask == 0 && g->target_bits[0] < 5 && n_bits >= 5) {
            const uint64_t namps = 1ull << n_bits;
            const uint64_t blocks = (namps + 256 * U - 1) / (256 * U);
            k_gate1_shfl<R, U><<<(unsigned)blocks, 256, 0, s>>>((C*)state, namps, (int)g->target_bits[0], m2);
            ctx->stats.amp_bytes += 2 * namps * sizeof(C);
            ctx->stats.kernel_launches++;
            return QSV_OK;
        }
        if ((rc = build_insert(ctx, g, n_bits, &ins, true))) return rc;
        const uint64_t npairs = 1ull << (n_bits - ins.n);
        const uint64_t blocks = (npairs + 256 * U - 1) / (256 * U);
        if (kind == QSV_PERMX)
            k_gate1<R, MODE_SWAP, U><<<(unsigned)blocks, 256, 0, s>>>((C*)state, npairs, ins, tmask, m2);
        else if (kind == QSV_DIAG1)
            k_gate1<R, MODE_DIAG, U><<<(unsigned)blocks, 256, 0, s>>>((C*)state, npairs, ins, tmask, m2);
        else
            k_gate1<R, MODE_DENSE, U><<<(unsigned)blocks, 256, 0, s>>>((C*)state, npairs, ins, tmask, m2);
        ctx->stats.amp_bytes += 4 * npairs * sizeof(C);
    }
    ctx->stats.kernel_launches++;
    return QSV_OK;
}

template <typename R>
static int launch_densek(qsv_ctx* ctx, void* state, int n_bits, const qsv_gate_t* g, const double* d_mat, cudaStream_t s) {
    typedef typename cplx_of<R>::type C;
    BitInsert ins;
    int rc;
    if ((rc = build_insert(ctx, g, n_bits, &ins, true))) return rc;
    TargetSet ts;
    ts.t = (int)g->n_targets;
    for (int j = 0; j < ts.t; ++j) ts.mask[j] = 1ull << g->target_bits[j];
    const uint64_t ngroups = 1ull << (n_bits - ins.n);
    const unsigned blocks = (unsigned)((ngroups + 127) / 128);
    switch (ts.t) {
        case 1: k_densek<R, 1><<<blocks, 128, 0, s>>>((C*)state, ngroups, ins, ts, d_mat); break;
        case 2: k_densek<R, 2><<<blocks, 128, 0, s>>>((C*)state, ngroups, ins, ts, d_mat); break;
        case 3: k_densek<R, 3><<<blocks, 128, 0, s>>>((C*)state, ngroups, ins, ts, d_mat); break;
        case 4: k_densek<R, 4><<<blocks, 128, 0, s>>>((C*)state, ngroups, ins, ts, d_mat); break;
        case 5: k_densek<R, 5><<<blocks, 128, 0, s>>>((C*)state, ngroups, ins, ts, d_mat); break;
        default: return qsv_fail(ctx, QSV_E_UNSUPPORTED, "dense gate with %d targets (max %d)", ts.t, QSV_MAX_TARGETS);
    }
    ctx->stats.kernel_launches++;
    ctx->stats.amp_bytes += 2 * (ngroups << ts.t) * sizeof(C);
    return QSV_OK;
}

template <typename R>
static int launch_rows2(qsv_ctx* ctx, void* state, uint64_t batch, uint64_t r0, uint64_t r1, const double* mat,
                        bool swap, cudaStream_t s) {
    typedef typename cplx_of<R>::type C;
    Mat2 m2;
    memset(&m2, 0, sizeof(m2));
    if (!swap) memcpy(m2.m, mat, sizeof(m2.m));
    const unsigned blocks = (unsigned)((batch + 255) / 256);
    if (swap)
        k_rows2<R, MODE_SWAP><<<blocks, 256, 0, s>>>((C*)state, batch, r0, r1, m2);
    else
        k_rows2<R, MODE_DENSE><<<blocks, 256, 0, s>>>((C*)state, batch, r0, r1, m2);
    ctx->stats.kernel_launches++;
    ctx->stats.amp_bytes += 4 * batch * sizeof(C);
    return QSV_OK;
}

// one gate, matrix given on the host (immediate) or already on the device (program)
static int apply_one(qsv_ctx* ctx, void* state, int n_bits, int batch_bits, const qsv_gate_t* g, const double* h_mat,
                     const double* d_mat, cudaStream_t s) {
    int rc = QSV_OK;
    const bool dbl = ctx->dtype == QSV_C128;
    switch (g->kind) {
        case QSV_DENSE1:
        case QSV_PERMX:
        case QSV_DIAG1:
            if (g->n_targets != 1) return qsv_fail(ctx, QSV_E_INVALID, "kind %u needs exactly one target", g->kind);
            rc = dbl ? launch_gate1<double>(ctx, state, n_bits, g, h_mat, s) : launch_gate1<float>(ctx, state, n_bits, g, h_mat, s);
            break;
        case QSV_DENSEK:
            if (g->n_targets < 1 || g->n_targets > QSV_MAX_TARGETS)
                return qsv_fail(ctx, QSV_E_UNSUPPORTED, "dense gate with %u targets (max %d)", g->n_targets, QSV_MAX_TARGETS);
            rc = dbl ? launch_densek<double>(ctx, state, n_bits, g, d_mat, s) : launch_densek<float>(ctx, state, n_bits, g, d_mat, s);
            break;
        case QSV_ROWS2: {
            const uint64_t dim = 1ull << (n_bits - batch_bits), batch = 1ull << batch_bits;
            if (g->row0 >= dim || g->row1 >= dim || g->row0 == g->row1)
                return qsv_fail(ctx, QSV_E_INVALID, "two-level rows (%llu,%llu) invalid for dim %llu", (unsigned long long)g->row0,
                                (unsigned long long)g->row1, (unsigned long long)dim);
            const bool swap = h_mat && is_zero(h_mat) && is_one(h_mat + 2) && is_one(h_mat + 4) && is_zero(h_mat + 6);
            rc = dbl ? launch_rows2<double>(ctx, state, batch, g->row0, g->row1, h_mat, swap, s)
                     : launch_rows2<float>(ctx, state, batch, g->row0, g->row1, h_mat, swap, s);
            break;
        }
        default:
            return qsv_fail(ctx, QSV_E_INVALID, "unknown gate kind %u", g->kind);
    }
    if (rc == QSV_OK) {
        ctx->stats.gates_applied++;
        QSV_CUDA(ctx, cudaGetLastError());
    }
    return rc;
}

// used by the fused plan for gates it cannot fuse (multi-target dense, row gates)
int qsv_apply_planned_single(qsv_ctx* ctx, void* state, size_t gate_index, cudaStream_t s) {
    const qsv_gate_t& g = ctx->gates[gate_index];
    const double* h = ctx->pool.data() + 2 * g.matrix_offset;
    const double* d = ctx->d_pool ? ctx->d_pool + 2 * g.matrix_offset : nullptr;
    return apply_one(ctx, state, ctx->prog_bits, ctx->prog_batch_bits, &g, g.kind == QSV_PERMX ? nullptr : h, d, s);
}

extern "C" {

const char* qsv_version(void) { return "qsv 0.1 (sm_100a)"; }

const char* qsv_last_error(const qsv_ctx* ctx) { return ctx ? ctx->err : g_create_err; }

int qsv_create(qsv_ctx** out, int device, int dtype) {
    if (!out) return qsv_fail(nullptr, QSV_E_INVALID, "out is NULL");
    if (dtype != QSV_C128 && dtype != QSV_C64) return qsv_fail(nullptr, QSV_E_INVALID, "dtype must be QSV_C128 or QSV_C64");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return qsv_fail(nullptr, QSV_E_CUDA, "no CUDA device available (%s); this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return qsv_fail(nullptr, QSV_E_INVALID, "device %d out of range [0,%d)", device, ndev);
    qsv_ctx* ctx = new qsv_ctx();
    ctx->device = device;
    ctx->dtype = dtype;
    memset(&ctx->stats, 0, sizeof(ctx->stats));
    e = cudaSetDevice(device);
    if (e == cudaSuccess) {
        ctx->d_mat_bytes = (size_t)64 * 64 * 16;
        e = cudaMalloc(&ctx->d_mat, ctx->d_mat_bytes);
    }
    if (e == cudaSuccess) {
        ctx->d_red_bytes = (size_t)(2 * CMP_BLOCKS + 2) * sizeof(double);
        e = cudaMalloc(&ctx->d_red, ctx->d_red_bytes);
    }
    if (e != cudaSuccess) {
        qsv_fail(nullptr, QSV_E_CUDA, "context set-up failed: %s", cudaGetErrorString(e));
        qsv_destroy(ctx);
        return QSV_E_CUDA;
    }
    *out = ctx;
    return QSV_OK;
}

void qsv_destroy(qsv_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->plan) sweep_plan_free(ctx->plan);
    if (ctx->rows_plan) rowslab_free(ctx->rows_plan);
    if (ctx->d_mat) cudaFree(ctx->d_mat);
    if (ctx->d_red) cudaFree(ctx->d_red);
    if (ctx->d_pool) cudaFree(ctx->d_pool);
    delete ctx;
}

int qsv_get_stats(const qsv_ctx* ctx, qsv_stats_t* out) {
    if (!ctx || !out) return QSV_E_INVALID;
    *out = ctx->stats;
    return QSV_OK;
}

int qsv_reset_stats(qsv_ctx* ctx) {
    if (!ctx) return QSV_E_INVALID;
    memset(&ctx->stats, 0, sizeof(ctx->stats));
    return QSV_OK;
}

int qsv_apply_gate(qsv_ctx* ctx, void* state, int n_bits, const qsv_gate_t* gate, const double* matrix, void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!state || !gate) return qsv_fail(ctx, QSV_E_INVALID, "state/gate is NULL");
    if (n_bits < 1 || n_bits > 62) return qsv_fail(ctx, QSV_E_INVALID, "n_bits %d out of range", n_bits);
    if (gate->kind == QSV_ROWS2) return qsv_fail(ctx, QSV_E_INVALID, "use qsv_apply_rows for row gates");
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    const double* d_mat = nullptr;
    if (gate->kind == QSV_DENSEK) {
        if (!matrix) return qsv_fail(ctx, QSV_E_INVALID, "matrix required");
        if (gate->n_targets < 1 || gate->n_targets > QSV_MAX_TARGETS)
            return qsv_fail(ctx, QSV_E_UNSUPPORTED, "dense gate with %u targets (max %d)", gate->n_targets, QSV_MAX_TARGETS);
        const size_t bytes = ((size_t)16 << (2 * gate->n_targets));
        // pageable source: the copy is staged before the call returns, so `matrix` may be reused
        QSV_CUDA(ctx, cudaMemcpyAsync(ctx->d_mat, matrix, bytes, cudaMemcpyHostToDevice, s));
        d_mat = ctx->d_mat;
    }
    return apply_one(ctx, state, n_bits, 0, gate, matrix, d_mat, s);
}

int qsv_apply_rows(qsv_ctx* ctx, void* state, uint64_t dim, uint64_t batch, const uint64_t* rows, int m,
                   const double* matrix, void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!state || !rows || !matrix) return qsv_fail(ctx, QSV_E_INVALID, "NULL argument");
    if (m < 1 || m > 16) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "row gate with m=%d rows (supported 1..16)", m);
    if (batch == 0) return QSV_OK;
    RowSet rs;
    rs.m = m;
    for (int i = 0; i < m; ++i) {
        if (rows[i] >= dim) return qsv_fail(ctx, QSV_E_INVALID, "row %llu >= dim %llu", (unsigned long long)rows[i], (unsigned long long)dim);
        for (int k = 0; k < i; ++k)
            if (rows[k] == rows[i]) return qsv_fail(ctx, QSV_E_INVALID, "duplicate row %llu", (unsigned long long)rows[i]);
        rs.row[i] = rows[i];
    }
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = QSV_OK;
    if (m == 2) {
        const bool swap = is_zero(matrix) && is_one(matrix + 2) && is_one(matrix + 4) && is_zero(matrix + 6);
        rc = ctx->dtype == QSV_C128 ? launch_rows2<double>(ctx, state, batch, rows[0], rows[1], matrix, swap, s)
                                    : launch_rows2<float>(ctx, state, batch, rows[0], rows[1], matrix, swap, s);
    } else {
        QSV_CUDA(ctx, cudaMemcpyAsync(ctx->d_mat, matrix, (size_t)16 * m * m, cudaMemcpyHostToDevice, s));
        const unsigned blocks = (unsigned)((batch + 255) / 256);
        if (ctx->dtype == QSV_C128)
            k_rows<double><<<blocks, 256, 0, s>>>((double2*)state, batch, rs, ctx->d_mat);
        else
            k_rows<float><<<blocks, 256, 0, s>>>((float2*)state, batch, rs, ctx->d_mat);
        ctx->stats.kernel_launches++;
        ctx->stats.amp_bytes += 2 * (uint64_t)m * batch * amp_size(ctx->dtype);
    }
    if (rc) return rc;
    ctx->stats.gates_applied++;
    QSV_CUDA(ctx, cudaGetLastError());
    return QSV_OK;
}

int qsv_upload_program(qsv_ctx* ctx, int n_bits, int batch_bits, const qsv_gate_t* gates, uint64_t n_gates,
                       const double* pool, uint64_t pool_elems) {
    if (!ctx) return QSV_E_INVALID;
    if (n_bits < 1 || n_bits > 62 || batch_bits < 0 || batch_bits > n_bits)
        return qsv_fail(ctx, QSV_E_INVALID, "n_bits=%d batch_bits=%d out of range", n_bits, batch_bits);
    if (n_gates && !gates) return qsv_fail(ctx, QSV_E_INVALID, "gates is NULL");
    for (uint64_t i = 0; i < n_gates; ++i) {
        const qsv_gate_t& g = gates[i];
        if (g.kind > QSV_ROWS2) return qsv_fail(ctx, QSV_E_INVALID, "gate %llu: unknown kind %u", (unsigned long long)i, g.kind);
        if (g.kind != QSV_ROWS2 && (g.n_targets < 1 || g.n_targets > QSV_MAX_TARGETS))  // (before the shift below uses it)
            return qsv_fail(ctx, QSV_E_UNSUPPORTED, "gate %llu: %u targets (max %d)", (unsigned long long)i, g.n_targets, QSV_MAX_TARGETS);
        uint64_t need = 0;
        if (g.kind == QSV_ROWS2) need = 4;
        else if (g.kind != QSV_PERMX) need = 1ull << (2 * g.n_targets);
        if (need && (g.matrix_offset > pool_elems || need > pool_elems - g.matrix_offset))
            return qsv_fail(ctx, QSV_E_INVALID, "gate %llu: matrix outside the pool", (unsigned long long)i);
        if (need && !pool) return qsv_fail(ctx, QSV_E_INVALID, "pool is NULL");
        if (g.kind == QSV_ROWS2) {
            const uint64_t dim = 1ull << (n_bits - batch_bits);
            if (g.row0 >= dim || g.row1 >= dim || g.row0 == g.row1)
                return qsv_fail(ctx, QSV_E_INVALID, "gate %llu: rows (%llu, %llu) must be distinct and below %llu", (unsigned long long)i,
                                (unsigned long long)g.row0, (unsigned long long)g.row1, (unsigned long long)dim);
        }
        if (g.kind == QSV_DIAG1) {
            // the fused path folds diagonals into a phase polynomial (angles only): a non-unit modulus would be applied by the
            // per-gate kernels and dropped by the fused ones, so it is refused here for both
            const double* m = pool + 2 * g.matrix_offset;
            const double r0 = m[0] * m[0] + m[1] * m[1], r1 = m[6] * m[6] + m[7] * m[7];
            if (!(std::fabs(r0 - 1.0) <= 1e-6) || !(std::fabs(r1 - 1.0) <= 1e-6) || m[2] != 0 || m[3] != 0 || m[4] != 0 || m[5] != 0)
                return qsv_fail(ctx, QSV_E_INVALID, "gate %llu: a diagonal record needs unit-modulus entries and a zero off-diagonal", (unsigned long long)i);
        }
        if (g.kind != QSV_ROWS2) {
            BitInsert ins;
            int rc = build_insert(ctx, &g, n_bits, &ins, true);
            if (rc) return rc;
            for (uint32_t j = 0; j < g.n_targets; ++j)
                if (g.target_bits[j] < batch_bits) return qsv_fail(ctx, QSV_E_INVALID, "gate %llu targets a batch bit", (unsigned long long)i);
            if (g.ctrl_mask & ((1ull << batch_bits) - 1ull)) return qsv_fail(ctx, QSV_E_INVALID, "gate %llu controls on a batch bit", (unsigned long long)i);
        }
    }
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->plan) { sweep_plan_free(ctx->plan); ctx->plan = nullptr; }
    if (ctx->rows_plan) { rowslab_free(ctx->rows_plan); ctx->rows_plan = nullptr; }
    if (ctx->d_pool) { cudaFree(ctx->d_pool); ctx->d_pool = nullptr; }
    ctx->gates.assign(gates, gates + n_gates);
    ctx->pool.assign(pool, pool + 2 * pool_elems);
    if (pool_elems) {
        QSV_CUDA(ctx, cudaMalloc(&ctx->d_pool, pool_elems * 16));
        QSV_CUDA(ctx, cudaMemcpy(ctx->d_pool, pool, pool_elems * 16, cudaMemcpyHostToDevice));
    }
    ctx->prog_bits = n_bits;
    ctx->prog_batch_bits = batch_bits;
    ctx->has_program = true;
    return QSV_OK;
}

int qsv_fuse(qsv_ctx* ctx, int max_tile_bits, uint64_t* out_sweeps) {
    if (!ctx) return QSV_E_INVALID;
    if (!ctx->has_program) return qsv_fail(ctx, QSV_E_NOPROGRAM, "qsv_fuse before qsv_upload_program");
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->rows_plan) { rowslab_free(ctx->rows_plan); ctx->rows_plan = nullptr; }
    {
        // streams made only of two-level / fully controlled gates: column-slab kernel, one pass for the whole program
        static const bool no_rows = getenv("QSV_NO_ROWSLAB") != nullptr;  // debugging aid
        RowSlabPlan* rp = nullptr;
        const int rc = no_rows ? QSV_E_UNSUPPORTED : rowslab_build(ctx, &rp);
        if (rc == QSV_OK) {
            if (ctx->plan) { sweep_plan_free(ctx->plan); ctx->plan = nullptr; }
            ctx->rows_plan = rp;
            if (out_sweeps) *out_sweeps = 1;
            return QSV_OK;
        }
        if (rc != QSV_E_UNSUPPORTED) return rc;
    }
    return sweep_plan_build(ctx, max_tile_bits, out_sweeps);
}

int qsv_fuse_to_layout(qsv_ctx* ctx, int max_tile_bits, const int* final_pos, int carry_exchange, uint64_t* out_sweeps) {
    if (!ctx) return QSV_E_INVALID;
    if (!ctx->has_program) return qsv_fail(ctx, QSV_E_NOPROGRAM, "qsv_fuse_to_layout before qsv_upload_program");
    if (!final_pos) return qsv_fail(ctx, QSV_E_INVALID, "final_pos is NULL");
    uint64_t seen = 0;
    for (int b = 0; b < ctx->prog_bits; ++b) {
        if (final_pos[b] < 0 || final_pos[b] >= ctx->prog_bits || (seen & (1ull << final_pos[b])))
            return qsv_fail(ctx, QSV_E_INVALID, "final_pos is not a permutation of the index bits");
        seen |= 1ull << final_pos[b];
        if (b < ctx->prog_batch_bits && final_pos[b] != b) return qsv_fail(ctx, QSV_E_INVALID, "batch bits cannot move");
    }
    for (const qsv_gate_t& g : ctx->gates)
        if (g.kind == QSV_ROWS2 || g.kind == QSV_DENSEK)
            return qsv_fail(ctx, QSV_E_UNSUPPORTED, "unsupported: a requested layout needs a program of one-target gates");
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->rows_plan) { rowslab_free(ctx->rows_plan); ctx->rows_plan = nullptr; }
    return sweep_plan_build(ctx, max_tile_bits, out_sweeps, final_pos, carry_exchange != 0);
}

int qsv_run_exchange(qsv_ctx* ctx, void* state, void* spare, void* const* peer_out, int g_bits, int rank, void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!ctx->plan) return qsv_fail(ctx, QSV_E_NOPROGRAM, "qsv_run_exchange before qsv_fuse_to_layout");
    if (!state || !peer_out) return qsv_fail(ctx, QSV_E_INVALID, "NULL argument");
    if (g_bits < 1 || g_bits > 3) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "g_bits=%d: 2, 4 or 8 ranks are supported", g_bits);
    if (rank < 0 || rank >= (1 << g_bits)) return qsv_fail(ctx, QSV_E_INVALID, "rank %d out of range", rank);
    uint64_t peers[8] = {0};
    for (int i = 0; i < (1 << g_bits); ++i) {
        if (!peer_out[i]) return qsv_fail(ctx, QSV_E_INVALID, "peer buffer %d is NULL", i);
        peers[i] = (uint64_t)peer_out[i];
    }
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    int flag = 0;
    return sweep_plan_run(ctx, state, spare, &flag, (cudaStream_t)stream, g_bits, rank, peers);
}

int qsv_plan_info(qsv_ctx* ctx, uint64_t* out_sweeps, uint64_t* out_of_place_sweeps) {
    if (!ctx) return QSV_E_INVALID;
    uint64_t s = 0, o = 0;
    if (ctx->rows_plan) s = 1;
    else if (ctx->plan) sweep_plan_count(ctx->plan, &s, &o);
    else return qsv_fail(ctx, QSV_E_NOPROGRAM, "qsv_plan_info before qsv_fuse");
    if (out_sweeps) *out_sweeps = s;
    if (out_of_place_sweeps) *out_of_place_sweeps = o;
    return QSV_OK;
}

int qsv_set_profiling(qsv_ctx* ctx, int on) {
    if (!ctx) return QSV_E_INVALID;
    ctx->profile = on != 0;
    return QSV_OK;
}

int qsv_get_sweep_profile(qsv_ctx* ctx, float* ms, int cap, int* n_out) {
    if (!ctx) return QSV_E_INVALID;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    return sweep_plan_collect_profile(ctx, ms, cap, n_out);
}

int qsv_run(qsv_ctx* ctx, void* state, void* spare, int* result_in_spare, void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!ctx->has_program) return qsv_fail(ctx, QSV_E_NOPROGRAM, "qsv_run before qsv_upload_program");
    if (!state) return qsv_fail(ctx, QSV_E_INVALID, "state is NULL");
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    if (result_in_spare) *result_in_spare = 0;
    if (ctx->rows_plan) return rowslab_run(ctx, ctx->rows_plan, state, s);
    if (ctx->plan) return sweep_plan_run(ctx, state, spare, result_in_spare, s);
    for (size_t i = 0; i < ctx->gates.size(); ++i) {
        const qsv_gate_t& g = ctx->gates[i];
        const double* h = ctx->pool.data() + 2 * g.matrix_offset;
        const double* d = ctx->d_pool ? ctx->d_pool + 2 * g.matrix_offset : nullptr;
        int rc = apply_one(ctx, state, ctx->prog_bits, ctx->prog_batch_bits, &g, g.kind == QSV_PERMX ? nullptr : h, d, s);
        if (rc) return rc;
    }
    return QSV_OK;
}

int qsv_init_basis_columns(qsv_ctx* ctx, void* state, int n_bits, int batch_bits, uint64_t col0, uint64_t ancilla_mask,
                           void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!state) return qsv_fail(ctx, QSV_E_INVALID, "state is NULL");
    if (n_bits < 1 || n_bits > 62 || batch_bits < 0 || batch_bits > n_bits) return qsv_fail(ctx, QSV_E_INVALID, "bad sizes");
    const int row_bits = n_bits - batch_bits;
    if (row_bits < 64 && (ancilla_mask >> row_bits)) return qsv_fail(ctx, QSV_E_INVALID, "ancilla mask exceeds the row bits");
    const int free_bits = row_bits - __builtin_popcountll(ancilla_mask);
    const uint64_t ncols = 1ull << batch_bits;
    if (col0 + ncols > (1ull << free_bits)) return qsv_fail(ctx, QSV_E_INVALID, "columns [%llu,%llu) exceed 2^%d basis states",
                                                            (unsigned long long)col0, (unsigned long long)(col0 + ncols), free_bits);
    BitInsert anc;
    anc.n = 0;
    anc.value = 0;
    for (int b = 0; b < row_bits; ++b)
        if (ancilla_mask & (1ull << b)) anc.pos[anc.n++] = (uint8_t)b;
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    QSV_CUDA(ctx, cudaMemsetAsync(state, 0, amp_size(ctx->dtype) << n_bits, s));
    const unsigned blocks = (unsigned)((ncols + 255) / 256);
    if (ctx->dtype == QSV_C128)
        k_basis<double><<<blocks, 256, 0, s>>>((double2*)state, ncols, batch_bits, col0, anc);
    else
        k_basis<float><<<blocks, 256, 0, s>>>((float2*)state, ncols, batch_bits, col0, anc);
    ctx->stats.kernel_launches++;
    QSV_CUDA(ctx, cudaGetLastError());
    return QSV_OK;
}

int qsv_project_ancilla(qsv_ctx* ctx, const void* in, void* out, int n_bits, int batch_bits, uint64_t ancilla_mask,
                        void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!in || !out) return qsv_fail(ctx, QSV_E_INVALID, "NULL buffer");
    if (n_bits < 1 || n_bits > 62 || batch_bits < 0 || batch_bits > n_bits) return qsv_fail(ctx, QSV_E_INVALID, "bad sizes");
    const int row_bits = n_bits - batch_bits;
    if (row_bits < 64 && (ancilla_mask >> row_bits)) return qsv_fail(ctx, QSV_E_INVALID, "ancilla mask exceeds the row bits");
    BitInsert anc;
    anc.n = 0;
    anc.value = 0;
    for (int b = 0; b < row_bits; ++b)
        if (ancilla_mask & (1ull << b)) anc.pos[anc.n++] = (uint8_t)(b + batch_bits);
    const uint64_t nout = 1ull << (n_bits - anc.n);
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    const unsigned blocks = (unsigned)std::min<uint64_t>((nout + 255) / 256, 148 * 16);
    if (ctx->dtype == QSV_C128)
        k_project<double><<<blocks, 256, 0, s>>>((const double2*)in, (double2*)out, nout, anc);
    else
        k_project<float><<<blocks, 256, 0, s>>>((const float2*)in, (float2*)out, nout, anc);
    ctx->stats.kernel_launches++;
    QSV_CUDA(ctx, cudaGetLastError());
    return QSV_OK;
}

static int psi_args(qsv_ctx* ctx, int n_bits, int batch_bits, uint64_t ancilla_mask, const double* psi, int shift, BitInsert* anc, Psi* p) {
    if (n_bits < 1 || n_bits > 62 || batch_bits < 0 || batch_bits > n_bits) return qsv_fail(ctx, QSV_E_INVALID, "bad sizes");
    const int row_bits = n_bits - batch_bits;
    if (row_bits < 64 && (ancilla_mask >> row_bits)) return qsv_fail(ctx, QSV_E_INVALID, "ancilla mask exceeds the row bits");
    if (!psi) return qsv_fail(ctx, QSV_E_INVALID, "psi is NULL");
    const double nrm = psi[0] * psi[0] + psi[1] * psi[1] + psi[2] * psi[2] + psi[3] * psi[3];
    if (!(std::fabs(nrm - 1.0) <= 1e-8)) return qsv_fail(ctx, QSV_E_INVALID, "psi is not normalised (|psi|^2 = %g)", nrm);
    if (__builtin_popcountll(ancilla_mask) > 16) return qsv_fail(ctx, QSV_E_TOOBIG, "more than 16 ancillas in a general state");
    anc->n = 0;
    anc->value = 0;
    for (int b = 0; b < row_bits; ++b)
        if (ancilla_mask & (1ull << b)) anc->pos[anc->n++] = (uint8_t)(b + shift);
    p->a_re = psi[0]; p->a_im = psi[1]; p->b_re = psi[2]; p->b_im = psi[3];
    return QSV_OK;
}

int qsv_init_basis_columns_psi(qsv_ctx* ctx, void* state, int n_bits, int batch_bits, uint64_t col0, uint64_t ancilla_mask,
                               const double psi[4], void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!state) return qsv_fail(ctx, QSV_E_INVALID, "state is NULL");
    BitInsert anc;
    Psi p;
    int rc = psi_args(ctx, n_bits, batch_bits, ancilla_mask, psi, 0, &anc, &p);
    if (rc) return rc;
    const int free_bits = n_bits - batch_bits - anc.n;
    const uint64_t ncols = 1ull << batch_bits;
    if (col0 + ncols > (1ull << free_bits)) return qsv_fail(ctx, QSV_E_INVALID, "columns [%llu,%llu) exceed 2^%d basis states",
                                                            (unsigned long long)col0, (unsigned long long)(col0 + ncols), free_bits);
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    QSV_CUDA(ctx, cudaMemsetAsync(state, 0, amp_size(ctx->dtype) << n_bits, s));
    const uint64_t work = ncols << anc.n;
    const unsigned blocks = (unsigned)((work + 255) / 256);
    if (ctx->dtype == QSV_C128)
        k_basis_psi<double><<<blocks, 256, 0, s>>>((double2*)state, ncols, batch_bits, col0, anc, p);
    else
        k_basis_psi<float><<<blocks, 256, 0, s>>>((float2*)state, ncols, batch_bits, col0, anc, p);
    ctx->stats.kernel_launches++;
    QSV_CUDA(ctx, cudaGetLastError());
    return QSV_OK;
}

int qsv_project_ancilla_psi(qsv_ctx* ctx, const void* in, void* out, int n_bits, int batch_bits, uint64_t ancilla_mask,
                            const double psi[4], void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!in || !out) return qsv_fail(ctx, QSV_E_INVALID, "NULL buffer");
    BitInsert anc;
    Psi p;
    int rc = psi_args(ctx, n_bits, batch_bits, ancilla_mask, psi, batch_bits, &anc, &p);
    if (rc) return rc;
    const uint64_t nout = 1ull << (n_bits - anc.n);
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    const unsigned blocks = (unsigned)std::min<uint64_t>((nout + 255) / 256, 148 * 16);
    if (ctx->dtype == QSV_C128)
        k_project_psi<double><<<blocks, 256, 0, s>>>((const double2*)in, (double2*)out, nout, anc, p);
    else
        k_project_psi<float><<<blocks, 256, 0, s>>>((const float2*)in, (float2*)out, nout, anc, p);
    ctx->stats.kernel_launches++;
    QSV_CUDA(ctx, cudaGetLastError());
    return QSV_OK;
}

int qsv_permute_bits(qsv_ctx* ctx, const void* in, void* out, int n_bits, const int* perm, void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!in || !out || !perm) return qsv_fail(ctx, QSV_E_INVALID, "NULL argument");
    if (in == out) return qsv_fail(ctx, QSV_E_INVALID, "qsv_permute_bits is out of place");
    if (n_bits < 1 || n_bits > 62) return qsv_fail(ctx, QSV_E_INVALID, "bad n_bits");
    BitPerm bp;
    bp.n = n_bits;
    uint64_t seen = 0;
    for (int i = 0; i < n_bits; ++i) {
        if (perm[i] < 0 || perm[i] >= n_bits || (seen & (1ull << perm[i]))) return qsv_fail(ctx, QSV_E_INVALID, "perm is not a permutation");
        seen |= 1ull << perm[i];
        bp.to[i] = (uint8_t)perm[i];
    }
    int fixed_low = 0;
    while (fixed_low < n_bits && perm[fixed_low] == fixed_low) ++fixed_low;
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint64_t n = 1ull << n_bits;
    {
        // large states: one gate-free tile sweep with a permuting store (coalesced both ways, HBM speed)
        const int rc = sweep_permute_run(ctx, in, out, n_bits, perm, s);
        if (rc == QSV_OK) {
            ctx->stats.kernel_launches++;
            ctx->stats.sweeps++;
            ctx->stats.amp_bytes += 2 * n * amp_size(ctx->dtype);
            return QSV_OK;
        }
        if (rc != QSV_E_UNSUPPORTED) return rc;
    }
    const unsigned blocks = (unsigned)std::min<uint64_t>((n + 255) / 256, 148 * 32);
    if (ctx->dtype == QSV_C128)
        k_permute_bits<double><<<blocks, 256, 0, s>>>((const double2*)in, (double2*)out, n, bp, fixed_low);
    else
        k_permute_bits<float><<<blocks, 256, 0, s>>>((const float2*)in, (float2*)out, n, bp, fixed_low);
    ctx->stats.kernel_launches++;
    ctx->stats.sweeps++;
    ctx->stats.amp_bytes += 2 * n * amp_size(ctx->dtype);
    QSV_CUDA(ctx, cudaGetLastError());
    return QSV_OK;
}

int qsv_permute_exchange(qsv_ctx* ctx, const void* in, void* const* peer_out, int g_bits, int rank, int n_bits, const int* perm,
                         void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!in || !peer_out || !perm) return qsv_fail(ctx, QSV_E_INVALID, "NULL argument");
    if (g_bits < 1 || g_bits > 3) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "g_bits=%d: 2, 4 or 8 ranks are supported", g_bits);
    if (rank < 0 || rank >= (1 << g_bits)) return qsv_fail(ctx, QSV_E_INVALID, "rank %d out of range", rank);
    if (n_bits < 9 || n_bits > 62 || n_bits < 2 * g_bits) return qsv_fail(ctx, QSV_E_UNSUPPORTED, "local shard of 2^%d amplitudes is too small for the fused exchange", n_bits);
    uint64_t seen = 0;
    uint64_t peers[8] = {0};
    for (int i = 0; i < n_bits; ++i) {
        if (perm[i] < 0 || perm[i] >= n_bits || (seen & (1ull << perm[i]))) return qsv_fail(ctx, QSV_E_INVALID, "perm is not a permutation");
        seen |= 1ull << perm[i];
    }
    for (int i = 0; i < (1 << g_bits); ++i) {
        if (!peer_out[i]) return qsv_fail(ctx, QSV_E_INVALID, "peer buffer %d is NULL", i);
        if (peer_out[i] == in) return qsv_fail(ctx, QSV_E_INVALID, "the exchange is out of place: peer buffer %d aliases the input", i);
        peers[i] = (uint64_t)peer_out[i];
    }
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = sweep_permute_run(ctx, in, peer_out[rank], n_bits, perm, s, g_bits, rank, peers);
    if (rc == QSV_E_UNSUPPORTED) return qsv_fail(ctx, rc, "unsupported: permutation needs more tile bits than available");
    if (rc) return rc;
    ctx->stats.kernel_launches++;
    ctx->stats.sweeps++;
    ctx->stats.amp_bytes += 2 * (amp_size(ctx->dtype) << n_bits);
    return QSV_OK;
}

int qsv_compare(qsv_ctx* ctx, const void* a, const void* b, uint64_t n_amps, double rtol, double atol, double* out_max,
                uint64_t* out_bad, void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (!a || !b) return qsv_fail(ctx, QSV_E_INVALID, "NULL buffer");
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    double res[2] = {0.0, 0.0};
    if (n_amps) {
        const int blocks = (int)std::min<uint64_t>((n_amps + 255) / 256, CMP_BLOCKS);
        double* partial = ctx->d_red;
        double* result = ctx->d_red + 2 * CMP_BLOCKS;
        if (ctx->dtype == QSV_C128)
            k_compare<double><<<blocks, 256, 0, s>>>((const double2*)a, (const double2*)b, n_amps, rtol, atol, partial);
        else
            k_compare<float><<<blocks, 256, 0, s>>>((const float2*)a, (const float2*)b, n_amps, rtol, atol, partial);
        k_compare_final<<<1, 256, 0, s>>>(partial, blocks, result);
        ctx->stats.kernel_launches += 2;
        QSV_CUDA(ctx, cudaGetLastError());
        QSV_CUDA(ctx, cudaMemcpyAsync(res, result, sizeof(res), cudaMemcpyDeviceToHost, s));
        QSV_CUDA(ctx, cudaStreamSynchronize(s));
    }
    if (out_max) *out_max = res[0];
    if (out_bad) *out_bad = (uint64_t)res[1];
    return QSV_OK;
}


int qsv_allprop(qsv_ctx* ctx, const void* a, const void* b, uint64_t n_amps, double rtol, double atol, int* out_ok,
                double* out_ratio, void* stream) {
    if (!ctx) return QSV_E_INVALID;
    if (n_amps && (!a || !b)) return qsv_fail(ctx, QSV_E_INVALID, "NULL buffer");
    cudaStream_t s = (cudaStream_t)stream;
    QSV_CUDA(ctx, cudaSetDevice(ctx->device));
    const bool dbl = ctx->dtype == QSV_C128;
    const size_t amp = amp_size(ctx->dtype);
    double ratio[2] = {0.0, 0.0};
    double res[2] = {0.0, 0.0};
    if (n_amps) {
        const int blocks = (int)std::min<uint64_t>((n_amps + 255) / 256, CMP_BLOCKS);
        double* partial = ctx->d_red;
        double* result = ctx->d_red + 2 * CMP_BLOCKS;
        unsigned long long* d_idx = reinterpret_cast<unsigned long long*>(result);
        unsigned long long idx = ~0ull;
        QSV_CUDA(ctx, cudaMemsetAsync(d_idx, 0xFF, sizeof(unsigned long long), s));
        if (dbl) k_first_nonzero<double><<<blocks, 256, 0, s>>>((const double2*)b, n_amps, 1e-8, d_idx);
        else k_first_nonzero<float><<<blocks, 256, 0, s>>>((const float2*)b, n_amps, 1e-8, d_idx);
        QSV_CUDA(ctx, cudaGetLastError());
        QSV_CUDA(ctx, cudaMemcpyAsync(&idx, d_idx, sizeof(idx), cudaMemcpyDeviceToHost, s));
        QSV_CUDA(ctx, cudaStreamSynchronize(s));
        ctx->stats.kernel_launches++;
        if (idx != ~0ull) {  // ratio = a[idx] / b[idx], as mfun.py:50
            double av[2], bv[2];
            if (dbl) {
                QSV_CUDA(ctx, cudaMemcpyAsync(av, (const char*)a + idx * amp, 16, cudaMemcpyDeviceToHost, s));
                QSV_CUDA(ctx, cudaMemcpyAsync(bv, (const char*)b + idx * amp, 16, cudaMemcpyDeviceToHost, s));
                QSV_CUDA(ctx, cudaStreamSynchronize(s));
            } else {
                float af[2], bf[2];
                QSV_CUDA(ctx, cudaMemcpyAsync(af, (const char*)a + idx * amp, 8, cudaMemcpyDeviceToHost, s));
                QSV_CUDA(ctx, cudaMemcpyAsync(bf, (const char*)b + idx * amp, 8, cudaMemcpyDeviceToHost, s));
                QSV_CUDA(ctx, cudaStreamSynchronize(s));
                av[0] = af[0]; av[1] = af[1]; bv[0] = bf[0]; bv[1] = bf[1];
            }
            const double den = bv[0] * bv[0] + bv[1] * bv[1];
            ratio[0] = (av[0] * bv[0] + av[1] * bv[1]) / den;
            ratio[1] = (av[1] * bv[0] - av[0] * bv[1]) / den;
        }
        // b ~ 0 everywhere: proportional iff a ~ 0 too (ratio 0), which the scaled compare with lambda = 0 checks
        if (dbl)
            k_compare_scaled<double><<<blocks, 256, 0, s>>>((const double2*)a, (const double2*)b, n_amps, ratio[0], ratio[1], rtol, atol, partial);
        else
            k_compare_scaled<float><<<blocks, 256, 0, s>>>((const float2*)a, (const float2*)b, n_amps, ratio[0], ratio[1], rtol, atol, partial);
        k_compare_final<<<1, 256, 0, s>>>(partial, blocks, result);
        ctx->stats.kernel_launches += 2;
        QSV_CUDA(ctx, cudaGetLastError());
        QSV_CUDA(ctx, cudaMemcpyAsync(res, result, sizeof(res), cudaMemcpyDeviceToHost, s));
        QSV_CUDA(ctx, cudaStreamSynchronize(s));
    }
    if (out_ok) *out_ok = res[1] == 0.0;
    if (out_ratio) {
        out_ratio[0] = ratio[0];
        out_ratio[1] = ratio[1];
    }
    return QSV_OK;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free compare.
//
// Same reduction as qsv_compare (k_compare + k_compare_final), but its scratch -- one (max, count)
// partial per block plus the two results -- lives in the fixed array below instead of memory that
// qsv_create obtained from cudaMalloc.  It is meant for hosts that cannot go to the CUDA allocator
// at the moment they need the answer: a simulation that has already sized its state vectors to
// fill the device (so any further cudaMalloc fails), or a service thread that must not call
// cudaMalloc/cudaFree because those calls synchronise the whole device and stall every other
// stream.  The array is part of the library image: each GPU that loads libqsv.so keeps its own
// 16 KiB copy until the process exits.
//
// `g_noalloc_scratch` as written in host code is only a handle for the runtime, not a device
// address, so the address a kernel can use is looked up once per device with
// cudaGetSymbolAddress (with that device current) and cached in g_noalloc_ptr[].
// ------------------------------------------------------------------------------------------------
#define NOALLOC_BLOCKS 1024
#define NOALLOC_MAX_DEVICES 64
static __device__ __align__(256) double g_noalloc_scratch[2 * NOALLOC_BLOCKS + 2];
static double* g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // per device: address of the array, NULL until first use

int qsv_compare_noalloc(int device, int dtype, const void* a, const void* b, uint64_t n_amps, double rtol, double atol,
                        double* out_max, uint64_t* out_bad, void* stream) {
    if (dtype != QSV_C128 && dtype != QSV_C64) return qsv_fail(nullptr, QSV_E_INVALID, "qsv_compare_noalloc: dtype must be QSV_C128 or QSV_C64");
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return qsv_fail(nullptr, QSV_E_INVALID, "qsv_compare_noalloc: device %d out of range", device);
    if (n_amps > QSV_NOALLOC_MAX_AMPS)
        return qsv_fail(nullptr, QSV_E_TOOBIG,
                        "qsv_compare_noalloc: %llu amplitudes exceed the fixed-scratch limit of %llu (QSV_NOALLOC_MAX_AMPS); "
                        "use qsv_compare with a context for larger arrays",
                        (unsigned long long)n_amps, (unsigned long long)QSV_NOALLOC_MAX_AMPS);
    if (n_amps && (!a || !b)) return qsv_fail(nullptr, QSV_E_INVALID, "qsv_compare_noalloc: NULL buffer");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return qsv_fail(nullptr, QSV_E_CUDA, "qsv_compare_noalloc: cudaSetDevice(%d) failed: %s", device, cudaGetErrorString(e));
    if (!g_noalloc_ptr[device]) {
        // first call on this device: resolve the device address of the static array (do this before any stream capture)
        double* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return qsv_fail(nullptr, QSV_E_CUDA, "qsv_compare_noalloc: cudaGetSymbolAddress failed on device %d: %s", device, cudaGetErrorString(e));
        g_noalloc_ptr[device] = p;
    }
    cudaStream_t s = (cudaStream_t)stream;
    double res[2] = {0.0, 0.0};
    if (n_amps) {
        double* partial = g_noalloc_ptr[device];
        double* result = partial + 2 * NOALLOC_BLOCKS;
        const int blocks = (int)std::min<uint64_t>((n_amps + 255) / 256, NOALLOC_BLOCKS);
        // every slot that k_compare_final reads is written by k_compare in this same call, so nothing
        // left over from an earlier call needs clearing
        if (dtype == QSV_C128)
            k_compare<double><<<blocks, 256, 0, s>>>((const double2*)a, (const double2*)b, n_amps, rtol, atol, partial);
        else
            k_compare<float><<<blocks, 256, 0, s>>>((const float2*)a, (const float2*)b, n_amps, rtol, atol, partial);
        k_compare_final<<<1, 256, 0, s>>>(partial, blocks, result);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaMemcpyAsync(res, result, sizeof(res), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) return qsv_fail(nullptr, QSV_E_CUDA, "qsv_compare_noalloc: %s", cudaGetErrorString(e));
    }
    if (out_max) *out_max = res[0];
    if (out_bad) *out_bad = (uint64_t)res[1];
    return QSV_OK;
}

}  // extern "C"
