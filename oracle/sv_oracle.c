/* ORACLE -- TEST INFRASTRUCTURE / CPU BASELINE ONLY.  Never linked into libqsv.so.
 *
 * Plain-C (OpenMP) restatement of the statevector semantics of the reference's gate
 * action: for every index x whose control bits equal their activation values,
 * out[x|targets=r] = sum_c M[r][c] in[x|targets=c]  (quompiler/construct/unitary.py:75-76
 * with the core of quompiler/construct/qontroller.py:112-120).  It executes the same
 * qsv_gate_t records the GPU executes, so bench.py can time "the same work on the host
 * cores".  Pinned against the reference through tests/test_oracle_golden.py
 * (test_c_oracle_matches_numpy_oracle).
 */
#include <complex.h>
#include <omp.h>
#include <stdint.h>
#include <string.h>

#include "qsv.h"

typedef double complex cd;

static inline uint64_t expand(uint64_t i, const int* pos, int n, uint64_t value) {
    for (int k = 0; k < n; ++k) {
        const int p = pos[k];
        const uint64_t low = i & ((1ull << p) - 1ull);
        i = ((i >> p) << (p + 1)) | low;
    }
    return i | value;
}

static void apply1(cd* st, int n_bits, const qsv_gate_t* g, const cd* m) {
    int pos[64], n = 0;
    const int t = g->target_bits[0];
    const uint64_t sel = g->ctrl_mask | (1ull << t);
    for (int b = 0; b < 64; ++b)
        if (sel & (1ull << b)) pos[n++] = b;
    const uint64_t npairs = 1ull << (n_bits - n), tm = 1ull << t;
    if (g->kind == QSV_PERMX) {
#pragma omp parallel for schedule(static)
        for (uint64_t i = 0; i < npairs; ++i) {
            const uint64_t x = expand(i, pos, n, g->ctrl_value);
            const cd a = st[x];
            st[x] = st[x | tm];
            st[x | tm] = a;
        }
        return;
    }
    const cd m00 = m[0], m01 = m[1], m10 = m[2], m11 = m[3];
#pragma omp parallel for schedule(static)
    for (uint64_t i = 0; i < npairs; ++i) {
        const uint64_t x = expand(i, pos, n, g->ctrl_value);
        const cd a0 = st[x], a1 = st[x | tm];
        st[x] = m00 * a0 + m01 * a1;
        st[x | tm] = m10 * a0 + m11 * a1;
    }
}

static void applyk(cd* st, int n_bits, const qsv_gate_t* g, const cd* m) {
    int pos[64], n = 0;
    const int t = (int)g->n_targets, D = 1 << t;
    uint64_t sel = g->ctrl_mask;
    for (int j = 0; j < t; ++j) sel |= 1ull << g->target_bits[j];
    for (int b = 0; b < 64; ++b)
        if (sel & (1ull << b)) pos[n++] = b;
    uint64_t off[32];
    for (int v = 0; v < D; ++v) {
        off[v] = 0;
        for (int j = 0; j < t; ++j)
            if ((v >> (t - 1 - j)) & 1) off[v] |= 1ull << g->target_bits[j];
    }
    const uint64_t ngroups = 1ull << (n_bits - n);
#pragma omp parallel for schedule(static)
    for (uint64_t i = 0; i < ngroups; ++i) {
        const uint64_t x = expand(i, pos, n, g->ctrl_value);
        cd in[32], out[32];
        for (int v = 0; v < D; ++v) in[v] = st[x | off[v]];
        for (int r = 0; r < D; ++r) {
            cd acc = 0;
            for (int c = 0; c < D; ++c) acc += m[r * D + c] * in[c];
            out[r] = acc;
        }
        for (int v = 0; v < D; ++v) st[x | off[v]] = out[v];
    }
}

/* Thread count of the parallel loops.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the timed CPU
 * baseline must not inherit that by accident, so bench.py sets the count explicitly and reports what it got. */
void orc_set_threads(int n) {
    if (n > 0) omp_set_num_threads(n);
}
int orc_get_threads(void) { return omp_get_max_threads(); }

/* Applies `n_gates` records (time order) to a complex128 state of 2^n_bits amplitudes. */
int orc_run(double* state, int n_bits, int batch_bits, const qsv_gate_t* gates, uint64_t n_gates, const double* pool) {
    cd* st = (cd*)state;
    for (uint64_t i = 0; i < n_gates; ++i) {
        const qsv_gate_t* g = &gates[i];
        const cd* m = (const cd*)pool + g->matrix_offset;
        if (g->kind == QSV_ROWS2) {
            const uint64_t B = 1ull << batch_bits;
            cd* r0 = st + g->row0 * B;
            cd* r1 = st + g->row1 * B;
#pragma omp parallel for schedule(static)
            for (uint64_t j = 0; j < B; ++j) {
                const cd a0 = r0[j], a1 = r1[j];
                r0[j] = m[0] * a0 + m[1] * a1;
                r1[j] = m[2] * a0 + m[3] * a1;
            }
        } else if (g->n_targets == 1) {
            apply1(st, n_bits, g, m);
        } else if (g->n_targets <= 5) {
            applyk(st, n_bits, g, m);
        } else {
            return -1;
        }
    }
    return 0;
}
