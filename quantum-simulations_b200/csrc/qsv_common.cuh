// Shared declarations of libqsv: context, complex helpers, index expansion.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>

#include "qsv.h"

#define QSV_MAX_INSERT 48  // max (controls + targets) of one gate

// ---- complex helpers ------------------------------------------------------------------------
template <typename R> struct cplx_of;
template <> struct cplx_of<double> { typedef double2 type; };
template <> struct cplx_of<float>  { typedef float2 type; };

template <typename C, typename R>
__host__ __device__ __forceinline__ C cmake(R x, R y) { C c; c.x = x; c.y = y; return c; }

// a*b, written so that multiplication by an exact (1,0) returns a unchanged
template <typename C>
__host__ __device__ __forceinline__ C cmul(const C a, const C b) {
    C r;
    r.x = a.x * b.x - a.y * b.y;
    r.y = a.x * b.y + a.y * b.x;
    return r;
}
// acc + a*b
template <typename C>
__host__ __device__ __forceinline__ C cfma(const C a, const C b, const C acc) {
    C r;
    r.x = fma(a.x, b.x, fma(-a.y, b.y, acc.x));
    r.y = fma(a.x, b.y, fma(a.y, b.x, acc.y));
    return r;
}

// ---- index expansion: insert zero bits at sorted positions, then OR a constant ----------------
struct BitInsert {
    int n;
    uint8_t pos[QSV_MAX_INSERT];  // ascending
    uint64_t value;               // OR-ed after expansion (control activation values)
};

__host__ __device__ __forceinline__ uint64_t bit_expand(uint64_t i, const BitInsert& ins) {
    for (int k = 0; k < ins.n; ++k) {
        const int p = ins.pos[k];
        const uint64_t low = i & ((1ull << p) - 1ull);
        i = ((i >> p) << (p + 1)) | low;
    }
    return i | ins.value;
}

struct Mat2 { double m[8]; };  // row-major (re,im) x 4

// ---- fused-plan types (qsv_sweep_ops.cuh) -----------------------------------------------------------
struct SweepPlan;    // opaque here (qsv_sweep_ops.cuh)
struct RowSlabPlan;  // opaque here (qsv_rows.cu)

struct qsv_ctx {
    int device = 0;
    int dtype = QSV_C128;
    char err[512] = {0};
    qsv_stats_t stats;
    // device scratch
    double* d_mat = nullptr;       // staged matrix for immediate dense-k / rows gates
    size_t d_mat_bytes = 0;
    double* d_red = nullptr;       // reduction partials for qsv_compare
    size_t d_red_bytes = 0;
    // program
    bool has_program = false;
    int prog_bits = 0, prog_batch_bits = 0;
    std::vector<qsv_gate_t> gates;
    std::vector<double> pool;      // host copy, complex128 interleaved
    double* d_pool = nullptr;      // device copy
    // fused plan
    SweepPlan* plan = nullptr;
    RowSlabPlan* rows_plan = nullptr;  // set instead of `plan` when the whole program is row-pair operations
    bool profile = false;          // record CUDA events around every fused sweep
};

int qsv_fail(qsv_ctx* ctx, int code, const char* fmt, ...);
#define QSV_CUDA(ctx, call)                                                                      \
    do {                                                                                         \
        cudaError_t e__ = (call);                                                                \
        if (e__ != cudaSuccess)                                                                  \
            return qsv_fail((ctx), QSV_E_CUDA, "%s failed: %s (%s:%d)", #call,                 \
                            cudaGetErrorString(e__), __FILE__, __LINE__);                        \
    } while (0)

static inline size_t amp_size(int dtype) { return dtype == QSV_C128 ? 16 : 8; }

// implemented in qsv_sweep.cu
void sweep_plan_free(SweepPlan* p);
int sweep_plan_build(qsv_ctx* ctx, int max_tile_bits, uint64_t* out_sweeps, const int* final_pos = nullptr, bool carry_exchange = false);
// xg > 0: the plan's last sweep stores into the peers' buffers (peers[0 .. 2^xg)), see SweepDesc::xg
int sweep_plan_run(qsv_ctx* ctx, void* state, void* spare, int* result_in_spare, cudaStream_t stream, int xg = 0, int xrank = 0,
                   const uint64_t* peers = nullptr);
int sweep_plan_collect_profile(qsv_ctx* ctx, float* ms, int cap, int* n_out);
void sweep_plan_count(const SweepPlan* plan, uint64_t* sweeps, uint64_t* out_of_place);
// implemented in qsv_rows.cu
void rowslab_free(RowSlabPlan* p);
int rowslab_build(qsv_ctx* ctx, RowSlabPlan** out);
int rowslab_run(qsv_ctx* ctx, RowSlabPlan* plan, void* state, cudaStream_t stream);
int sweep_permute_run(qsv_ctx* ctx, const void* in, void* out, int n_bits, const int* perm, cudaStream_t stream, int xg = 0,
                      int xrank = 0, const uint64_t* peers = nullptr);
