/* qsv_hostcheck.h -- TEST INFRASTRUCTURE exported by libqsv_hostcheck.so, a library of its own built from the product's
 * objects plus csrc/qsv_hostcheck.cu; the product library libqsv.so does not contain it.
 *
 * One entry point that interprets the FUSED PLAN of a gate program on the host, so that the fusion pass
 * (host code: stage selection, phase polynomials, SWAP relabelling, launch budgets, FFT-shape conversion,
 * every address table the kernel consumes) can be checked against the oracle on a machine without a GPU.
 * The per-op arithmetic is the same __host__ __device__ source the CUDA kernel k_sweep is compiled from;
 * the tile loop is a plain sequential restatement of the kernel's.
 *
 * It is NOT a compute path and not a fallback: nothing in the quantum-simulations_b200 package calls it
 * (only tests/ do), it never touches a CUDA device, and it refuses registers above 22 bits.
 * No counterpart in the reference (the reference has no fusion pass).
 */
#ifndef QSV_HOSTCHECK_H
#define QSV_HOSTCHECK_H
#include "qsv.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Build the fused plan for `gates` (one-target kinds only: QSV_DENSE1 / QSV_PERMX / QSV_DIAG1, same record
 * layout and pool as qsv_upload_program) over an n_bits register, tile size max_tile_bits (0 = default), and
 * run it on the HOST buffer `state` (2^n_bits interleaved complex values of `dtype`), in place.
 * With state == NULL only the plan is built (n_bits up to 40): the counts below describe a full-size plan.
 * out_sweeps / out_fft_sweeps (may be NULL): number of fused launches / of those in the FFT-shaped image.
 * Returns a QSV_* status; on failure the message is copied to err[0..err_len). */
int qsv_hostcheck_fused(int dtype, int n_bits, const qsv_gate_t* gates, uint64_t n_gates, const double* pool, uint64_t pool_elems,
                        int max_tile_bits, void* state, uint64_t* out_sweeps, uint64_t* out_fft_sweeps, char* err, uint64_t err_len);

/* The same with a requested end layout (qsv_fuse_to_layout): the result is left with logical bit b at physical position
 * final_pos[b]; carry_exchange != 0 additionally demands that the last sweep can carry an exchange (out-of-place store
 * in a cp.async image) and fails otherwise.  final_pos == NULL: canonical layout.                                    */
int qsv_hostcheck_fused_layout(int dtype, int n_bits, const qsv_gate_t* gates, uint64_t n_gates, const double* pool, uint64_t pool_elems,
                               int max_tile_bits, const int* final_pos, int carry_exchange, void* state, uint64_t* out_sweeps,
                               uint64_t* out_fft_sweeps, char* err, uint64_t err_len);

#ifdef __cplusplus
}
#endif
#endif
