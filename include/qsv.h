/*
 * qsv.h -- C ABI of the B200 statevector engine (libqsv.so).
 *
 * The reference (wanshoupu/quantum-simulations, package `quompiler`) is pure
 * Python and has NO FFI of its own: its plug-in boundary for this path is the
 * Python ABC `CircuitBuilder` (quompiler/circuits/qbuilder.py:9-49) driven by
 * `QRenderer.render` (quompiler/circuits/render.py:12-32), and its arithmetic is
 * dense numpy products (quompiler/construct/cgate.py:104-121,
 * quompiler/construct/unitary.py:56-77).  Each entry point below therefore cites
 * the reference code whose work it replaces.  The Python binding a maintainer
 * would add on the reference side is a ctypes stub; see INTEGRATION.md.
 *
 * Conventions
 *  - A state is a device array of `2^n_bits` complex amplitudes, interleaved
 *    (re, im), complex128 (16 B) or complex64 (8 B), index bit 0 = least
 *    significant.  A batch of 2^b states is stored batch-innermost, i.e. as ONE
 *    array with n_bits = n_qubits + b whose low b bits no gate touches.
 *  - Qubit at sorted-register position p of N owns logical bit N-1-p
 *    (big-endian, quompiler/construct/qontroller.py:7-15); the caller adds b.
 *  - Gate action (quompiler/construct/unitary.py:75-76 + qontroller.py:112-120):
 *    for every index x with (x & ctrl_mask) == ctrl_value,
 *        out[x | targets=r] = sum_c M[r][c] * in[x | targets=c],
 *    M = matrix()*phase() row-major complex128, row/col bits taken over
 *    target_bits[] with target_bits[0] the most significant.
 *  - All pointers named `state`, `in`, `out`, `a`, `b` are DEVICE pointers owned
 *    by the caller; `matrix`, `gates`, `pool`, `rows`, `perm` and every `out_*`
 *    scalar are HOST pointers.  `stream` is a cudaStream_t passed as void*.
 *  - Every function returns 0 on success or a negative QSV_E_* code; the message
 *    is available from qsv_last_error().  No exception crosses the ABI.
 *  - A context is bound to one device and is not thread-safe; use one context
 *    per host thread / stream.
 */
#ifndef QSV_H
#define QSV_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QSV_OK            0
#define QSV_E_INVALID    -1   /* bad argument                                   */
#define QSV_E_CUDA       -2   /* a CUDA runtime call failed                     */
#define QSV_E_NOPROGRAM  -3   /* qsv_run before qsv_upload_program              */
#define QSV_E_UNSUPPORTED -4  /* valid but not implemented for these sizes      */
#define QSV_E_TOOBIG     -5   /* exceeds a documented fixed capacity            */

#define QSV_C128 0
#define QSV_C64  1

/* gate kinds */
#define QSV_DENSE1   0  /* controlled dense 2x2 on one target bit               */
#define QSV_PERMX    1  /* controlled X: pure amplitude swap, no arithmetic     */
#define QSV_DIAG1    2  /* controlled diag(d0,d1) on one target bit             */
#define QSV_DENSEK   3  /* controlled dense 2^t x 2^t, 2 <= t <= QSV_MAX_TARGETS */
#define QSV_ROWS2    4  /* two-level (Givens) unitary on rows row0,row1 of a
                           [dim][batch] array; dim need not be a power of two   */

#define QSV_MAX_TARGETS 5

/* POD gate record, 64 bytes.  The flattened form of one CtrlGate
 * (quompiler/construct/cgate.py:24-52) or two-level UnitaryM
 * (quompiler/construct/unitary.py:26-46).                                       */
typedef struct qsv_gate {
    uint32_t kind;
    uint32_t n_targets;
    uint8_t  target_bits[8];   /* bit positions, most significant matrix bit first */
    uint64_t ctrl_mask;
    uint64_t ctrl_value;
    uint64_t row0, row1;       /* QSV_ROWS2 only                                   */
    uint64_t matrix_offset;    /* in complex128 elements into the matrix pool      */
    uint64_t reserved;
} qsv_gate_t;

typedef struct qsv_stats {
    uint64_t kernel_launches;  /* kernels of this library launched via this context */
    uint64_t sweeps;           /* passes over the full state                        */
    uint64_t amp_bytes;        /* algorithmic bytes: 2 * touched amplitudes * size  */
    uint64_t gates_applied;    /* source gates executed                             */
    uint64_t fused_sweeps;     /* sweeps executed by the fused tile kernel          */
    uint64_t reserved[3];
} qsv_stats_t;

typedef struct qsv_ctx qsv_ctx;

/* Library / context ------------------------------------------------------------ */
const char* qsv_version(void);
/* message of the last failure on `ctx` (or of the last failed qsv_create if NULL) */
const char* qsv_last_error(const qsv_ctx* ctx);
/* Creates a context on CUDA device `device` for amplitudes of type `dtype`.
 * Allocates a small device scratch area (reduction partials, staged matrices).
 * Replaces: nothing in the reference (numpy needs no context).                    */
int  qsv_create(qsv_ctx** out, int device, int dtype);
void qsv_destroy(qsv_ctx* ctx);
int  qsv_get_stats(const qsv_ctx* ctx, qsv_stats_t* out);
int  qsv_reset_stats(qsv_ctx* ctx);

/* Single gates ------------------------------------------------------------------ */
/* Applies one gate in place.  `matrix` = host complex128 row-major (2^t)^2 values
 * (ignored for QSV_PERMX).  Replaces one factor of the dense product in
 * CtrlGate.__matmul__ (quompiler/construct/cgate.py:104-121) together with the
 * lifts it needs: expand/np.kron (cgate.py:195-218) and promote/ctrl_expand
 * (cgate.py:307-331, quompiler/utils/inter_product.py:366-437) are implicit in the
 * stride and control-mask index math.                                              */
int qsv_apply_gate(qsv_ctx* ctx, void* state, int n_bits, const qsv_gate_t* gate,
                   const double* matrix, void* stream);

/* Applies an m x m unitary to rows `rows[0..m)` of a row-major [dim][batch] array
 * (m <= 16).  Replaces UnitaryM.inflate + dense product for a UnitaryM with an
 * arbitrary core (quompiler/construct/unitary.py:56-77); dim may be any size
 * (the reference's tests use dimension 3).                                         */
int qsv_apply_rows(qsv_ctx* ctx, void* state, uint64_t dim, uint64_t batch,
                   const uint64_t* rows, int m, const double* matrix, void* stream);

/* Programs ----------------------------------------------------------------------- */
/* Copies a gate stream (circuit TIME order: gates[0] acts first, i.e. the order
 * BytecodeRevIter yields, quompiler/construct/bytecode.py:69-88) and its matrix
 * pool to the device.  QSV_ROWS2 records address rows of a [2^(n_bits-batch_bits)]
 * [2^batch_bits] array.                                                            */
int qsv_upload_program(qsv_ctx* ctx, int n_bits, int batch_bits, const qsv_gate_t* gates,
                       uint64_t n_gates, const double* pool, uint64_t pool_elems);
/* Gate-fusion pass: packs the uploaded stream into shared-memory-staged tile
 * sweeps of at most `max_tile_bits` local bits each.  No reference counterpart
 * (nearest analogue: the peephole SlidingWindowCombiner,
 * quompiler/optimize/basic_optimizer.py:11-41).  Returns the number of sweeps via
 * out_sweeps (may be NULL).                                                        */
int qsv_fuse(qsv_ctx* ctx, int max_tile_bits, uint64_t* out_sweeps);
/* Runs the uploaded program on `state` (fused plan if qsv_fuse was called, else one
 * kernel per gate).  `spare` is a second device buffer of the same size used by
 * out-of-place (bit-permuting) sweeps, or NULL to forbid them.  On return
 * *result_in_spare (may be NULL) is 1 if the final state lives in `spare`.
 * Replaces functools.reduce(lambda a, b: a @ b, leaves) applied to a state
 * (tests/qompiler/circuits/test_qompiler.py:79-82).                                 */
int qsv_run(qsv_ctx* ctx, void* state, void* spare, int* result_in_spare, void* stream);

/* qsv_fuse with a requested END layout: after the program, logical index bit b sits at physical position
 * final_pos[b] (a permutation of 0..n_bits-1 that keeps the batch bits; identity = what qsv_fuse delivers).  The
 * correcting permutation rides on a compute sweep's out-of-place store where the tile allows, else one gate-free
 * permuting sweep is appended.  A sharded run asks for the layout its next exchange step needs (victim bits parked in
 * the top positions).  With carry_exchange != 0 the plan's LAST sweep is guaranteed to store out of place, in a kernel
 * image whose store can fan out over peer buffers: qsv_run_exchange then makes that store the exchange itself.
 * One-target gate kinds only.  No reference counterpart (the reference has neither layouts nor ranks).           */
int qsv_fuse_to_layout(qsv_ctx* ctx, int max_tile_bits, const int* final_pos, int carry_exchange, uint64_t* out_sweeps);
/* qsv_run for a plan built with carry_exchange: every sweep but the last runs locally between `state` and `spare`
 * (pass both); the last one stores each amplitude straight into the rank that owns it after the exchange "the g_bits
 * global bits trade places with the top g_bits local bits": the top g_bits of the (permuted) output index select
 * peer_out[r], this rank's id takes their place in the address there (as in qsv_permute_exchange).  peer_out[r] =
 * peer-mapped device address of rank r's destination buffer (peer_out[rank] = this rank's own).  The caller brackets
 * the call with barriers across the ranks: nobody may still read its destination buffer, and all stores must have
 * landed before anyone uses the result.  ONE kernel computes the segment's last gates and performs the collective.  */
int qsv_run_exchange(qsv_ctx* ctx, void* state, void* spare, void* const* peer_out, int g_bits, int rank, void* stream);

/* Shape of the fused plan built by qsv_fuse: number of fused sweeps and how many of them store out of place (each of
 * those makes qsv_run's result change buffers once).  Ranks of a sharded run use it to agree, before the first run, on
 * which of their two buffers will hold a segment's result.  No reference counterpart.                          */
int qsv_plan_info(qsv_ctx* ctx, uint64_t* out_sweeps, uint64_t* out_of_place_sweeps);

/* Profiling: when on, every fused sweep launched by qsv_run is bracketed by CUDA events on the
 * launch stream; qsv_get_sweep_profile waits for them and returns their durations in launch
 * order (*n_out = how many were recorded, at most `cap` are written).  No reference
 * counterpart (the reference has no timers on this path).                                */
int qsv_set_profiling(qsv_ctx* ctx, int on);
int qsv_get_sweep_profile(qsv_ctx* ctx, float* ms, int cap, int* n_out);

/* State preparation / read-out ---------------------------------------------------- */
/* Fills a [2^(n_bits-batch_bits)][2^batch_bits] array with basis columns: column j
 * holds |e(col0 + j)>, where e() deposits the bits of its argument into the row-index
 * bits NOT in ancilla_mask (ancillas start in |0>, cgate.py:289-305).  Replaces
 * np.eye(...) (quompiler/construct/unitary.py:75).                                  */
int qsv_init_basis_columns(qsv_ctx* ctx, void* state, int n_bits, int batch_bits,
                           uint64_t col0, uint64_t ancilla_mask, void* stream);
/* Gathers the rows whose ancilla bits are all 0: out is
 * [2^(row_bits - popcount(ancilla_mask))][2^batch_bits].  Replaces CtrlGate.dela /
 * trace / qproject with |0> (cgate.py:228-305, inter_product.py:318-363).           */
int qsv_project_ancilla(qsv_ctx* ctx, const void* in, void* out, int n_bits, int batch_bits,
                        uint64_t ancilla_mask, void* stream);
/* The same two steps with every ancilla in the single-qubit pure state psi = (psi[0] + i psi[1]) |0> +
 * (psi[2] + i psi[3]) |1> instead of |0>: column j starts as |e(col0 + j)> (x) psi (x) ... (x) psi, and the read-out
 * contracts every ancilla with <psi|: out[r] = sum_v conj(prod_i psi_{v_i}) in[r, ancillas = v].  Together they are
 * the sandwich <psi| U |psi> of CtrlGate.dela(state) / trace(qubit, state) for a state other than |0>
 * (cgate.py:228-305; qproject, inter_product.py:357-363).  At most 16 ancillas.          */
int qsv_init_basis_columns_psi(qsv_ctx* ctx, void* state, int n_bits, int batch_bits, uint64_t col0,
                               uint64_t ancilla_mask, const double psi[4], void* stream);
int qsv_project_ancilla_psi(qsv_ctx* ctx, const void* in, void* out, int n_bits, int batch_bits,
                            uint64_t ancilla_mask, const double psi[4], void* stream);
/* out[y] = in[x] where bit perm[i] of y equals bit i of x.  Replaces
 * Permuter.bitsortall + np.ix_ in CtrlGate.sorted (cgate.py:181-183,
 * quompiler/utils/permute.py:60-121); also the local half of a qubit remap.         */
int qsv_permute_bits(qsv_ctx* ctx, const void* in, void* out, int n_bits, const int* perm,
                     void* stream);
/* Multi-GPU exchange fused with the local bit permutation (one kernel, no NCCL): the shard `in` of 2^n_bits
 * amplitudes on rank `rank` is bit-permuted (perm as in qsv_permute_bits) and every result is stored straight into
 * the destination rank's buffer over NVLink: the top g_bits bits of the permuted index select the rank
 * (peer_out[0 .. 2^g_bits), peer-mapped device pointers to every rank's destination buffer, peer_out[rank] being this
 * rank's own), and `rank` takes their place in the address there -- i.e. the all-to-all "global bits <-> top local
 * bits" of the sharded state vector.  The caller synchronises the ranks before the destination buffers are read
 * (and before they are written again).  No reference counterpart (the reference is single process).            */
int qsv_permute_exchange(qsv_ctx* ctx, const void* in, void* const* peer_out, int g_bits, int rank, int n_bits,
                         const int* perm, void* stream);

/* *out_max = max_i |a[i] - b[i]| (complex modulus) and *out_bad = number of entries
 * with |a-b| > atol + rtol*|b|.  Synchronises `stream`.  Replaces np.allclose(v, u)
 * (tests/qompiler/circuits/test_qompiler.py:82, quompiler/utils/mfun.py:34-52).      */
int qsv_compare(qsv_ctx* ctx, const void* a, const void* b, uint64_t n_amps, double rtol,
                double atol, double* out_max, uint64_t* out_bad, void* stream);

/* Equality up to one complex scalar: *out_ok = 1 iff a == lambda * b within |a - lambda b| <= atol + rtol |lambda b|,
 * lambda = a[i]/b[i] at the first i with |b[i]| > 1e-8 (lambda = 0 if there is none), returned in out_ratio[2] (re, im).
 * Synchronises `stream`.  Replaces allprop (quompiler/utils/mfun.py:34-52), the check the reference's PRINCIPAL-level
 * builder tests use (tests/qompiler/circuits/test_cirq_circuit.py:52).                                           */
int qsv_allprop(qsv_ctx* ctx, const void* a, const void* b, uint64_t n_amps, double rtol, double atol,
                int* out_ok, double* out_ratio, void* stream);

/* Same result as qsv_compare, for callers that must not allocate.  No counterpart in the
 * reference.  It needs no context and calls no cudaMalloc/cudaFree (nor any routine that does):
 * its scratch is a fixed array inside the library, about 16 KiB on every GPU that loads
 * libqsv.so, held until the process exits.  Intended for hosts that have already given all device
 * memory to their state vectors, or that may not enter the CUDA allocator because it synchronises
 * the whole device.
 *  - At most QSV_NOALLOC_MAX_AMPS amplitudes per call; larger sizes return QSV_E_TOOBIG (use
 *    qsv_compare).  n_amps = 0 returns max = 0, bad = 0.
 *  - Makes `device` current.  The first call on each device looks up the array's device address
 *    (a one-off cudaGetSymbolAddress); make that first call before starting any stream capture.
 *    Not thread-safe against another first call on the same device.
 *  - That first call is also where the CUDA runtime may create the device's context and load this
 *    library's code image (kernels and the 16 KiB array) onto the device -- work the RUNTIME does with
 *    device memory of its own.  On a device that is already full it can therefore fail with an
 *    out-of-memory error even though this function allocates nothing: make one warm-up call per
 *    device at start-up, before handing the memory out (n_amps = 0 is enough; it performs the whole
 *    set-up and returns max = 0, bad = 0).  Later calls do no set-up.
 *  - All calls on a device share the one array.  Each call overwrites every slot it later
 *    reads, so contents left by the previous call are harmless; but two calls whose kernels
 *    overlap on the GPU (different streams, or different host threads) race on it and give
 *    wrong answers -- serialise them.
 *  - Synchronises `stream` before returning.  Errors: message via qsv_last_error(NULL).            */
#define QSV_NOALLOC_MAX_AMPS (1ull << 28)
int qsv_compare_noalloc(int device, int dtype, const void* a, const void* b, uint64_t n_amps,
                        double rtol, double atol, double* out_max, uint64_t* out_bad, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* QSV_H */
