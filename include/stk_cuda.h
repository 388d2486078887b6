/* stk_cuda.h — C-ABI of libstk_cuda.so, the B200-native sampling backend for
 * Stochaskell (the `Language.CUDA` counterpart of `Language.CC`).
 *
 * What this boundary replaces in the reference (davidar/stochaskell):
 *   - compileCC / runCC            stochaskell/src/Language/CC.hs:459-498
 *       (IR -> C++ text -> g++ -> one process per transition, state through
 *        stdin/stdout text).  Here: IR -> CUDA text (stkc / Language.CUDA) ->
 *        NVRTC sm_100a cubin (SHA1-keyed cache like CC.hs:471-479) -> one
 *        resident module, state resident in HBM, thousands of chains per call.
 *   - runStep / simulate           stochaskell/src/Data/Program.hs:1350-1359, 1038
 *   - hmcStan / hmcStanInit / runStan  stochaskell/src/Language/Stan.hs:443-500
 *       (static-HMC semantics of cmdstan .../hmc/static/base_static_hmc.hpp:42-72)
 *
 * There is no FFI in the reference (its boundary is process + text), so these
 * entry points are what `foreign import ccall` in haskell/Language/CUDA.hs
 * binds; see INTEGRATION.md for the reference-side stubs.
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error; the message is in
 *     stk_last_error() (thread-local).  No exception crosses the boundary.
 *   - handles are opaque; host pointers are borrowed for the call only; device
 *     memory is owned by the library.
 *   - one stk_ctx per GPU; a ctx is not thread-safe (serialise per ctx).
 *   - chain state is structure-of-arrays: field f with per-chain length L is
 *     stored [L][n_chains] (element-major, chain-minor) so that loads of one
 *     parameter across neighbouring chains coalesce.  Host-side get/set use
 *     the same [L][n_chains] layout.
 *   - there is NO CPU fallback: without a CUDA device every compute entry
 *     point fails with STK_ERR_CUDA.
 */
#ifndef STK_CUDA_H
#define STK_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct stk_ctx stk_ctx;
typedef struct stk_module stk_module;
typedef struct stk_chains stk_chains;
typedef void* stk_stream; /* cudaStream_t; NULL = the ctx's own stream */

#define STK_OK 0
#define STK_ERR_ARG (-1)     /* bad argument / unknown field / size mismatch  */
#define STK_ERR_CUDA (-2)    /* CUDA runtime / driver error, or no device     */
#define STK_ERR_COMPILE (-3) /* NVRTC compilation failed (log in last_error)  */
#define STK_ERR_IO (-4)      /* cache directory / file error                  */
#define STK_ERR_UNSUPPORTED (-5)

#define STK_ABI_VERSION 1

/* -------- diagnostics ---------------------------------------------------- */
const char* stk_last_error(void);
int stk_abi_version(void);
/* number of kernel launches issued by this library since load (bench.py's
 * "gpu_launches" claim is read from here, not guessed). */
int64_t stk_launch_count(void);

/* -------- context (one per GPU) ------------------------------------------ */
int stk_ctx_create(int device, stk_ctx** out);
int stk_ctx_destroy(stk_ctx* ctx);
/* the ctx's stream as a cudaStream_t (so a harness can record CUDA events on
 * the stream the kernels are launched on). */
int stk_ctx_stream(stk_ctx* ctx, stk_stream* out);

/* -------- modules: compiled models --------------------------------------- */
/* Compile generated CUDA source to an sm_100a cubin with NVRTC, without
 * touching a device (mirrors the "does it build" half of compileCC,
 * CC.hs:481-492).  The cubin is written to
 * <cache_dir>/stk_<sha1(source)>.cubin and its path copied to path_out.
 * A cache hit skips compilation (CC.hs:476-479).  include_dir is where
 * stk_device.cuh lives (NULL = the directory baked in at build time). */
int stk_compile(const char* cuda_src, const char* cache_dir,
                const char* include_dir, char* path_out, size_t path_cap);

/* The Language.CUDA lowering (stkc): Stochaskell IR text -> CUDA source of a module.
 * Replaces ccProgram / ccDAG / ccNode / ccPNode (CC.hs:123-432).  `ir_text` is what
 * `showPBlock` prints (Program.hs:181-187) — the text Language.CC embeds at the top of
 * every generated file (CC.hs:400-402) and runStep traces (Program.hs:1351).
 * `input_types` lists the kernel inputs' types in showType syntax (Expression.hs:437-453),
 * separated by `;` (empty = a model to forward-simulate: entry `simulate`, outputs out<k>);
 * a transition kernel  t -> P t  becomes kernel `step` over state fields in<k>...
 * (layout in csrc/stkc.cpp).  `options`: blank-separated key=value (name, veccap, block,
 * level).  Writes a NUL-terminated string; *needed (may be NULL) receives its size.
 * Unsupported IR constructs fail with STK_ERR_UNSUPPORTED naming the construct.  No GPU needed. */
int stk_ir_to_cuda(const char* ir_text, const char* input_types, const char* options,
                   char* out, size_t cap, size_t* needed);

/* Load a module on ctx's device.  `spec` is one of
 *   "builtin:<name>"   a hand-tuned model shipped with the library
 *   "<path>.cubin"     a cubin produced by stk_compile / nvcc
 *   anything else      CUDA source text: stk_compile + load
 */
int stk_module_load(stk_ctx* ctx, const char* spec, const char* cache_dir,
                    stk_module** out);
int stk_module_destroy(stk_module* mod);
/* the module's descriptor text (fields, kernels, launch geometry) */
int stk_module_describe(stk_module* mod, char* buf, size_t cap);
/* field / kernel / parameter lookup by name: index >= 0, or STK_ERR_ARG */
int stk_module_field(stk_module* mod, const char* name);
int stk_module_kernel(stk_module* mod, const char* name);

/* -------- chains: batched state ------------------------------------------ */
/* cudaMalloc the SoA state for n_chains chains.  `seed` keys the Philox
 * stream; draws of global chain c at transition t are a pure function of
 * (seed, c, t, slot), so results do not depend on how chains are sharded. */
int stk_chains_create(stk_module* mod, int64_t n_chains, uint64_t seed,
                      stk_chains** out);
int stk_chains_destroy(stk_chains* ch);
/* global index of this shard's first chain (multi-GPU: rank * n_chains) */
int stk_chains_set_offset(stk_chains* ch, int64_t first_global_chain);
/* capacity of a dynamically sized data field must be set before set_data */
int stk_set_data(stk_chains* ch, int field, const void* host, size_t bytes);
/* broadcast != 0: `host` holds ONE chain's value, replicated to all chains
 * (the hmcStanInit analogue, Stan.hs:498-500). */
int stk_set_state(stk_chains* ch, int field, const void* host, size_t bytes,
                  int broadcast);
int stk_get_state(stk_chains* ch, int field, void* host, size_t bytes);
/* raw device pointer of a field (zero-copy interop with torch tensors) */
int stk_device_ptr(stk_chains* ch, int field, void** dptr, size_t* bytes);
/* scalar tunables by name: "eps", "L", "rw_scale", ... (CudaMethod record,
 * mirroring StanMethod, Stan.hs:378-440) */
int stk_set_param(stk_chains* ch, const char* name, double value);
int stk_get_param(stk_chains* ch, const char* name, double* value);

/* -------- compute -------------------------------------------------------- */
/* joint log-density of each chain's current state (parity tier 1);
 * lp_host[n_chains].  Program.hs:787-812 (`lpdf`). */
int stk_logp(stk_chains* ch, double* lp_host);
/* log-density and gradient w.r.t. the unconstrained parameters;
 * grad_host is [P][n_chains]. */
int stk_logp_grad(stk_chains* ch, double* lp_host, double* grad_host);
/* n_transitions fused transitions of kernel `kernel_id` (MH / RJ / HMC) for
 * every chain, in ONE launch; asynchronous on `stream`. */
int stk_step(stk_chains* ch, int kernel_id, int64_t n_transitions,
             stk_stream stream);
/* forward simulation from the prior (simulate, Program.hs:1038) */
int stk_simulate(stk_chains* ch, stk_stream stream);
/* draws ring: keep every `thin`-th state of the next stk_step calls in a
 * device buffer [n_slots][P][n_chains]; n_slots = 0 disables. */
int stk_record(stk_chains* ch, int64_t n_slots, int64_t thin);
/* copy recorded draws [n_recorded][P][n_chains] to host; returns count in
 * *n_recorded (host may be NULL to query). */
int stk_get_draws(stk_chains* ch, double* host, size_t bytes,
                  int64_t* n_recorded);
/* per-parameter cross-chain diagnostics of the recorded draws on this GPU:
 * mean[P], var[P], rhat[P], ess[P] (any may be NULL).  `moments4` (may be
 * NULL) receives the additive [4][P] buffer (count, sum of chain means, sum
 * of squared chain means, sum of chain variances) that ranks all-reduce with
 * NCCL for the global R-hat. */
int stk_moments(stk_chains* ch, double* mean, double* var, double* rhat,
                double* ess, double* moments4);
/* the same diagnostics in two halves for multi-GPU runs: every rank fills the
 * additive buffer ([4][P] moments, then [P][max_lag+1] autocovariance sums;
 * bytes = 8*P*(4+max_lag+1); max_lag < 0 means n_recorded-1), ranks all-reduce
 * it (sum), and any rank finishes on the host. */
int stk_diag_partial(stk_chains* ch, int64_t max_lag, double* buf, size_t bytes);
int stk_diag_finish(const double* buf, int P, int64_t n_draws, int64_t max_lag,
                    double* mean, double* var, double* rhat, double* ess);
/* The same diagnostics over SEVERAL chain sets — the shards of one job, each on its own context / device (the
 * `devices` of CudaMethod; SURVEY 8e: chains are sharded contiguously, data replicated, this buffer is the one exchange).
 * One process driving several GPUs needs no collective library: every shard's additive buffer is computed on its device
 * and summed on the host.  With one process per GPU the ranks all-reduce stk_diag_partial's buffer instead (NCCL / gloo,
 * as bench.py does) and any rank calls stk_diag_finish.  All shards must have recorded the same number of draws of the
 * same module. */
int stk_moments_multi(stk_chains* const* shards, int n_shards, double* mean, double* var, double* rhat, double* ess);
/* One process per GPU (the reference-side driver launches one rank per device): the additive buffer is all-reduced by
 * NCCL over NVLink INSIDE the library (ncclAllReduce sum, fp64, on the context's stream; libnccl.so.2 is dlopen()ed on first
 * use, so the library has no link-time dependency on it).  Rank 0 calls stk_nccl_unique_id and sends the 128 bytes to the
 * other ranks by whatever means the host program has (MPI, a socket, torch.distributed); every rank then creates its
 * communicator and calls stk_moments_nccl collectively. */
typedef struct stk_comm stk_comm;
int stk_nccl_unique_id(void* id128);
int stk_nccl_comm_create(stk_ctx* ctx, int n_ranks, int rank, const void* id128, stk_comm** out);
int stk_nccl_comm_destroy(stk_comm* comm);
int stk_moments_nccl(stk_chains* ch, stk_comm* comm, double* mean, double* var, double* rhat, double* ess);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid allocation).  stk_diag_partial's additive buffer, computed by
 * variant kernels (csrc/stk_noalloc.cu) whose per-chain scratch (the [n_values][n_chains] chain means) is a static __device__
 * array of 512*1024 doubles (4 MiB, held by every GPU that loads the library) — the call allocates and frees nothing.
 *   d_draws  DEVICE pointer, [n_draws][n_values][n_chains] (the layout of stk_get_draws)
 *   d_buf    DEVICE pointer, bytes = 8*n_values*(4+max_lag+1) (max_lag < 0 or > n_draws-1 means n_draws-1), cleared here
 *   stream   cudaStream_t of the CURRENT device (NULL = its default stream); the memset and the two kernels are enqueued on
 *            it and the call returns without synchronising; read d_buf after the stream has reached this point and
 *            finish with stk_diag_finish.
 * Limits: n_values*n_chains <= 524288 (STK_ERR_UNSUPPORTED beyond it), n_draws <= 25600, n_values <= 65535.  The first call
 * on a device looks the array's address up (cudaGetSymbolAddress) — make it before starting a capture.  All calls on a device
 * share the one array, which keeps the last call's contents: calls that overlap on the GPU race, serialise them. */
int stk_diag_partial_noalloc(const double* d_draws, int64_t n_draws, int n_values, int64_t n_chains, int64_t max_lag,
                             double* d_buf, size_t bytes, stk_stream stream);
int stk_sync(stk_chains* ch);
/* doubles per recorded draw per chain (P_rec), valid after stk_record */
int stk_num_recorded_values(stk_chains* ch);

/* number of unconstrained parameters P and chains */
int stk_num_params(stk_chains* ch);
int64_t stk_num_chains(stk_chains* ch);

/* ---- wire formats either side of the HMC path (csrc/stk_io.cpp) ------------------------------------------
 * R dump: the text `stanDump` writes for data and initial values (Stan.hs:346-363) and CmdStan reads (`data file=`,
 * `init=`).  stk_rdump_parse replaces the reference's reliance on CmdStan's own reader: values come back column-major
 * as in the file, dims = {} for a scalar.  stk_rdump_format writes one `name <- value` line in stanDump's shape.
 * Stan CSV: what `runStan` reads back through `stanRead` (Stan.hs:365-371, 466-484): `#` comments, a header of column
 * names (`lp__`, `name.i.j`), one row per draw.  Errors: STK_ERR_ARG with stk_last_error(), e.g. "CSV file contains no
 * data" (the message of Stan.hs:371). */
typedef struct stk_rdump stk_rdump;
typedef struct stk_csv stk_csv;
int stk_rdump_parse(const char* text, stk_rdump** out);
int stk_rdump_count(const stk_rdump* d);
const char* stk_rdump_name(const stk_rdump* d, int i);
int stk_rdump_ndim(const stk_rdump* d, int i);
const long long* stk_rdump_dims(const stk_rdump* d, int i);
int stk_rdump_is_int(const stk_rdump* d, int i);
long long stk_rdump_size(const stk_rdump* d, int i);
const double* stk_rdump_values(const stk_rdump* d, int i);
void stk_rdump_free(stk_rdump* d);
int stk_rdump_format(const char* name, const double* values, int is_int, int ndim, const long long* dims,
                     char* buf, size_t cap, size_t* need);
int stk_stan_csv_write(const char* path, int n_cols, const char* const* col_names, const double* draws,
                       long long n_draws, long long stride_draw, long long stride_col, const char* comment);
int stk_stan_csv_read(const char* path, stk_csv** out);
int stk_csv_cols(const stk_csv* c);
long long stk_csv_rows(const stk_csv* c);
const char* stk_csv_col_name(const stk_csv* c, int i);
const double* stk_csv_values(const stk_csv* c);      /* row-major [rows][cols] */
const char* stk_csv_comments(const stk_csv* c);
void stk_csv_free(stk_csv* c);
/* writeNPy (Data/Expression/Const/IO.hs:28-45): NumPy 1.0 file, '<f8', C order, the reference's header padding byte for
 * byte; `values` row-major, dims = {} for a scalar.  readRealMatrix (IO.hs:22-26): headerless CSV of doubles into an
 * stk_csv (column names empty); ragged rows or non-numeric fields are STK_ERR_ARG. */
int stk_npy_write(const char* path, const double* values, int ndim, const long long* dims);
int stk_real_matrix_read(const char* path, stk_csv** out);

#ifdef __cplusplus
}
#endif
#endif /* STK_CUDA_H */
