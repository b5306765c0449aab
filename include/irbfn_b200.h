/*
 * irbfn_b200.h -- C ABI of the B200-native IRBFN evaluation library (libirbfn_b200.so).
 *
 * This is the drop-in boundary for the one hot path of hzheng40/irbfn: the batched
 * evaluation (forward + VJP) of the interpolating RBF network.  Every entry point
 * takes plain device pointers, sizes and a CUDA stream; no torch / XLA / Python types
 * appear.  An XLA FFI handler, a ctypes binding or a torch extension binds exactly these
 * symbols (INTEGRATION.md shows each).
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *   irbfn_forward          <- WCRBFNet.__call__            src/irbfn_mpc/model.py:169-198
 *                             (gate  _region_activation     src/irbfn_mpc/model.py:42-95,
 *                              layer RBFLayer.__call__      flax_rbf/flax_rbf.py:258-285,
 *                              basis gaussian / inverse_quadratic  flax_rbf.py:34-37,49-52,
 *                              blend model.py:187-193, readout nn.Dense model.py:196)
 *   irbfn_forward (hidden) <- DeeperWCRBFNet front end      src/irbfn_mpc/model.py:259-283
 *   irbfn_forward (gate)   <- ClusterWCRBFNet blend+readout src/irbfn_mpc/model.py:395-414
 *   irbfn_softmax_gate     <- softmax(Dense(x))             src/irbfn_mpc/model.py:402-406
 *   irbfn_region_gate      <- _region_activation alone      src/irbfn_mpc/model.py:42-95
 *   irbfn_backward         <- jax.value_and_grad(loss_fn) through the module,
 *                             scripts/train_nmpc.py:297,416; scripts/train_nmpc_frenet.py:388,416,448
 *                             (cotangent dy (B,O) in; dcenters, dlog_sigs, dkernel, dbias,
 *                              dx, dgate out)
 *
 * Conventions
 *   - All arrays are dense, row-major, fp32 unless stated.  Symbols: B queries, R regions
 *     (local networks), K kernels per region, D input features, O outputs, S split dims.
 *   - The library allocates nothing and keeps no global mutable state.  Scratch memory is
 *     caller-provided (irbfn_*_workspace_bytes tells how much).  Outputs are fully
 *     overwritten, never accumulated into.
 *   - All work is enqueued on the given stream; no device-wide synchronisation; every call is
 *     CUDA-graph capturable.
 *   - Return value: 0 on success, a negative IRBFN_ERR_* code otherwise.  Nothing is printed
 *     and nothing aborts.  irbfn_error_string() maps a code to text.
 *   - x == centre is handled on d^2, so gradients there are the finite analytic limit
 *     (the reference's sqrt formulation yields NaN).
 */
#ifndef IRBFN_B200_H_
#define IRBFN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IRBFN_ABI_VERSION 1

/* basis functions (flax_rbf.py:34-37, 49-52) */
#define IRBFN_BASIS_GAUSSIAN 0          /* phi = exp(-alpha^2)     */
#define IRBFN_BASIS_INVERSE_QUADRATIC 1 /* phi = 1 / (1 + alpha^2) */

/* compiled limits */
#define IRBFN_MAX_D 8   /* input features               */
#define IRBFN_MAX_S 8   /* split dimensions of the gate */

/* error codes */
#define IRBFN_OK 0
#define IRBFN_ERR_BAD_DIMS (-1)       /* a dimension is <= 0 or inconsistent            */
#define IRBFN_ERR_UNSUPPORTED_D (-2)  /* D > IRBFN_MAX_D                                */
#define IRBFN_ERR_BAD_BASIS (-3)      /* basis id is not one of IRBFN_BASIS_*           */
#define IRBFN_ERR_NULL_POINTER (-4)   /* a required pointer is NULL                     */
#define IRBFN_ERR_BAD_GATE (-5)       /* neither / both gate inputs, or S > D, S > MAX  */
#define IRBFN_ERR_WORKSPACE (-6)      /* workspace missing or too small                 */
#define IRBFN_ERR_LAUNCH (-7)         /* CUDA reported a launch / runtime error         */
#define IRBFN_ERR_NO_DEVICE (-8)      /* no usable sm_100 device is current             */
#define IRBFN_ERR_BAD_MASK (-9)       /* grad mask empty or asks for an output whose pointer is NULL */

/* which gradients irbfn_backward produces */
#define IRBFN_GRAD_CENTERS 0x01u
#define IRBFN_GRAD_LOG_SIGS 0x02u
#define IRBFN_GRAD_KERNEL 0x04u /* also produces dbias */
#define IRBFN_GRAD_X 0x08u
#define IRBFN_GRAD_GATE 0x10u   /* dgate (B,R); external-gate mode only */
#define IRBFN_GRAD_PARAMS (IRBFN_GRAD_CENTERS | IRBFN_GRAD_LOG_SIGS | IRBFN_GRAD_KERNEL)

typedef void* irbfn_stream_t; /* a cudaStream_t */

/*
 * Static description of the tanh region gate (module attributes lower_bounds, upper_bounds,
 * delta, dimension_ranges, activation_idx of model.py:113-125), packed flat.  The struct lives
 * on the host; the five arrays live on the device.
 *   n_d        = offset[d+1] - offset[d] bounds on split dimension d (d < num_split)
 *   g_d[i](x)  = ((tanh(delta[d]*(x_d - lower[offset[d]+i])) + 1)/2)
 *              * ((tanh(delta[d]*(upper[offset[d]+i] - x_d)) + 1)/2)
 *   gamma_r(x) = prod_d g_d[ranges[r*num_split + d]](x)   for r < num_ranges, else 0
 * Not normalised (model.py:88-93).
 */
typedef struct IrbfnGateDesc {
    int32_t num_split;      /* S = len(activation_idx); split dims are input columns 0..S-1 */
    int32_t num_ranges;     /* rows of dimension_ranges (normally == R)                      */
    int32_t num_bounds;     /* sum_d n_d                                                     */
    const float* lower;     /* device, (num_bounds)                                          */
    const float* upper;     /* device, (num_bounds)                                          */
    const int32_t* offset;  /* device, (S + 1)                                               */
    const float* delta;     /* device, (S)                                                   */
    const int32_t* ranges;  /* device, (num_ranges, S)                                       */
} IrbfnGateDesc;

int irbfn_abi_version(void);
const char* irbfn_error_string(int code);
/* After IRBFN_ERR_LAUNCH: which step failed and the CUDA error text.  Per host thread. */
const char* irbfn_last_error_detail(void);

/*
 * Forward:  y[b,:] = (sum_r gamma[b,r] * phi(x[b], centers[r], log_sigs[r])) @ W + bias
 *
 *   x         (B,D)    queries
 *   centers   (R,K,D)  params['rbf_list']['centers']
 *   log_sigs  (R,K)    params['rbf_list']['log_sigs']   (one scalar width per kernel)
 *   W, bias   (K,O),(O) params['linear']['kernel'|'bias'];  both may be NULL iff y is NULL
 *   gate      (B,R)    externally computed gate (ClusterWCRBFNet) or NULL
 *   gd        tanh region gate description or NULL.  Exactly one of gate / gd is given.
 *   y         (B,O)    output, or NULL
 *   hidden    (B,K)    blended activations h = sum_r gamma_r phi_r (DeeperWCRBFNet front
 *                      end, model.py:283), or NULL.  At least one of y / hidden is given.
 *   workspace / workspace_bytes : >= irbfn_forward_workspace_bytes(...) (may be 0 / NULL).
 */
size_t irbfn_forward_workspace_bytes(int64_t B, int R, int K, int D, int O, int fused_gate);

int irbfn_forward(const float* x, const float* centers, const float* log_sigs, const float* W,
                  const float* bias, const float* gate, const IrbfnGateDesc* gd, float* y,
                  float* hidden, int64_t B, int R, int K, int D, int O, int basis,
                  void* workspace, size_t workspace_bytes, irbfn_stream_t stream);

/*
 * Which kernel irbfn_forward runs for a shape (a host-side query; nothing is launched).  y_only: the call asks for
 * y alone (no hidden activations, not irbfn_forward_gate / _streamed).
 *   IRBFN_PATH_TILED  the query-stationary throughput kernel, one CTA per 1,024-query tile
 *   IRBFN_PATH_SPLIT  the same with every region's kernels split over CTAs and a fixed-order reduction of the
 *                     partial outputs (needs the workspace; without it the tiled kernel runs)
 *   IRBFN_PATH_SMALL  the latency kernel of pred_step (src/irbfn_mpc/irbfn_planner.py:29-32: one query per control
 *                     step; 400 in the notebooks): a CTA owns one or four queries and its threads share the
 *                     (region, kernel) pairs; ONE launch, no workspace.  Taken when y_only, O <= 16,
 *                     R*K <= 16,384 and B*R*K <= 5*2^20.
 * All three give the same y within the forward's tolerance (the summation order over kernels differs); each is
 * deterministic run to run.  Negative: an IRBFN_ERR_* code.
 */
#define IRBFN_PATH_TILED 0
#define IRBFN_PATH_SPLIT 1
#define IRBFN_PATH_SMALL 2
int irbfn_forward_path(int64_t B, int R, int K, int D, int O, int y_only);

/*
 * Training-step forward: irbfn_forward with the tanh gate, which also writes the gate gamma (B,R) it
 * computes on the fly (model.py:70-95).  Handing that buffer to irbfn_backward as its external `gate`
 * (gd = NULL) saves the backward's own gate pass -- what jax.value_and_grad's forward residuals would
 * hold for scripts/train_nmpc.py:297.  gate_out must not be NULL.
 */
int irbfn_forward_gate(const float* x, const float* centers, const float* log_sigs, const float* W,
                       const float* bias, const IrbfnGateDesc* gd, float* y, float* gate_out,
                       int64_t B, int R, int K, int D, int O, int basis,
                       void* workspace, size_t workspace_bytes, irbfn_stream_t stream);

/*
 * Forward over rows that are still being uploaded (the host-buffer path, bench.py's e2e): irbfn_forward launched
 * BEFORE its input has arrived.  The caller uploads x in pieces of queries_per_flag rows on other streams and, behind
 * each piece on the piece's stream, stores a non-zero word into the piece's flag (cuStreamWriteValue32 or a 4-byte
 * copy; irbfn_upload_streamed does both steps).  The persistent kernel hands its 1,024-query tiles out through a
 * counter and waits for the flag of a tile's piece before it touches the tile's rows, so the upload of later pieces
 * overlaps the evaluation of earlier ones inside ONE launch -- no per-chunk launches, ramps or tails.  y may be device
 * memory or pinned, device-mapped host memory (the kernel's stores then go straight over PCIe, as full 16-byte
 * vectors: no download pass).  Replaces the reference's implicit jax.device_put of the whole batch before wcrbf.apply
 * (irbfn_vs_explicit_frenet.ipynb, cells 8-13).
 *   ready    device uint32: ceil(B / queries_per_flag) flags followed by ONE tile-counter word; ALL ZERO when this call
 *            is enqueued (and not written by a copy stream before that)
 *   queries_per_flag   a multiple of 1,024
 *   status   optional device int, ZERO on entry; set to 1 if a piece did not arrive within ~2 s (e.g. under a
 *            profiler that serialises the kernel against the copy streams).  The whole launch is then given up -- every
 *            CTA stops at its next wait -- and y is undefined.  The word is sticky: a launch that finds it non-zero
 *            gives up at its first wait, until the caller clears it.
 *   O <= 16.  An external gate must be complete in device memory before the call (only x is streamed).
 */
int irbfn_forward_streamed(const float* x, const float* centers, const float* log_sigs, const float* W,
                           const float* bias, const float* gate, const IrbfnGateDesc* gd, float* y, int64_t B,
                           int R, int K, int D, int O, int basis, uint32_t* ready, int64_t queries_per_flag,
                           int* status, irbfn_stream_t stream);

/*
 * The upload side of irbfn_forward_streamed, for a caller that has no copy loop of its own: enqueues the
 * host->device copies of x (B,D) from PINNED host memory in pieces of queries_per_flag rows, alternating over
 * num_streams copy streams (one stream's per-copy set-up hides behind another's transfer), each piece followed on its
 * stream by a cuStreamWriteValue32 that sets the piece's flag.  None of the copy streams may be the stream
 * irbfn_forward_streamed runs on, and all must be ordered behind the clearing of ready[] (an event).
 */
int irbfn_upload_streamed(const float* x_host, float* x_dev, int64_t B, int D, uint32_t* ready,
                          int64_t queries_per_flag, const irbfn_stream_t* copy_streams, int num_streams);
/* 1 when the driver offers cuStreamWriteValue32 (irbfn_upload_streamed can run), else 0: ask BEFORE launching
 * irbfn_forward_streamed, whose tiles would otherwise wait for flags that never come. */
int irbfn_upload_streamed_available(void);

/*
 * irbfn_forward for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 *
 * Same arguments and result as irbfn_forward without the workspace pair.  The split-over-kernels partial outputs of
 * small batches (what irbfn_forward_workspace_bytes sizes) go to a file-scope `static __device__` array of
 * IRBFN_NOALLOC_SCRATCH_BYTES (8 MiB) inside the library; every GPU that loads the library holds those 8 MiB until the
 * process exits.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine), on any call.
 *   - IRBFN_ERR_WORKSPACE (irbfn_last_error_detail() names both sizes) when irbfn_forward_workspace_bytes(B, ...)
 *     exceeds IRBFN_NOALLOC_SCRATCH_BYTES: use irbfn_forward with a caller-owned workspace, or a smaller batch.
 *     (Batches of >= 4 tiles of 1,024 queries per SM need no scratch at all and always pass.)
 *   - every call on a device shares the one array and it keeps the last call's contents.  Nothing in it has to start
 *     at zero: each split writes its whole partial output before the fixed-order reduction reads it.  Calls that
 *     OVERLAP on the GPU (different streams, or replays of graphs that captured it) RACE on the array: serialise them.
 */
#define IRBFN_NOALLOC_SCRATCH_BYTES (8u << 20)
int irbfn_forward_noalloc(const float* x, const float* centers, const float* log_sigs, const float* W,
                          const float* bias, const float* gate, const IrbfnGateDesc* gd, float* y,
                          float* hidden, int64_t B, int R, int K, int D, int O, int basis,
                          irbfn_stream_t stream);

/*
 * Gate-sparse, region-binned forward (tanh gate only).  Same result as irbfn_forward within
 * an ABSOLUTE tolerance the caller chooses: a local network r is skipped for a query when
 *     gamma_r <= tol_abs / (R * max_o sum_k |W[k,o]|),
 * which bounds the dropped terms of model.py:187-196 by tol_abs per output (phi <= 1 for both bases), so
 *     |y[b,o] - y_dense[b,o]| <= tol_abs  (+ fp32 rounding)   for every b, o.
 * tol_abs = 0 skips only regions whose gate underflowed to exactly 0.  Queries are binned by their set of
 * non-negligible gate intervals (a counting sort in the workspace) so that a warp's 128 queries skip the same
 * regions; shapes whose parameters are staged in chunks stage only the regions a CTA needs.
 *   counters   optional device uint64[2]: (warp, region-chunk, region) evaluations skipped / done
 *   workspace  >= irbfn_forward_sparse_workspace_bytes(B, gd->num_bounds)
 * B < 2^31.  The dense entry point stays the reference for timing and roofline accounting.
 */
size_t irbfn_forward_sparse_workspace_bytes(int64_t B, int num_bounds);
int irbfn_forward_sparse(const float* x, const float* centers, const float* log_sigs, const float* W,
                         const float* bias, const IrbfnGateDesc* gd, float* y, int64_t B, int R, int K,
                         int D, int O, int basis, float tol_abs, uint64_t* counters, void* workspace,
                         size_t workspace_bytes, irbfn_stream_t stream);
/* The same with the GATE threshold given directly (a local network is skipped for a query when gamma_r <= gate_tol):
 * the forward of a gate-sparse training step, which hands the same threshold to irbfn_backward_sparse. */
int irbfn_forward_sparse_gate(const float* x, const float* centers, const float* log_sigs, const float* W,
                              const float* bias, const IrbfnGateDesc* gd, float* y, int64_t B, int R, int K,
                              int D, int O, int basis, float gate_tol, uint64_t* counters, void* workspace,
                              size_t workspace_bytes, irbfn_stream_t stream);

/*
 * Backward (VJP).  Recomputes phi from x and the parameters; nothing from the forward is saved.
 *
 *   dy        (B,O)    cotangent of y, or NULL (then W may be NULL too)
 *   dhidden   (B,K)    cotangent of the hidden activations (DeeperWCRBFNet front end), or NULL.
 *                      At least one of dy / dhidden is given; with both, the cotangent reaching
 *                      h is dy W^T + dhidden.
 *   dcenters  (R,K,D)  if grad_mask & IRBFN_GRAD_CENTERS
 *   dlog_sigs (R,K)    if grad_mask & IRBFN_GRAD_LOG_SIGS
 *   dW, dbias (K,O),(O) if grad_mask & IRBFN_GRAD_KERNEL
 *   dx        (B,D)    if grad_mask & IRBFN_GRAD_X.  With a tanh gate (gd) this includes the
 *                      gate's own dependence on x; with an external gate it is the RBF-layer
 *                      term only and dgate carries the rest.
 *   dgate     (B,R)    if grad_mask & IRBFN_GRAD_GATE (external gate only)
 * Parameter gradients are reduced over the batch deterministically (fixed partition, fixed
 * order, no atomics).
 */
size_t irbfn_backward_workspace_bytes(int64_t B, int R, int K, int D, int O, unsigned grad_mask);

int irbfn_backward(const float* x, const float* centers, const float* log_sigs, const float* W,
                   const float* gate, const IrbfnGateDesc* gd, const float* dy,
                   const float* dhidden, float* dcenters, float* dlog_sigs, float* dW, float* dbias,
                   float* dx, float* dgate, int64_t B, int R, int K, int D, int O, int basis,
                   unsigned grad_mask, void* workspace, size_t workspace_bytes,
                   irbfn_stream_t stream);

/*
 * Gate-sparse parameter gradients: irbfn_backward with one more argument.  The PARAMETER gradients
 * (dcenters, dlog_sigs, dW) leave out the terms of every (query b, region r) pair whose gate gamma[b,r] <= gate_tol;
 * each such term of jax.value_and_grad(loss_fn) (scripts/train_nmpc.py:297) carries the factor gamma[b,r], so
 *     |d theta - d theta(dense)| <= gate_tol * sum over dropped (b,r) of |d(phi_rk dh_bk)/d theta|,
 * and gate_tol = 0 drops only gates that are exactly 0 (bit-identical sums).  dbias and dgate are not affected.
 * dx (tanh gate): the queries are binned by their set of non-negligible gate intervals (as irbfn_forward_sparse does)
 * and a warp skips the regions whose gate is <= gate_tol for all of its 128 queries; every dropped term of dx, the
 * gate path included, carries the factor gamma, so dx moves by the order of gate_tol too.
 * The tanh gate of the shipped multi-region shapes is > 2^-24 for 4-5 of 12 regions on uniformly drawn queries.
 * The kernel's threads own (region, kernel) pairs and a query is a broadcast operand, so the skip is one
 * warp-uniform vote per (warp, query): no binning pass, same fixed summation order (deterministic).  Shapes without a
 * gate-sparse instantiation (one region, wide outputs, a dhidden cotangent) run the dense kernels: same result.
 * Same workspace as irbfn_backward.  gate_tol < 0 or NaN: IRBFN_ERR_BAD_DIMS.
 */
int irbfn_backward_sparse(const float* x, const float* centers, const float* log_sigs, const float* W,
                          const float* gate, const IrbfnGateDesc* gd, const float* dy,
                          const float* dhidden, float* dcenters, float* dlog_sigs, float* dW, float* dbias,
                          float* dx, float* dgate, int64_t B, int R, int K, int D, int O, int basis,
                          unsigned grad_mask, float gate_tol, void* workspace, size_t workspace_bytes,
                          irbfn_stream_t stream);

/*
 * Introspection (host only, no device needed): how irbfn_forward cuts the (R,K) parameter set into
 * shared-memory stages for a device with smem_limit_bytes of opt-in shared memory per block
 * (B200: 232448).  regions_per_stage == R and kernels_per_stage >= K means the whole parameter set
 * is staged once per CTA.  Returns IRBFN_ERR_BAD_DIMS if no plan exists.
 */
int irbfn_forward_plan(int R, int K, int D, int O, int smem_limit_bytes, int* regions_per_stage,
                       int* kernels_per_stage, int* threads_per_block, size_t* smem_bytes);

/* gamma (B,R) of the tanh region gate alone (model.py:42-95). */
int irbfn_region_gate(const float* x, const IrbfnGateDesc* gd, float* gamma, int64_t B, int R,
                      int D, irbfn_stream_t stream);

/* ClusterWCRBFNet gate (model.py:402-406): logits = x @ Wc + bc (B,R); gate = softmax(logits).
 * logits may be NULL. */
int irbfn_softmax_gate(const float* x, const float* Wc, const float* bc, float* gate,
                       float* logits, int64_t B, int D, int R, irbfn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Loss-side vehicle dynamics: the tail of the reference's loss_fn after
 * wcrbf.apply, fused into one launch that yields the loss terms AND the cotangent dloss/dy_pred
 * which irbfn_backward consumes.
 *
 *   mode IRBFN_LOSS_FRENET_ONEINT  <- train_step_oneint   scripts/train_nmpc_frenet.py:339-392
 *                                     (|y_pred - y|.mean() + 100 |aux(x,y_pred) - aux(x,y)|.mean(),
 *                                      aux = dynamic_frenet_onestep_aux, dynamics.py:293-383)
 *   mode IRBFN_LOSS_FRENET_FULLINT <- train_step_fullint  scripts/train_nmpc_frenet.py:395-420
 *                                     (integrate_frenet_mult, dynamics.py:284-290; O == 10)
 *   mode IRBFN_LOSS_ST_ONEINT      <- train_step_oneint   scripts/train_nmpc.py:258-300
 *                                     (optax.l2_loss terms; dynamic_st_onestep_aux, dynamics.py:103-187)
 *   mode IRBFN_LOSS_ST_FULLINT     <- train_step_fullint  scripts/train_nmpc.py:303-421 (O == 10)
 *
 *   x (B,D), y_pred (B,O), y (B,O) device; dyn_params: HOST array of 13 floats in the order of
 *   scripts/train_nmpc_frenet.py:322-338 (mu, m, I, lf, lr, C_Sf, C_Sr, h, dt, sv_max, a_max,
 *   s_max, v_max).  O is 2 or 10.  Means are taken over global_batch (>= B) samples so that shard
 *   cotangents / losses summed over ranks equal the one-device values.
 *   dy (B,O) device out; losses device out, 2 floats: {pred_loss, int_loss (weighted)} of this shard.
 *   Reduction is deterministic (fixed partition, fixed order, fp64).
 */
#define IRBFN_LOSS_FRENET_ONEINT 0
#define IRBFN_LOSS_FRENET_FULLINT 1
#define IRBFN_LOSS_ST_ONEINT 2
#define IRBFN_LOSS_ST_FULLINT 3

size_t irbfn_dynamics_loss_workspace_bytes(int64_t B);

int irbfn_dynamics_loss(int mode, const float* x, const float* y_pred, const float* y,
                        const float* dyn_params, float* dy, float* losses, int64_t B, int D, int O,
                        int64_t global_batch, void* workspace, size_t workspace_bytes,
                        irbfn_stream_t stream);

/*
 * The prediction term of the loss alone, any output width, with its cotangent, in one launch plus the
 * deterministic reduction: replaces `optax.l2_loss(pred_u, u).mean()` (scripts/train_nmpc.py:287;
 * IRBFN_PRED_LOSS_L2 = mean of 0.5 d^2) / `jnp.abs(pred_u - u).mean()` (scripts/train_nmpc_frenet.py:347)
 * and their autodiff transposes.  y_pred, y, dy are (B,O) device arrays; loss is ONE device float, this
 * shard's sum divided by global_batch * O.  Workspace: irbfn_prediction_loss_workspace_bytes(B, O).
 */
#define IRBFN_PRED_LOSS_L2 0
#define IRBFN_PRED_LOSS_L1 1

size_t irbfn_prediction_loss_workspace_bytes(int64_t B, int O);

int irbfn_prediction_loss(int kind, const float* y_pred, const float* y, float* dy, float* loss,
                          int64_t B, int O, int64_t global_batch, void* workspace, size_t workspace_bytes,
                          irbfn_stream_t stream);

/*
 * Optimiser update on a flat fp32 parameter buffer: optax.chain(clip_by_global_norm(max_norm), adam(lr, b1,
 * b2, eps)) as built at scripts/train_nmpc.py:231-236 and applied by state.apply_gradients (:299).  optax is a
 * dependency of the reference (pyproject.toml, unpinned), not part of it; the update follows optax's published
 * clip_by_global_norm / scale_by_adam / scale(-lr) definitions.
 *   params, grad, mu, nu: device arrays of n floats, updated in place (grad is left clipped).
 *   step_count: ONE device double, the Adam count; incremented by this call before use (so the call is
 *   CUDA-graph capturable).  max_norm <= 0 disables clipping.  grad_norm: device float out or NULL
 *   (the pre-clip global norm).  The norm is reduced in a fixed order in fp64: deterministic.
 *   Hyper-parameters are doubles and are rounded to float32 one by one, (1 - b) included, the way JAX
 *   rounds weak-typed Python scalars.
 */
size_t irbfn_clip_adam_workspace_bytes(int64_t n);

int irbfn_clip_adam(float* params, float* grad, float* mu, float* nu, int64_t n, double* step_count, double lr,
                    double b1, double b2, double eps, double max_norm, float* grad_norm, void* workspace,
                    size_t workspace_bytes, irbfn_stream_t stream);

/*
 * The tail of a data-parallel training step in ONE launch: [one-shot all-reduce of the flat gradient over
 * NVLink peer memory] -> global norm -> clip_by_global_norm -> Adam (scripts/train_nmpc.py:231-236,299), for
 * gradients of at most irbfn_fused_update_max_n() floats (the shipped NMPC shapes hold 10-20 k): at that size
 * ncclAllReduce + the two optimiser launches are pure latency.
 *
 *   peers    device array of `world` pointers: rank r's symmetric buffer mapped into this process
 *            (peers[rank] is this rank's own), each irbfn_fused_update_buffer_bytes(n + n_extra, world) bytes,
 *            zero-filled once before the first call.  This rank's backward writes its LOCAL gradient into
 *            peers[rank][0 .. n) and up to n_extra more floats (the loss) behind it; the words after the data
 *            are flags owned by the kernel.  NULL with world == 1: `reduced` then already holds the gradient.
 *   reduced  [n + n_extra] local: receives the sum over ranks (same summation order on every rank, so the
 *            replicated parameters stay bit-identical); its first n floats are clipped in place by the update
 *   apply_update  0: all-reduce only
 *   status   optional device int, set to 1 if a peer did not arrive within ~20 s
 * Every rank must make the same sequence of calls.  Graph-capturable (the step sequence number lives in the
 * buffer).  IRBFN_ERR_BAD_DIMS when n + n_extra exceeds the limit: use NCCL + irbfn_clip_adam.
 */
int64_t irbfn_fused_update_max_n(void);
size_t irbfn_fused_update_buffer_bytes(int64_t n_total, int world);
int irbfn_fused_update(float* params, float* mu, float* nu, float* reduced, int64_t n, int n_extra,
                       float* const* peers, int rank, int world, int apply_update, double* step_count,
                       double lr, double b1, double b2, double eps, double max_norm, float* grad_norm,
                       int* status, irbfn_stream_t stream);

/*
 * Plain rollouts (what irbfn_planner.py:212-214,510-518 and the losses call):
 *   IRBFN_DYN_FRENET_MULT         integrate_frenet_mult       dynamics.py:284-290  (B,18) -> (B,5,8)
 *   IRBFN_DYN_FRENET_ONESTEP_AUX  dynamic_frenet_onestep_aux  dynamics.py:293-383  (B,width>=8) -> (B,6)
 *   IRBFN_DYN_ST_MULT             integrate_st_mult           dynamics.py:94-100   (B,17) -> (B,5,7)
 *   IRBFN_DYN_ST_ONESTEP_AUX      dynamic_st_onestep_aux      dynamics.py:103-187  (B,width>=9) -> (B,7)
 * x_and_u (B,width) and states are device arrays; dyn_params is a HOST array of 13 floats.
 */
#define IRBFN_DYN_FRENET_MULT 0
#define IRBFN_DYN_FRENET_ONESTEP_AUX 1
#define IRBFN_DYN_ST_MULT 2
#define IRBFN_DYN_ST_ONESTEP_AUX 3

int irbfn_integrate(int model, const float* x_and_u, const float* dyn_params, float* states,
                    int64_t B, int width, irbfn_stream_t stream);

/*
 * Per-query Jacobian in one pass:  jac[b,o,j] = d y[b,o] / d x[b,j].
 * Replaces jax.jacrev(lambda x: model.apply(params, x)) over WCRBFNet.__call__
 * (src/irbfn_mpc/model.py:169-198); the reference has no call site.
 *   jac      (B,O,D)  out
 *   y        (B,O)    out or NULL (the forward value rides along for free)
 *   region_y (B,R,O)  out or NULL: sum_k W[k,o] phi_rk per region.  With an external gate the
 *                     Jacobian holds the RBF-layer term only (gate treated as constant); the caller
 *                     adds sum_r region_y[b,r,o] * dgate[b,r]/dx_j.  With the tanh gate (gd) the
 *                     gate's own dependence on x is included.
 * Other arguments as irbfn_forward.  No workspace.
 */
int irbfn_jacobian(const float* x, const float* centers, const float* log_sigs, const float* W,
                   const float* bias, const float* gate, const IrbfnGateDesc* gd, float* jac,
                   float* y, float* region_y, int64_t B, int R, int K, int D, int O, int basis,
                   irbfn_stream_t stream);

/*
 * Gate-sparse one-pass Jacobian (tanh gate): irbfn_jacobian with the terms of every (query, region) pair left out
 * for which one of the region's interval factors is <= gate_tol (then gamma <= gate_tol, and every term of J and y the
 * pair contributes carries the factor gamma).  Queries are binned by their set of non-negligible gate intervals
 * (irbfn_forward_sparse's passes) and a warp skips the regions all of its 64 queries drop -- gate and kernels alike.
 * No reference call site, like irbfn_jacobian.  Shapes the sparse kernel does not cover (one region,
 * more than 64 gate intervals, parameters larger than shared memory) run the dense kernel: same result.
 *   workspace >= irbfn_jacobian_sparse_workspace_bytes(B, gd->num_bounds)
 */
size_t irbfn_jacobian_sparse_workspace_bytes(int64_t B, int num_bounds);
int irbfn_jacobian_sparse(const float* x, const float* centers, const float* log_sigs, const float* W,
                          const float* bias, const IrbfnGateDesc* gd, float* jac, float* y, int64_t B,
                          int R, int K, int D, int O, int basis, float gate_tol, void* workspace,
                          size_t workspace_bytes, irbfn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Table-side kernels.
 *
 * irbfn_gather_rows: the training input pipeline on the device.  x_out[i,:] = X[index[i],:] and
 * y_out[i,:] = Y[index[i],:] in one launch -- replaces the host-side NumPy gather + per-step
 * host->device copy of scripts/train_nmpc_frenet.py:499-501 (batch_x = X[perm]).  X (N,D), Y (N,O)
 * (Y / y_out may be NULL with O == 0), index (n) int64, all on the device.  An index outside [0,N)
 * is never dereferenced: its output row is zero and *bad_index_flag (device int, caller-zeroed) is
 * set to 1.  Bit-exact.
 */
int irbfn_gather_rows(const float* X, const float* Y, const int64_t* index, float* x_out, float* y_out,
                      int64_t n, int64_t N, int D, int O, int* bad_index_flag, irbfn_stream_t stream);

/*
 * irbfn_lattice_queries: the sweep's query lattice built on the device -- np.meshgrid(*axes, indexing="ij")
 * over the listed columns, every other column held at fixed[col] (irbfn_vs_explicit_frenet.ipynb cell 8).
 * values: the axes' values concatenated; offset (num_axes + 1) int32 prefix sums; column (num_axes) int32 query
 * column of each axis (the first axis varies slowest); fixed (D); x (n, D) with n = product of the axis lengths.
 * All device pointers.  Bit-exact with the NumPy spelling on the same float32 axis values.
 */
int irbfn_lattice_queries(const float* values, const int32_t* offset, const int32_t* column, int num_axes,
                          const float* fixed, float* x, int64_t n, int D, irbfn_stream_t stream);

/*
 * irbfn_grid_index: per dimension d, index[b,d] = min(len_d - 1, searchsorted(keys_d, q[b,d],
 * side="right")) -- the grid lookup of ExplicitPlanner.plan (src/irbfn_mpc/explicit_planner.py:165-175).
 * keys: all dimensions' sorted key arrays concatenated; offset (num_dims + 1) int32 prefix sums;
 * q (B,num_dims); index (B,num_dims) int32.  Index-exact with numpy.searchsorted on the SAME float32 keys and
 * queries (NaN sorts last).  The reference bisects float64 table keys with float64 values: a value within
 * float32 rounding of a key can land in the neighbouring cell once both are narrowed to float32 -- callers
 * that need the reference's float64 cell boundaries use irbfn_grid_index_f64 below.
 * irbfn_nearest_row compares float32 SQUARED distances with a strict <, rows in increasing order (lowest
 * index wins exact ties of the squared distance); the reference takes argmin of float32 norms, where the
 * square root can merge two distinct squared distances into one value and hand the tie to the lower index.
 */
int irbfn_grid_index(const float* q, const float* keys, const int32_t* offset, int32_t* index,
                     int64_t B, int num_dims, irbfn_stream_t stream);

/*
 * irbfn_grid_index_f64: the same lookup on float64 keys and float64 values -- what explicit_planner.py:165-175
 * bisects (the table's keys are float64 arrays, the query a float64 NumPy row), so the cell is the reference's for
 * every input, including values within float32 rounding of a key.  HBM-bound either way.
 */
int irbfn_grid_index_f64(const double* q, const double* keys, const int32_t* offset, int32_t* index,
                         int64_t B, int num_dims, irbfn_stream_t stream);

/*
 * irbfn_nearest_row: index[b] = argmin_n || table[n,:] - q[b,:] ||_2 (first minimum wins), the
 * brute-force lookup get_closest_ind (src/irbfn_mpc/explicit_planner.py:21-25) and
 * irbfn_vs_explicit_frenet.ipynb cell 11, batched.  q (B,D), table (N,D), index (B) int64,
 * distance (B) Euclidean distance of the winner or NULL.  Distances are accumulated in fp32
 * (sequential FMA over the features); rows whose fp32 distances tie resolve to the lowest index.
 * Small batches split the table across CTAs; the partial winners live in the workspace.
 */
size_t irbfn_nearest_row_workspace_bytes(int64_t B, int64_t N);

int irbfn_nearest_row(const float* q, const float* table, int64_t* index, float* distance, int64_t B,
                      int64_t N, int D, void* workspace, size_t workspace_bytes, irbfn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* IRBFN_B200_H_ */
