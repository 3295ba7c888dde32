/*
 * multigrad_b200 -- C ABI of the sm_100a kernel library (libmgb200.so).
 *
 * Every entry point is `extern "C"`, takes raw device pointers plus a
 * `cudaStream_t` (passed as void*), runs on the *current* CUDA device and
 * returns 0 on success or a negative MGB_E* / positive cudaError_t code.
 * `mgb_last_error()` returns a human readable message for the calling thread.
 *
 * The reference (AlanPearl/multigrad) contains no native code at all
 * (SURVEY.md section 2.0); the comment above each declaration names the
 * Python component of the reference whose work the kernel replaces.
 */
#ifndef MULTIGRAD_B200_H_
#define MULTIGRAD_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGB_MAX_WORLD 16   /* ranks in one peer workspace (one NVSwitch box)  */
#define MGB_MAX_BINS 32    /* soft bins per model in the fused kernels        */
#define MGB_MAX_PARAMS 8   /* parameter-vector length of the scalar-link path */
#define MGB_MAX_BATCH 4096 /* parameter vectors per batched evaluation        */

/* error codes (negative; positive values are cudaError_t) */
#define MGB_OK 0
#define MGB_EINVAL (-1)
#define MGB_ETOOBIG (-2)
#define MGB_ETIMEOUT (-3)
#define MGB_ESTATE (-4)

const char* mgb_last_error(void);
int mgb_version(void);

/* ------------------------------------------------------------------------
 * Peer workspace (N1).  Replaces the MPI communicator plumbing of the
 * reference (multigrad/multigrad.py:15-27, 88-146): one cudaMalloc'ed block
 * per rank, exported with CUDA IPC so that every rank of a (sub)communicator
 * can load/store every other rank's block over NVLink from inside kernels.
 * Layout: [ctrl 4 KiB][LL slots: 2 parities x world x ll_units x 8 B]
 *         [bulk region: bulk_bytes][bulk region 2: bulk_bytes].
 * ---------------------------------------------------------------------- */
typedef struct MgbPeerTable {
  int32_t world;            /* ranks in this workspace                       */
  int32_t rank;             /* this rank                                     */
  uint32_t ll_units;        /* 8-byte LL slots per (parity, source rank)     */
  uint32_t spin_timeout_ms; /* watchdog for every in-kernel wait             */
  uint64_t bulk_bytes;      /* size of each of the two bulk regions          */
  void* base[MGB_MAX_WORLD];/* workspace base of every rank (peer-mapped)    */
} MgbPeerTable;

/* sizeof(MgbPeerTable, MgbErfBinModel, MgbStepArgs, MgbGroupArgs, MgbLinModel,
 * MgbLinArgs) -> out[0..n); returns how many there are.  Bindings compare
 * these with their own mirrors when they load the library. */
int mgb_abi_sizes(uint32_t* out, int n);

uint64_t mgb_ws_bytes(int world, uint32_t ll_units, uint64_t bulk_bytes);
int mgb_ws_alloc(uint64_t bytes, void** out_ptr);
int mgb_ws_free(void* ptr);
int mgb_ws_export(void* ptr, uint8_t handle_out[64]);
int mgb_ws_import(const uint8_t handle[64], void** out_ptr);
int mgb_ws_unimport(void* ptr);
uint64_t mgb_ws_bulk_offset(int world, uint32_t ll_units);
/* K7: device-to-device copy that accepts peer-mapped pointers (NVLink copy);
 * with the bulk region it replaces the pickled send/recv of util.scatter_nd
 * (multigrad/util.py:65-77). */
int mgb_peer_copy(void* dst, const void* src, uint64_t bytes, void* stream);
/* reads ctrl word `idx` (0 epoch, 1 error, 2 bulk epoch) of a local block */
int mgb_ws_read_ctrl(const void* local_base, int idx, uint32_t* out,
                     void* stream);

/* ------------------------------------------------------------------------
 * K5: stand-alone all-reduce (sum) over the peer workspace.  Replaces
 * reduce_sum's np.asarray + MPI Allreduce (multigrad/multigrad.py:149-183).
 * dtype: 0 = float32, 1 = float64.  In place on `data` (n elements).
 * Small messages use the one-shot ordered LL exchange (bit-identical result
 * on every rank); larger ones the two-shot reduce-scatter + all-gather over
 * peer loads/stores.  `algo`: 0 auto, 1 force LL, 2 force two-shot,
 * 3 force one-shot push (mid-size float32 messages).
 * ---------------------------------------------------------------------- */
int mgb_allreduce_sum(const MgbPeerTable* pt, void* data, uint64_t n,
                      int dtype, int algo, void* stream);
/* ------------------------------------------------------------------------
 * N1 (multicast) + K5 (NVLS): NVLink-switch multicast object of a
 * communicator and the in-switch all-reduce (multimem.ld_reduce /
 * multimem.st) on it.  Set-up order: root mgb_mc_create -> fd to the peers
 * (unix socket) -> peers mgb_mc_import; all mgb_mc_add_device; barrier; all
 * mgb_mc_bind; barrier.  No counterpart in the reference.
 * ---------------------------------------------------------------------- */
int mgb_mc_supported(int* out);
int mgb_mc_round_size(int world, uint64_t bytes, uint64_t* out);
int mgb_mc_create(int world, uint64_t bytes, uint64_t* handle_out, int* fd_out);
int mgb_mc_import(int fd, uint64_t* handle_out);
int mgb_mc_add_device(uint64_t handle);
int mgb_mc_bind(uint64_t handle, uint64_t bytes, void** uc_out, void** mc_out,
                uint64_t* mem_out);
int mgb_mc_release(uint64_t handle, uint64_t mem, void* uc, void* mc,
                   uint64_t bytes);
/* float32 sum in place on `data`; `uc`/`mc` are this rank's unicast and the
 * multicast view of the bound buffer (mc_bytes each); pt supplies the flags */
int mgb_allreduce_sum_nvls(const MgbPeerTable* pt, void* uc, void* mc,
                           uint64_t mc_bytes, float* data, uint64_t n,
                           void* stream);

/* push `n` floats from `root` to every rank over NVLink (replaces a pickled
 * parameter bcast like multigrad/bfgs.py:72-77 or mpi4jax.bcast,
 * multigrad/mpi4jax/multigrad.py:52).  Ready-first: every non-root rank
 * announces itself (one tagged word) and waits on the GPU for the data; the
 * root pushes once all have announced, so no collective can overrun a slow
 * rank whatever the mix of collectives and roots.  n <= ll_units - 1. */
int mgb_bcast_f32(const MgbPeerTable* pt, float* data, uint32_t n, int root,
                  void* stream);

/* ------------------------------------------------------------------------
 * K1/K2/K3: fused erf-soft-binned statistic.  One launch evaluates, for the
 * local shard x[0..n), the partial per-bin sums and their parameter
 * Jacobian, reduces them over CTAs and (if pt->world > 1) over GPUs inside
 * the kernel, and finishes with the loss / chain rule / optimiser epilogue.
 * Replaces OnePointModel._vjp (multigrad/multigrad.py:508-538: forward,
 * Allreduce, dL/ds, VJP, Allreduce, second loss evaluation) and, in the
 * optimiser modes, one iteration of _adam_optimizer (multigrad/adam.py:52-68)
 * including the bound transforms (adam.py:172-239) or of simple_grad_descent
 * (multigrad/util.py:100-114).
 * ---------------------------------------------------------------------- */
typedef struct MgbErfBinModel {
  const float* x;        /* object coordinate (e.g. log10 halo mass) [n]     */
  const float* w;        /* optional per-object weight [n] or NULL           */
  int64_t n;             /* objects in this rank's shard                     */
  int32_t nbins;         /* K <= MGB_MAX_BINS                                */
  int32_t nparams;       /* P <= MGB_MAX_PARAMS                              */
  int32_t mu_idx;        /* mu_h    = x_h + params[mu_idx]                   */
  int32_t sig_idx;       /* sigma   = link(params[sig_idx])                  */
  int32_t sigma_link;    /* 0: sigma = p, 1: sigma = 10**p                   */
  int32_t loss_kind;     /* 0: mean (log10(s+eps)-log10(t+eps))^2
                            1: mean (s-t)^2   2: sum ((s-t)*inv_err)^2       */
  float loss_eps;
  float x0;              /* centring constant: the kernel works on x - x0;
                            pick a value near the bulk of the data (any value
                            is valid, 0 disables centring)                  */
  int32_t step_flags;    /* MGB_STEP_* bits ORed into EVERY launch for this
                            model, whichever entry point builds its arguments
                            (e.g. MGB_STEP_STATIC_SPLIT for bit-reproducible
                            host evaluations and L-BFGS-B sessions)         */
  int32_t reserved0;
  double edges[MGB_MAX_BINS + 1]; /* fp64: the scaled edge constants are
                            formed in double, and exactly uniform spacing (as
                            from linspace) enables the Gaussian recurrence    */
  float norm[MGB_MAX_BINS];    /* s_k = norm_k * sum_h (Phi_{k+1}-Phi_k)     */
  float target[MGB_MAX_BINS];
  float inv_err[MGB_MAX_BINS];
} MgbErfBinModel;

enum {
  MGB_MODE_SUMSTATS = 0, /* out_sumstats[B][K]                               */
  MGB_MODE_SUMSTATS_JAC = 1, /* + out_jac[B][K][P]                           */
  MGB_MODE_LOSS_GRAD = 2,    /* + out_loss[B], out_grad[B][P]                */
  MGB_MODE_ADAM = 3,         /* + Adam update of the optimiser state         */
  MGB_MODE_GD = 4            /* + fixed-step gradient-descent update         */
};

/* MgbStepArgs.flags.  By default the last ~10% of the object tiles are handed
 * out through an atomic counter (evens out the SMs: 2-3% faster); which warp
 * adds which of those tiles varies, so results differ in the last bits from
 * run to run (the replicas on different GPUs still agree bit for bit).
 * STATIC_SPLIT: fully static tile assignment, bit-reproducible results (also
 * selected for every launch by the environment variable
 * MULTIGRAD_B200_DETERMINISTIC=1). */
#define MGB_STEP_STATIC_SPLIT 1
/* CONST_KEY (stochastic links, csrc/linkbin.cu): the optimiser modes use
 * rng_step instead of the running step index for the random key, i.e. the
 * same draws at every step (const_randkey of multigrad/multigrad.py:291-297). */
#define MGB_STEP_CONST_KEY 2

typedef struct MgbStepArgs {
  int32_t mode;
  int32_t batch;            /* B parameter vectors (optimiser modes: 1)      */
  int32_t reduce;           /* 1: sum over the peer workspace, 0: local only */
  int32_t jac_at_params;    /* bounded Adam: reference-compatible Jacobian   */
  const float* params;      /* [B][P] evaluation points (modes 0-2)          */
  const float* gate;        /* optional device flag: the launch is a no-op
                               when *gate == 0 (BFGS worker pipeline)        */
  float* out_sumstats;      /* [B][K]   or NULL                              */
  float* out_jac;           /* [B][K][P] or NULL                             */
  float* out_loss;          /* [B]      or NULL                              */
  float* out_grad;          /* [B][P]   or NULL                              */
  /* optimiser state, device resident, modes 3/4 */
  float* opt_p;             /* [P] current (bounded) parameters              */
  float* opt_u;             /* [P] unbounded parameters                      */
  float* opt_m;             /* [P] */
  float* opt_v;             /* [P] */
  int32_t* opt_step;        /* [1] number of updates applied so far          */
  float* traj;              /* [(nsteps+1)][P]; row step+1 is written (Adam),
                               row step (pre-update params) for GD           */
  float* loss_hist;         /* [nsteps]; entry step is written               */
  int32_t traj_rows;        /* capacity of traj in rows (bounds check)       */
  int32_t step_base;        /* row index = *opt_step - step_base (resume)    */
  float lr, b1, b2, eps;
  int32_t bound_kind[MGB_MAX_PARAMS]; /* 0 none 1 both 2 low-only 3 high-only */
  float bound_lo[MGB_MAX_PARAMS];
  float bound_hi[MGB_MAX_PARAMS];
  /* Host hand-shake (pre-launched host-driven steps, mgb_erfbin_adam_host):
   * when non-NULL the grid starts working only once the low 31 bits of
   * *wait_flag (pinned host memory, polled over PCIe) reach wait_value;
   * bit 31 set = cancelled, the grid exits without touching anything; after
   * wait_timeout_ms without either it exits as if cancelled. */
  const uint32_t* wait_flag;
  uint32_t wait_value;
  uint32_t wait_timeout_ms;
  /* Persistent fit (modes 3/4): nsteps > 1 runs that many optimiser steps in
   * ONE launch (all CTAs co-resident; between steps they wait on the device
   * for the elected CTA that reduces, exchanges over NVLink and updates the
   * parameters) -- the analogue of the reference's collectives inside one
   * compiled lax.scan (multigrad/mpi4jax/multigrad.py:38-58).  0 or 1: one
   * step per launch. */
  int32_t nsteps;
  int32_t flags;            /* MGB_STEP_* bits                               */
  /* Optional device buffer of %globaltimer stamps (tools/step_trace.py):
   * per step 8 + 2*grid uint64 words; NULL = no tracing. */
  uint64_t* trace;
  /* step index mixed into the device-side random key of a stochastic link
   * (evaluation modes; the optimiser modes use *opt_step) */
  uint32_t rng_step;
  uint32_t reserved1;
} MgbStepArgs;

/* scratch needed by mgb_erfbin_step: bytes for `scratch` given the batch */
uint64_t mgb_erfbin_scratch_bytes(int nbins, int batch);
int mgb_erfbin_grid(int nbins);
/* thread-block cluster size the fused step is launched with (diagnostics) */
int mgb_erfbin_cluster(int nbins, int weighted);
/* `scratch` must be zero-filled once before its first use. */
int mgb_erfbin_step(const MgbErfBinModel* model, const MgbStepArgs* args,
                    const MgbPeerTable* pt, void* scratch,
                    uint64_t scratch_bytes, void* stream);

/* Host-driven evaluation in one call: params (pinned host) -> device, fused
 * loss+gradient launch, (loss, grad) -> pinned host, stream sync.  The body of
 * one L-BFGS-B line-search evaluation (reference multigrad/bfgs.py:72-77 plus
 * multigrad.py:508-538).  out_host receives 1 + nparams floats. */
int mgb_erfbin_eval_host(const MgbErfBinModel* model, const MgbPeerTable* pt,
                         const float* params_host, float* out_host,
                         float* d_params, float* d_out, void* scratch,
                         uint64_t scratch_bytes, void* stream);
/* Zero-copy flavour: the kernel reads params_host and writes out_host (both
 * pinned host memory) directly over PCIe, the host polls the result words:
 * one launch, no copy-engine operation, no stream synchronise. */
int mgb_erfbin_eval_host_zc(const MgbErfBinModel* model,
                            const MgbPeerTable* pt, const float* params_host,
                            float* out_host, void* scratch,
                            uint64_t scratch_bytes, void* stream,
                            uint32_t timeout_ms);
/* Host-driven Adam (optimiser state on the host, as in the reference's
 * multigrad/adam.py:52-68) without a Python round trip per step: nsteps x
 * [zero-copy evaluation, Adam update in host memory].  params_host/out_host
 * pinned; traj ((nsteps+1) x P, row 0 = start) and loss_hist may be NULL. */
/* Session of host-driven evaluations with one grid always pre-launched and
 * waiting on the GPU (hides the launch latency for drivers that return to
 * Python between evaluations, e.g. scipy's L-BFGS-B).  op: 0 begin, 1 eval
 * (params_host -> out_host[0..P]), 2 end.  out_host as for
 * mgb_erfbin_adam_host; *seq is the caller-owned evaluation counter. */
int mgb_erfbin_host_session(const MgbErfBinModel* model,
                            const MgbPeerTable* pt, float* params_host,
                            float* out_host, void* scratch,
                            uint64_t scratch_bytes, void* stream,
                            uint32_t timeout_ms, int32_t op, uint32_t* seq);
#define MGB_HOST_FLAG_WORD 16  /* out_host needs MGB_HOST_FLAG_WORD + 1 words */
int mgb_erfbin_adam_host(const MgbErfBinModel* model, const MgbPeerTable* pt,
                         float* params_host, float* out_host, float* m,
                         float* v, int32_t step0, int32_t nsteps, float lr,
                         float b1, float b2, float eps, float* traj,
                         float* loss_hist, void* scratch,
                         uint64_t scratch_bytes, void* stream,
                         uint32_t timeout_ms, void* flush_buf,
                         uint64_t flush_bytes, double* timed_seconds);
/* flush_buf (optional, benchmarking): device buffer larger than L2 that is
 * overwritten in front of every step; *timed_seconds (optional) receives the
 * sum over steps of [params published -> loss+grad read back], which leaves
 * the wait for that fill out. */

/* Same fused step, for callers that cannot allocate scratch (a plug-in that
 * only receives its input/output buffers from a host runtime, a process whose
 * device memory is owned by another pool, a phase in which calling the CUDA
 * allocator is off limits).  No counterpart in the reference: the reference
 * has no native code and never manages device memory itself.
 * The scratch is a built-in __device__ array of MGB_STATIC_SCRATCH_BYTES that
 * every GPU loading the library keeps until exit; requests that need more
 * (see mgb_erfbin_scratch_bytes) are refused with MGB_ETOOBIG.  The array is
 * shared by all calls on a GPU: calls that overlap there race.  The first
 * call on a device resolves the array's device address; make it before any
 * stream capture.  Calls no cudaMalloc/cudaFree. */
#define MGB_STATIC_SCRATCH_BYTES (1u << 20)
int mgb_erfbin_step_static(const MgbErfBinModel* model,
                           const MgbStepArgs* args, const MgbPeerTable* pt,
                           void* stream);

/* ------------------------------------------------------------------------
 * K1/K2/K3 for user-defined links (csrc/linkbin.cu, compiled per model from a
 * header generated by multigrad_b200/link.py): mu_h = x_h + shift(x_h, f_h;
 * p), sigma_h = sigma(x_h, f_h; p) with up to MGB_LINK_MAX_FEAT per-object
 * feature columns.  Replaces jax.vjp of an arbitrary user sumstats function
 * (OnePointModel._vjp, multigrad/multigrad.py:508-538) for the soft-binned
 * family; the per-step random key (multigrad/adam.py:60-62) is derived on the
 * device from (seed, step).  Entry points live in the per-model library:
 * mgb_linkbin_step / mgb_linkbin_scratch_bytes / mgb_linkbin_nparams / ...
 * ---------------------------------------------------------------------- */
#define MGB_LINK_MAX_FEAT 4
typedef struct MgbLinkModel {
  const float* x;        /* object coordinate [n]                            */
  const float* w;        /* optional per-object weight [n] or NULL           */
  const float* feat[MGB_LINK_MAX_FEAT]; /* feature columns [n] each          */
  int64_t n;
  int64_t index0;        /* global index of this shard's first object (RNG)  */
  uint64_t seed;         /* random seed of a stochastic link                 */
  int32_t nbins, nparams, nfeat, loss_kind;
  float loss_eps, x0;
  double edges[MGB_MAX_BINS + 1];
  float norm[MGB_MAX_BINS];
  float target[MGB_MAX_BINS];
  float inv_err[MGB_MAX_BINS];
} MgbLinkModel;
uint64_t mgb_linkbin_scratch_bytes(int nbins);
/* `scratch` must be zero-filled once before its first use. */
int mgb_linkbin_step(const MgbLinkModel* model, const MgbStepArgs* args,
                     const MgbPeerTable* pt, void* scratch,
                     uint64_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------
 * Linear-link soft-binned statistic on the tensor cores (tcgen05 + TMEM +
 * TMA).  mu_h = x_h + sum_j W[h][j] theta_j with per-object bf16 features W
 * (row-major [n][pfeat], pfeat = 64 or 128, 16-byte aligned), sigma =
 * link(params[nfeat]).  params = (theta_0 .. theta_{nfeat-1}, p_sigma).
 * One launch streams W once (TMA), runs the forward product W.theta and the
 * Jacobian contraction G^T.W as UMMA tiles accumulating in TMEM, evaluates
 * erfc/exp on the CUDA cores, reduces over CTAs and GPUs and finishes with
 * the same loss / gradient / Adam epilogue as mgb_erfbin_step.
 * Replaces jax.vjp of a user sumstats function with a linear predictor
 * (OnePointModel._vjp, multigrad/multigrad.py:508-538) + adam.py:52-68.
 * ---------------------------------------------------------------------- */
#define MGB_LIN_MAX_BINS 11
#define MGB_LIN_MAX_FEAT 128

typedef struct MgbLinModel {
  const float* x;        /* [n] object coordinate (16-byte aligned)          */
  const void* w;         /* [n][pfeat] bf16 features                         */
  const float* ow;       /* [n] per-object weights or NULL (16-byte aligned) */
  int64_t n;
  int32_t nbins;         /* K <= MGB_LIN_MAX_BINS                            */
  int32_t nfeat;         /* features in use (<= pfeat); nparams = nfeat + 1  */
  int32_t pfeat;         /* padded row length of W: 64 or 128                */
  int32_t sigma_link;    /* 0: sigma = p, 1: sigma = 10**p                   */
  int32_t loss_kind;     /* as in MgbErfBinModel                             */
  float loss_eps;
  float x0;              /* centring constant                                */
  float edges[MGB_LIN_MAX_BINS + 1];
  float norm[MGB_LIN_MAX_BINS];
  float target[MGB_LIN_MAX_BINS];
  float inv_err[MGB_LIN_MAX_BINS];
} MgbLinModel;

typedef struct MgbLinArgs {
  int32_t mode;             /* MGB_MODE_SUMSTATS_JAC .. MGB_MODE_GD          */
  int32_t reduce;           /* 1: sum over the peer workspace                */
  int32_t jac_at_params;
  int32_t traj_rows;
  int32_t step_base;
  float lr, b1, b2, eps;
  const float* params;      /* [nfeat+1] (modes 1-2)                         */
  float* out_sumstats;      /* [K]                                           */
  float* out_jac;           /* [K][nfeat+1] or NULL                          */
  float* out_loss;          /* [1] or NULL                                   */
  float* out_grad;          /* [nfeat+1] or NULL                             */
  float* opt_p;             /* optimiser state, [nfeat+1] each               */
  float* opt_u;
  float* opt_m;
  float* opt_v;
  int32_t* opt_step;
  float* traj;              /* [traj_rows][nfeat+1]                          */
  float* loss_hist;
  const int32_t* bound_kind; /* device arrays [nfeat+1] or NULL              */
  const float* bound_lo;
  const float* bound_hi;
} MgbLinArgs;

uint64_t mgb_linlink_scratch_bytes(int nbins, int pfeat);
/* `scratch` must be zero-filled once before its first use. */
int mgb_linlink_step(const MgbLinModel* model, const MgbLinArgs* args,
                     const MgbPeerTable* pt, void* scratch,
                     uint64_t scratch_bytes, void* stream);
/* building-block check used by the tests (one tile through both UMMAs) */
int mgb_linlink_selftest(const void* w_bf16, const float* theta,
                         const float* gmat, float* out_mu, float* out_j,
                         int* status_dev, int P, void* stream);

/* ------------------------------------------------------------------------
 * K4 + K3: OnePointGroup step.  The member models' kernels leave their
 * (loss, grad) on the device; this one launch sums the local contributions,
 * all-reduces 1 + P doubles over the *main* communicator's peer workspace and
 * applies the optimiser update (args->mode = MGB_MODE_ADAM / MGB_MODE_GD) or
 * just writes the totals (MGB_MODE_LOSS_GRAD -> args->out_loss/out_grad).
 * Replaces the pickled allgathers + host sum of the reference
 * (multigrad/multigrad.py:560-581).  main_pt may be NULL (single process).
 * ---------------------------------------------------------------------- */
#define MGB_MAX_GROUP_MODELS 8
typedef struct MgbGroupArgs {
  int32_t n_models;         /* models evaluated by this rank                 */
  int32_t nparams;          /* P <= MGB_MAX_PARAMS, shared by all models     */
  int32_t active[MGB_MAX_GROUP_MODELS]; /* 1: this rank is the root of the
                               model's sub-communicator and contributes      */
  const float* loss[MGB_MAX_GROUP_MODELS];  /* [1] device                    */
  const float* grad[MGB_MAX_GROUP_MODELS];  /* [P] device                    */
} MgbGroupArgs;
int mgb_group_step(const MgbGroupArgs* g, const MgbStepArgs* args,
                   const MgbPeerTable* main_pt, void* stream);

/* ------------------------------------------------------------------------
 * K3 stand-alone: Adam / bound transforms on device vectors for the generic
 * (user-defined torch model) path.  Replaces jax_opt.adam's opt_update and
 * the jitted per-scalar transforms (multigrad/adam.py:52-68, 192-239).
 * ---------------------------------------------------------------------- */
int mgb_adam_update(float* params, const float* grad, float* m, float* v,
                    int32_t step, int32_t n, float lr, float b1, float b2,
                    float eps, void* stream);

/* ------------------------------------------------------------------------
 * K8: pointer jumping (indices = indices[indices] to a fixed point), the
 * device version of diffdesi_experimental.find_ultimate_top_indices
 * (multigrad/diffdesi_experimental/util.py:18-28).  `rounds_out` receives
 * the number of rounds; returns MGB_ESTATE if `max_rounds` is exceeded.
 * ---------------------------------------------------------------------- */
int mgb_pointer_jump(int64_t* indices, int64_t* tmp, int64_t n,
                     int32_t max_rounds, int32_t* rounds_out,
                     int32_t* changed_flag_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MULTIGRAD_B200_H_ */
