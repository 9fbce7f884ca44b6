/*
 * vtrace_b200.h -- C ABI of libvtrace_b200.so
 *
 * B200 (sm_100a) implementation of IMPALA's learner-side off-policy
 * correction: V-trace plus the three IMPALA loss terms, forward and backward.
 * Citations are to the reference tree (Zhehui-Huang/scalable_agent).
 *
 * The reference has no FFI boundary for this path: it is Python that adds
 * TensorFlow graph nodes (vtrace.py, experiment.py:324-343) and the authors
 * name "a custom TensorFlow operation" as the intended upgrade
 * (experiment.py:387-388).  This header is that custom-op boundary.  Each
 * entry point below cites the reference function it replaces.
 *
 * Conventions
 *   - All tensors are C-contiguous, time-major: logits [T,B,A] (A fastest),
 *     per-step tensors [T,B], bootstrap [B].  float = IEEE fp32, actions int32.
 *   - Every pointer is a DEVICE pointer on the current device unless the
 *     parameter comment says "host".  The caller owns every buffer, including
 *     workspaces; the library never allocates, frees or keeps a pointer.
 *   - Work is enqueued on `stream` (a cudaStream_t passed as void*); no entry
 *     point synchronises the host.
 *   - Return value: 0 = VTR_OK; negative = argument error detected before any
 *     launch (VTR_ERR_*); positive = the cudaError_t of a failed launch.
 *     vtr_last_error() returns a thread-local message for the last failure.
 *     Nothing aborts, exits or throws.
 *   - Pointers documented "may be NULL" switch the corresponding output off.
 *   - Out-of-range actions are clamped to [0, A-1] for memory safety; the
 *     results for such steps are unspecified (TF raises on CPU, vtrace.py:67).
 *
 * State the library keeps between calls (device memory compiled into the
 * library, per GPU; none of it can change a result)
 *   - Counter sets for the fused pass: 512 sets of 256 uint32, zero at load.
 *     A launch that finishes its loss reduction in the kernel takes the next
 *     of 256 sets in turn (tile cursor and tickets of the per-SM kernel, the
 *     "who finished last" ticket of the others) and its last CTA sets it back
 *     to zero, so no call needs a memset.  A set comes round again 256 fused
 *     calls later: keep fewer than 256 fused calls pending per GPU.  A call
 *     captured into a CUDA graph keeps one of the other 256 sets for good;
 *     once those are used up, captured calls zero counters in the workspace
 *     header (memset node) or leave the last stage to a second launch.
 *   - The per-SM speed table of the per-SM kernel (VTR_PATH_SM_TILE): each
 *     tile records how fast its SM streamed, the next call sizes its tiles
 *     from that.  Only the partition of the batch depends on it; every output
 *     is independent of the partition.  Concurrent callers on two streams
 *     update it without synchronisation; the worst that can happen is a less
 *     even partition on the next call.
 *   - The 200 KiB array of vtr_from_logits_fused_noalloc_f32 (see there).
 *
 * How the fused pass is launched: with programmatic stream serialization
 * (cudaLaunchAttributeProgrammaticStreamSerialization).  Its CTAs may be
 * placed on SMs while the previous kernel of the stream drains and do their
 * set-up there; they touch no tensor before griddepcontrol.wait, i.e. before
 * that kernel has completed and its writes are visible.  Stream order as the
 * caller sees it is the ordinary one.  VTR_NO_PDL=1 in the environment makes
 * the launches plain ones.
 */
#ifndef VTRACE_B200_H_
#define VTRACE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VTR_VERSION 200 /* major*10000 + minor*100 + patch */

#define VTR_OK 0
#define VTR_ERR_INVALID_ARGUMENT (-1) /* NULL required pointer, T/B/A <= 0 ... */
#define VTR_ERR_UNSUPPORTED (-2)      /* A above VTR_MAX_ACTIONS, bad mode    */
#define VTR_ERR_WORKSPACE (-3)        /* workspace missing or too small       */
#define VTR_ERR_MISALIGNED (-4)       /* pointer not aligned to its element   */

#define VTR_MAX_ACTIONS 256

/* flags for vtr_fused_params.flags / vtr_from_importance_weights_f32 */
#define VTR_FLAG_NO_CLIP_RHO 1u    /* clip_rho_threshold=None    (vtrace.py:236-239) */
#define VTR_FLAG_NO_CLIP_PG_RHO 2u /* clip_pg_rho_threshold=None (vtrace.py:270-274) */

/* reward clipping fused into the prologue (experiment.py:377-382) */
#define VTR_REWARD_CLIP_NONE 0
#define VTR_REWARD_CLIP_ABS_ONE 1         /* clip(r, -1, 1)            :378     */
#define VTR_REWARD_CLIP_SOFT_ASYMMETRIC 2 /* tanh(r/5)*5, x0.3 if r<0  :380-382 */

/* kernel selection for the fused entry point (0 = library decides) */
#define VTR_PATH_AUTO 0
#define VTR_PATH_COLUMN 1  /* 32-column tiles, one ring per tile (any alignment) */
#define VTR_PATH_SM_TILE 2 /* one wide tile per SM, shared ring (16-B aligned)  */
#define VTR_PATH_ITEM 3    /* one thread per (t, b) item, warp-shuffle affine scan
                              over T: tiny B, T <= 256, compiled-in A            */
#define VTR_PATH_CHUNK 4   /* T split over a thread-block cluster, chunk aggregates
                              of the recurrence through distributed shared
                              memory: small and mid-sized B, T <= 104            */

int vtr_version(void);

/* Thread-local, never NULL; valid until the next failing call on this thread. */
const char* vtr_last_error(void);

/*
 * Arguments of the fused forward+backward pass.
 *
 * Replaces, in one launch:
 *   vtrace.from_logits                        vtrace.py:71-161
 *     log_probs_from_logits_and_actions x2    vtrace.py:45-68, 143-146
 *     from_importance_weights                 vtrace.py:164-280
 *   compute_policy_gradient_loss              experiment.py:338-343
 *   compute_baseline_loss(vs - values)        experiment.py:324-328, 404-405
 *   compute_entropy_loss                      experiment.py:331-335
 *   total = pg + baseline_cost*bl + entropy_cost*ent   experiment.py:401-407
 *   the TF autodiff backward of that total    experiment.py:415
 *   optionally the reward clip / discount glue         experiment.py:377-384
 */
typedef struct vtr_fused_params {
  int32_t T; /* unroll length          (experiment.py:65)  */
  int32_t B; /* batch                  (experiment.py:64)  */
  int32_t A; /* number of actions, 1..VTR_MAX_ACTIONS       */
  uint32_t flags; /* VTR_FLAG_*                              */

  /* inputs */
  const float* behaviour_policy_logits; /* [T,B,A]                        */
  const float* target_policy_logits;    /* [T,B,A]                        */
  const int32_t* actions;               /* [T,B]                          */
  const float* discounts;               /* [T,B]; may be NULL iff done set */
  const float* rewards;                 /* [T,B]                          */
  const float* values;                  /* [T,B]                          */
  const float* bootstrap_value;         /* [B]                            */
  const uint8_t* done; /* [T,B] bool, may be NULL.  If set, discounts =
                          (!done)*discounting (experiment.py:384) and the
                          `discounts` pointer is ignored.                 */

  float clip_rho_threshold;    /* ignored with VTR_FLAG_NO_CLIP_RHO       */
  float clip_pg_rho_threshold; /* ignored with VTR_FLAG_NO_CLIP_PG_RHO    */
  float baseline_cost;         /* experiment.py:71                        */
  float entropy_cost;          /* experiment.py:70                        */
  float discounting;           /* experiment.py:72, used only with `done` */
  int32_t reward_clip_mode;    /* VTR_REWARD_CLIP_*                       */
  int32_t path;                /* VTR_PATH_*                              */
  int32_t reserved;

  /* V-trace outputs, each [T,B]; any may be NULL (VTraceFromLogitsReturns,
     vtrace.py:37-40) */
  float* vs;
  float* pg_advantages;
  float* log_rhos;
  float* behaviour_action_log_probs;
  float* target_action_log_probs;

  /* gradients of the total loss; both NULL = forward only.
     d_target_logits = (softmax - onehot(a))*pg_adv
                       + entropy_cost * p*(log p + H)          [T,B,A]
     d_values        = baseline_cost * (values - vs)            [T,B]   */
  float* d_target_logits;
  float* d_values;

  /* losses, device memory; may be NULL (then no loss is computed).
     losses[0]=policy-gradient, [1]=baseline (unweighted 0.5*sum sq),
     [2]=entropy (unweighted, = -sum H), [3]=weighted total.            */
  float* losses;      /* float[4]  */
  double* losses_f64; /* double[4], may be NULL */

  /* Optional, for a batch sharded over the GPUs of one box: publish the four
     loss scalars to every peer as the last thing the pass does, exactly as
     vtr_exchange_publish_f32 would (same buffers, same seq rules; see
     below).  The per-SM kernel does the peer stores itself from the tile that
     finishes the reduction; on the other kernels the library appends the
     publish launch.  NULL / 0 = off.  Needs `losses`. */
  void* const* exchange_peer_buffers; /* HOST array of exchange_world device pointers */
  int32_t exchange_rank;
  int32_t exchange_world;
  uint32_t exchange_seq;
  /* Which step's sums `exchange_sums` receives: 0 = this call's (the launch
     then ends only when the slowest rank has published: every step pays the
     rendezvous), 1 = the previous call's, seq - 1 (they are normally in this
     rank's buffer already, so no rank waits for another and the ranks may
     drift by a step; nothing is written when seq == 1).  The last step's sums
     are then fetched with vtr_exchange_collect_f32. */
  uint32_t exchange_lag;
  /* float[4] or NULL: if given, the pass also does vtr_exchange_collect_f32's
     part (wait for every rank's scalars of step seq - exchange_lag, at most
     2 s, and write their rank-ordered sum here), so a step needs no exchange
     call at all. */
  float* exchange_sums;

  /* float[1] or NULL: the weighted total (= losses[3]) once more, in a buffer
     of its own - for a caller whose framework wants the differentiable total
     and the constant component losses in separate tensors (an autograd
     output may not alias a non-differentiable one).  Needs `losses`. */
  float* total_loss;
} vtr_fused_params;

/* Bytes of device workspace that are enough for vtr_from_logits_fused_f32 at
   (T,B,A) whichever kernel runs: an 8 KB header (group partials of the per-SM
   kernel; counters only in the fallback described above, which the library
   zeroes itself) followed by the loss partials of the deterministic staged
   reduction, one double[3] per 32-column tile, per 4-column granule, per
   column or per CTA depending on the kernel.  The call itself checks against
   what the kernel it picks really writes, which can be less.  Returns 0 on
   invalid sizes. */
size_t vtr_fused_workspace_bytes(int32_t T, int32_t B, int32_t A);

/* The fused pass.  `workspace` must hold vtr_fused_workspace_bytes() bytes,
   8-byte aligned; it is required only when p->losses is set. */
int vtr_from_logits_fused_f32(const vtr_fused_params* p, void* workspace,
                              size_t workspace_bytes, void* stream);

/* What vtr_from_logits_fused_f32 would launch for `p` on the current device,
   without launching anything: stores the number of kernels in *num_kernels
   and writes a one-line description (kernel names, grid, ring depth, shared
   memory) to `text`.  `have_workspace` = whether a workspace would be passed.
   Host-side only; for benchmarks and logs. */
int vtr_fused_describe(const vtr_fused_params* p, int have_workspace, int* num_kernels,
                       char* text, size_t text_bytes);

/* Most batch columns the allocation-free entry point below accepts. */
#define VTR_NOALLOC_MAX_B 262144

/* No counterpart in the reference: the same fused pass for callers that cannot
   allocate device memory, and so cannot supply the workspace above -- an op
   inside a framework that hands it inputs and pre-declared outputs but no
   temporary allocator, or a learner running in a process where cudaMalloc is
   forbidden after start-up because the memory pool owns the whole GPU and a
   cudaMalloc would fail or stall the other streams by synchronising the
   device.  The tile partials go to a 200 KiB array built into the library
   (every GPU that loads the library holds it until the process exits); the
   caller still passes all inputs and outputs in `p`.  Per-column outputs are
   what vtr_from_logits_fused_f32 gives; the losses may differ from its in the
   last bit, because once a batch is too wide for the preferred kernel's
   partials to fit the array another kernel, with another summation order,
   takes over (8192 < B/4, i.e. B > 32768 on the per-SM kernel).
   - p->B must be <= VTR_NOALLOC_MAX_B, else VTR_ERR_UNSUPPORTED; so is an
     explicit p->path whose kernel's partials do not fit the array.
   - With p->losses and p->losses_f64 both NULL the array is not used.
   - No cudaMalloc/cudaFree or allocating library routine is called.
   - The first call on each device looks the array's address up (with that
     device current); make that first call before any stream capture.
   - All calls on a device share the one array, which keeps the previous call's
     contents.  Nothing needs zeroing (each call rewrites every slot it reads),
     but two calls that overlap on the GPU -- different streams, no ordering
     between them -- race on it and corrupt each other's losses.  Serialise
     them, or use the workspace variant. */
int vtr_from_logits_fused_noalloc_f32(const vtr_fused_params* p, void* stream);

/* Second stage of the loss reduction on its own: sums `num_partials` rows of
   double[3] in a fixed order.  (vtr_from_logits_fused_f32 calls this
   internally; exposed for callers that shard a batch over several calls.)
   Replaces the tf.reduce_sum calls at experiment.py:328,335,343. */
int vtr_reduce_losses(const double* partials, int32_t num_partials,
                      float baseline_cost, float entropy_cost, float* losses,
                      double* losses_f64, void* stream);

/*
 * vtrace.from_importance_weights (vtrace.py:164-280) on its own.
 * Shapes: values, rewards [T,B,C]; log_rhos, discounts [T,B,C] or [T,B,1];
 * bootstrap_value [B,C]; C = 1 is the plain rank-2 case.  `bcast_mask` bit 0
 * set = log_rhos has trailing size 1, bit 1 = discounts has trailing size 1
 * (the broadcasting case of vtrace.py:178-180, vtrace_test.py:249-259).
 * Outputs vs, pg_advantages [T,B,C].
 */
int vtr_from_importance_weights_f32(const float* log_rhos, const float* discounts,
                                    const float* rewards, const float* values,
                                    const float* bootstrap_value, int32_t T,
                                    int32_t B, int32_t C, uint32_t bcast_mask,
                                    float clip_rho_threshold,
                                    float clip_pg_rho_threshold, uint32_t flags,
                                    float* vs, float* pg_advantages, void* stream);

/* vtrace.log_probs_from_logits_and_actions (vtrace.py:45-68): rows = T*B. */
int vtr_log_probs_from_logits_and_actions_f32(const float* policy_logits,
                                              const int32_t* actions,
                                              int64_t rows, int32_t A,
                                              float* log_probs, void* stream);

/* Workspace for the three stand-alone loss helpers below, for `rows` rows
   (or `rows` elements for the baseline loss). */
size_t vtr_loss_workspace_bytes(int64_t rows);

/* compute_policy_gradient_loss (experiment.py:338-343).
   loss = sum_rows CE(logits, a) * advantages.  loss: device float[1]. */
int vtr_policy_gradient_loss_f32(const float* logits, const int32_t* actions,
                                 const float* advantages, int64_t rows, int32_t A,
                                 float* loss, void* workspace,
                                 size_t workspace_bytes, void* stream);
/* its backward: d_logits = grad_out * (softmax - onehot(a)) * advantages.
   grad_out: device float[1] (the upstream scalar gradient). */
int vtr_policy_gradient_loss_bwd_f32(const float* logits, const int32_t* actions,
                                     const float* advantages,
                                     const float* grad_out, int64_t rows,
                                     int32_t A, float* d_logits, void* stream);

/* compute_entropy_loss (experiment.py:331-335): loss = sum_rows sum_a p log p. */
int vtr_entropy_loss_f32(const float* logits, int64_t rows, int32_t A, float* loss,
                         void* workspace, size_t workspace_bytes, void* stream);
/* its backward: d_logits = grad_out * p * (log p + H). */
int vtr_entropy_loss_bwd_f32(const float* logits, const float* grad_out,
                             int64_t rows, int32_t A, float* d_logits,
                             void* stream);

/* compute_baseline_loss (experiment.py:324-328): loss = 0.5 * sum adv^2. */
int vtr_baseline_loss_f32(const float* advantages, int64_t n, float* loss,
                          void* workspace, size_t workspace_bytes, void* stream);
/* its backward: d_advantages = grad_out * advantages. */
int vtr_baseline_loss_bwd_f32(const float* advantages, const float* grad_out,
                              int64_t n, float* d_advantages, void* stream);

/*
 * Cross-GPU sum of a few scalars (the loss scalars of a batch sharded over
 * the GPUs of one box) through NVLink peer memory, without a collective
 * library call.  No counterpart in the reference, which has a single learner
 * (experiment.py:491).  Each rank owns a zero-initialised buffer of
 * vtr_exchange_buffer_bytes(world) bytes that every peer can store to (e.g.
 * torch symmetric memory); `peer_buffers` is a HOST array of the `world`
 * device pointers to those buffers as mapped on this rank, its own included.
 * Per step, with `seq` = 1, 2, 3, ... identical on all ranks:
 *   vtr_exchange_publish_f32: stores values[0..n), each as one 8-byte word
 *     {value, seq}, into slot [seq % 4][rank] of every peer's buffer (n <= 6);
 *     a word says for itself which step it belongs to, so there is no flag
 *     and no fence between values and flag;
 *   vtr_exchange_collect_f32: waits (at most 2 s, then writes NaN) until the
 *     n words of all `world` slots of step seq in this rank's buffer carry
 *     seq, and writes the sums of their values, accumulated in double in rank
 *     order, to sums[0..n).
 * A slot is reused four steps later: a rank must have collected step seq - 2
 * (itself or through exchange_sums) before it publishes step seq; it may
 * publish one step without collecting the one before (exchange_lag = 1).
 */
#define VTR_EXCHANGE_MAX_RANKS 16
#define VTR_EXCHANGE_SLOT_FLOATS 16
#define VTR_EXCHANGE_SLOTS 4
size_t vtr_exchange_buffer_bytes(int32_t world);
int vtr_exchange_publish_f32(const float* values, int32_t n, void* const* peer_buffers, int32_t rank,
                             int32_t world, uint32_t seq, void* stream);
int vtr_exchange_collect_f32(const void* my_buffer, int32_t n, int32_t world, uint32_t seq,
                             float* sums, void* stream);

/*
 * The optimiser step that follows the path: experiment.py:410-415,
 *   learning_rate = tf.train.polynomial_decay(lr, num_env_frames, total_frames, 0)
 *   tf.train.RMSPropOptimizer(learning_rate, decay, momentum, epsilon).minimize(total_loss)
 * with TensorFlow 1.9.0's ApplyRMSProp:
 *   ms <- ms + (1 - decay)(g^2 - ms);  mom <- momentum mom + lr g / sqrt(ms + epsilon);  var <- var - mom.
 * The slots start as tf.train.RMSPropOptimizer creates them: ms = 1, mom = 0
 * (the caller initialises them).  All buffers float32, 16-byte aligned.
 */
/* experiment.py:411-412 (cycle=False); plain host arithmetic in double. */
double vtr_polynomial_decay(double learning_rate, int64_t global_step, int64_t decay_steps,
                            double end_learning_rate, double power);
/* experiment.py:413-415 on one GPU, in place.  `mom` may be NULL when momentum
   == 0 (the reference's default, experiment.py:94): the slot would only ever
   hold the last update. */
int vtr_rmsprop_f32(float* var, float* ms, float* mom, const float* grad, int64_t n,
                    float learning_rate, float decay, float momentum, float epsilon, void* stream);
/*
 * The same step for a learner whose batch is sharded over the GPUs of one box
 * (no counterpart in the reference, which has a single learner,
 * experiment.py:491): the gradient is the sum of the ranks' gradients.  One
 * launch per rank does reduce-scatter, update and all-gather over NVLink peer
 * memory.  Every rank owns three buffers its peers can access (e.g. torch
 * symmetric memory): the full parameters [n], its full local gradient [n], and
 * vtr_rmsprop_sync_bytes() of zero-initialised sync words; `*_peers` are HOST
 * arrays of the `world` device pointers to them as mapped on this rank, own
 * included.  The slots are sharded: rank r holds only ms/mom of its slice
 * [lo, hi) of the parameters (vtr_rmsprop_slice; element 0 of `ms_slice` is
 * parameter lo).  Call once per step on every rank with the same n, seq = 1,
 * 2, 3, ... and hyper-parameters; when the call has completed on a rank, that
 * rank's parameter buffer holds the updated values of all slices and its
 * gradient buffer may be overwritten.  Gradients are summed in rank order, so
 * all ranks hold the same bits.
 *
 * A peer that does not show up must not hang the GPU: a rank waits at most the
 * peer time-out (vtr_set_peer_timeout_ms, 20 s unless changed) and then gives
 * the step up.  The decision is taken once per launch, so a rank's slice is
 * updated completely or not at all; the rank that gives up sets word 33 of its
 * sync buffer (sticky) and, if `host_status` is given, stores seq there - a
 * uint32 in pinned host memory that the caller can look at without
 * synchronising the device - and marks its flags in every peer's sync buffer,
 * so that a peer waiting for it, or arriving later, gives the step up too and
 * flags it the same way.  After a flagged step the ranks' parameters may
 * differ; the caller has to re-synchronise them (e.g. broadcast).
 */
int vtr_rmsprop_slice(int64_t n, int32_t world, int32_t rank, int64_t* lo, int64_t* hi);
size_t vtr_rmsprop_sync_bytes(void);
int vtr_rmsprop_allreduce_f32(void* const* var_peers, void* const* grad_peers, void* const* sync_peers,
                              float* ms_slice, float* mom_slice, int64_t n, int32_t rank, int32_t world,
                              uint32_t seq, float learning_rate, float decay, float momentum,
                              float epsilon, uint32_t* host_status, void* stream);
/* How long vtr_rmsprop_allreduce_f32 waits for a peer (process-wide; read at
   launch).  Learner ranks drift apart by much more than a kernel's length now
   and then - a late actor batch, a checkpoint - so keep it well above that. */
void vtr_set_peer_timeout_ms(uint64_t milliseconds);
uint64_t vtr_get_peer_timeout_ms(void);

/*
 * The policy and baseline heads that produce the path's inputs
 * (experiment.py:199-203: snt.Linear(num_actions) and snt.Linear(1) on the
 * core output, applied over [T*B, K] by snt.BatchApply at :237; K = 256):
 *   policy_logits[r, a] = sum_k core_output[r, k] w_policy[k, a] + b_policy[a]
 *   baseline[r]         = sum_k core_output[r, k] w_baseline[k]  + b_baseline
 * One tensor-core kernel (tcgen05.mma, accumulators in TMEM, tensor-map TMA
 * loads), fp32-accurate: every product is formed as three TF32 products of
 * the operands' high and low halves, so the logits are good to ~1e-6 and the
 * V-trace outputs computed from them keep their 1e-5 bound.
 * K must be a multiple of 32 up to 256 and A at most 31.  The weights are
 * packed once per update (w_policy [K, A] as Sonnet stores it, b_policy [A],
 * w_baseline [K], b_baseline [1]) into `packed`, a 128-byte aligned device
 * buffer of vtr_policy_head_packed_bytes(K, A) bytes; core_output is [rows, K],
 * 16-byte aligned; policy_logits [rows, A] and baseline [rows] are laid out as
 * the fused pass takes them (rows = T*B, time-major).
 */
size_t vtr_policy_head_packed_bytes(int32_t K, int32_t A);
int vtr_policy_head_prepare_f32(const float* w_policy, const float* b_policy, const float* w_baseline,
                                const float* b_baseline, int32_t K, int32_t A, void* packed, void* stream);
int vtr_policy_head_f32(const float* core_output, const void* packed, int64_t rows, int32_t K, int32_t A,
                        float* policy_logits, float* baseline, void* stream);

/* x[i] *= *scale for i < n, skipped entirely when *scale == 1.  `scale` is a
   device float[1].  Used to apply the upstream gradient of the total loss
   (experiment.py:415) to the gradients the fused pass already produced,
   without a host round trip to look at its value. */
int vtr_scale_inplace_f32(float* x, const float* scale, int64_t n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VTRACE_B200_H_ */
