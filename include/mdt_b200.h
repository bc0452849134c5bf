/*
 * mdt_b200.h -- C ABI of the B200-native VAE-HPO trial step.
 *
 * The reference (ORNL/MultiDistTorch) has no FFI: its hot path is ~60 lines of
 * Python (vae-hpo.py:19-92,122-131) that call PyTorch.  This header is the
 * boundary a maintainer binds instead of those calls (ctypes stub in
 * INTEGRATION.md).  Every entry point names the reference lines it replaces.
 *
 * Conventions
 *   - plain pointers and sizes only; every `float*` is DEVICE memory unless the
 *     name ends in `_host`; `stream` is a cudaStream_t passed as void*
 *     (NULL = legacy default stream); `nccl_comm` is an ncclComm_t as void*.
 *   - every function returns 0 on success or a negative mdt_status; the message
 *     is available from mdt_last_error() (thread-local).  Nothing throws.
 *   - work is enqueued on the engine's stream and is asynchronous unless the
 *     function name ends in `_host` (those synchronise the stream once).
 *   - one process per GPU, one engine per trial; not re-entrant.
 *   - there is no CPU path: without a CUDA device every call fails with
 *     MDT_ERR_CUDA.
 *
 * Flat parameter layout (all four bound buffers share it), row-major fp32:
 *   W1[H,D] b1[H] W2x[2Z,H] b2x[2Z] W3[H,Z] b3[H] W4[D,H] b4[D]
 * W2x = fc21.weight stacked over fc22.weight (mu rows first), b2x likewise;
 * every segment starts on a 32-float (128-byte) boundary; gaps stay zero.
 * mdt_param_layout() returns the offsets.
 */
#ifndef MDT_B200_H
#define MDT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MDT_VERSION 200 /* 0.2.0 */
#define MDT_NUM_SEGMENTS 8

typedef enum mdt_status {
  MDT_OK = 0,
  MDT_ERR_INVALID = -1,     /* bad argument / shape */
  MDT_ERR_CUDA = -2,        /* CUDA runtime or driver error (message has the code) */
  MDT_ERR_NCCL = -3,        /* NCCL error or NCCL symbols not found in the process */
  MDT_ERR_UNSUPPORTED = -4, /* feature not available in this build / on this device */
  MDT_ERR_STATE = -5        /* call order (e.g. step before bind_buffers) */
} mdt_status;

typedef enum mdt_gemm_mode {
  MDT_GEMM_FP32 = 0,   /* fp32 FFMA tiles: the tight-tolerance parity path */
  MDT_GEMM_TF32 = 1,   /* tcgen05.mma kind::tf32, TMA-staged, TMEM accumulators */
  MDT_GEMM_TF32X3 = 2, /* 3-pass split-TF32 on tcgen05 (fp32-grade accuracy) */
  MDT_GEMM_BF16 = 3    /* tcgen05.mma kind::f16 on bf16 copies of activations and weights (written by the producing
                          epilogues), fp32 accumulate, fp32 master weights and Adam: BASELINE config 4 "bf16 GEMMs" */
} mdt_gemm_mode;

/* How a trial group of n > 1 ranks sums its gradients (the DDP reducer of ref:vae-hpo.py:130). */
typedef enum mdt_exchange {
  MDT_EXCHANGE_AUTO = 0,         /* peer memory when mdt_engine_set_peers was called (one- or two-shot by size), else NCCL */
  MDT_EXCHANGE_NCCL = 1,         /* ncclAllReduce on the subgroup communicator even if peers are mapped */
  MDT_EXCHANGE_P2P_ONE_SHOT = 2, /* every rank reads every peer's whole gradient inside the optimizer kernel */
  MDT_EXCHANGE_P2P_TWO_SHOT = 3  /* reduce own slice, gather the reduced slices inside the optimizer kernel */
} mdt_exchange;

typedef enum mdt_act { MDT_ACT_NONE = 0, MDT_ACT_RELU = 1, MDT_ACT_SIGMOID = 2 } mdt_act;

/* Model/trial configuration: VAE(in_dim - hidden - latent), ref:vae-hpo.py:20-27. */
typedef struct mdt_cfg {
  int32_t in_dim;     /* D, 784 */
  int32_t hidden;     /* H, 400 (4096 in BASELINE config 5) */
  int32_t latent;     /* Z, 20 */
  int32_t max_batch;  /* largest per-rank batch any call will pass */
  int32_t gemm_mode;  /* mdt_gemm_mode */
  int32_t use_graph;  /* 1: capture each (batch size) step into a CUDA graph */
  int32_t exchange;   /* mdt_exchange; 0 = automatic */
  int32_t reserved;   /* 0 */
} mdt_cfg;

/* SGD hyper-parameters as torch.optim.SGD keeps them (BASELINE north_star (c): "fused per-trial Adam/SGD update"). */
typedef struct mdt_sgd {
  double lr, momentum, dampening, weight_decay;
  int32_t nesterov;
  int32_t reserved;
} mdt_sgd;

/* Where mdt_train_steps finds its batches: a device-resident pool of fp32 batches, or a device-resident uint8 image
 * table plus the sampler's index order (ref:vae-hpo.py:135-150: datasets.MNIST + ToTensor + DistributedSampler). */
typedef struct mdt_batch_source {
  const float* x_pool;      /* [n_batches][batch][in_dim] fp32, or NULL to use images_u8 */
  const uint8_t* images_u8; /* [n_images][in_dim] uint8; x = value / 255 as transforms.ToTensor() does */
  const int64_t* indices;   /* [n_batches * batch] device int64 sample indices, batch-major */
  int64_t n_batches;        /* step i of a call reads batch (first_batch + i) % n_batches */
  int64_t n_images;
} mdt_batch_source;

/* Adam hyper-parameters as the Python doubles torch.optim.Adam keeps
 * (ref:vae-hpo.py:131: lr=1e-3, defaults beta=(0.9,0.999), eps=1e-8). */
typedef struct mdt_adam {
  double lr, beta1, beta2, eps;
} mdt_adam;

typedef struct mdt_engine mdt_engine;

int mdt_version(void);
const char* mdt_last_error(void);

/* Offsets/sizes (in floats) of the 8 segments and the padded total.
 * Replaces: the implicit model.parameters() order of ref:vae-hpo.py:23-27,131. */
int mdt_param_layout(const mdt_cfg* cfg, int64_t offsets[MDT_NUM_SEGMENTS], int64_t numels[MDT_NUM_SEGMENTS],
                     int64_t* total);

/* ---- engine: one trial on one GPU ------------------------------------------------ */

/* Replaces: VAE().to(device) + DistributedDataParallel(model, process_group=group)
 * construction, ref:vae-hpo.py:129-130.  `nccl_comm` is the communicator of the
 * trial's subgroup (ref:utils.py:157) or NULL when group_size == 1. */
int mdt_engine_create(const mdt_cfg* cfg, void* nccl_comm, int group_size, int group_rank, void* stream,
                      mdt_engine** out);
void mdt_engine_destroy(mdt_engine* e);

/* The stream the engine enqueues on (its own blocking stream when created with
 * stream == NULL): record timing events / order other work against this one. */
int mdt_engine_stream(mdt_engine* e, void** stream);

/* Caller-owned flat buffers (layout above), 128-byte aligned, `numel` == layout total.
 * Replaces: model.parameters() / p.grad / Adam state tensors, ref:vae-hpo.py:131. */
int mdt_engine_bind_buffers(mdt_engine* e, float* params, float* grads, float* exp_avg, float* exp_avg_sq,
                            size_t numel);

/* DDP constructor semantics: broadcast params from group-rank 0 (ref:vae-hpo.py:130). */
int mdt_engine_broadcast_params(mdt_engine* e);

/* Device-resident per-trial state read by the kernels (so a captured step needs no
 * host-side scalars): Adam hyper-parameters + completed step count, Philox seed/offset.
 * Replaces: optim.Adam(..., lr=1e-3) state and the CUDA default generator. */
int mdt_engine_set_adam(mdt_engine* e, const mdt_adam* h, int64_t completed_steps);
int mdt_engine_set_rng(mdt_engine* e, uint64_t philox_seed, uint64_t philox_offset);
/* optim.SGD(model.parameters(), lr, momentum, dampening, weight_decay, nesterov) + .step() in place of Adam
 * (north_star (c); the reference itself only instantiates Adam, ref:vae-hpo.py:131).  exp_avg holds the momentum
 * buffer, exp_avg_sq is unused.  The update runs as one flat kernel after the gradient exchange (NCCL for groups). */
int mdt_engine_set_sgd(mdt_engine* e, const mdt_sgd* h, int64_t completed_steps);
int mdt_engine_get_state(mdt_engine* e, int64_t* completed_steps, uint64_t* philox_offset); /* syncs */

/* Stage buffer the step reads x from ([max_batch, D] fp32).  Writing the batch here
 * (H2D copy, or a device-side sampler) skips the D2D copy mdt_train_step would do. */
int mdt_engine_input_buffer(mdt_engine* e, float** x_stage);

/* recon, mu, logvar = model(x): ref:vae-hpo.py:42-45 (encode 29-31, reparameterize
 * 33-36, decode 38-40).  eps_or_null == NULL draws eps from Philox exactly like
 * torch.randn_like on CUDA and advances the offset.  Outputs may be NULL.  Keeps
 * the activations for mdt_backward. */
int mdt_forward(mdt_engine* e, const float* x, int batch, const float* eps_or_null, float* recon, float* mu,
                float* logvar);

/* loss.backward() for upstream gradients produced elsewhere (compat mode, where the
 * script's own loss_function ran in torch): ref:vae-hpo.py:72.  Fills `grads`. */
int mdt_backward(mdt_engine* e, const float* d_recon, const float* d_mu, const float* d_logvar);

/* Gradient averaging over the trial group: the DDP reducer inside
 * loss.backward(), ref:vae-hpo.py:72,130.  SUM only; 1/n is applied by the Adam
 * kernel (mdt_adam_step grad_scale). */
int mdt_allreduce_grads(mdt_engine* e);

/* optimizer.step(): ref:vae-hpo.py:74.  Uses/advances the device-resident step count. */
int mdt_adam_step(mdt_engine* e, float grad_scale);

/* One whole training step, ref:vae-hpo.py:68-74:
 *   forward -> BCE+KLD loss -> backward -> group all-reduce -> Adam.
 * `loss_out` (device, may be NULL) receives the summed loss of this rank's batch. */
int mdt_train_step(mdt_engine* e, const float* x, int batch, const float* eps_or_null, float* loss_out);

/* The inner loop of train(), ref:vae-hpo.py:67-74, for `n_steps` consecutive batches of a device-resident source:
 * step i trains on batch (first_batch + i) % src->n_batches.  The steps are enqueued as CUDA graphs of several steps
 * each (the batch index lives in device memory and advances by itself, so the graphs are launched unmodified); with
 * images_u8 the gather + /255 of mdt_gather_u8 happens inside each step's first kernel.  `losses_out` (device,
 * n_steps floats, or NULL; n_steps <= 4096 when given) receives the per-step losses; the host mailbox of
 * mdt_train_step_host_wait is fed as well.  Asynchronous. */
int mdt_train_steps(mdt_engine* e, const mdt_batch_source* src, int batch, int64_t first_batch, int n_steps,
                    float* losses_out);

/* SURVEY section 8f-3 (several independent trials on one GPU): one step of each of `n_trials` engines, every engine on
 * its own stream / CUDA graph so their kernels overlap on idle SMs.  xs[i] is trial i's device batch; eps comes from
 * each engine's Philox state; losses stay in the engines (mdt_engine_workspace "loss").  Equivalent to calling
 * mdt_train_step on each engine; it exists to take the per-call host overhead out of the HPO driver's loop.
 * Replaces: nothing in the reference (it runs one trial per rank, ref:vae-hpo.py:200-202). */
int mdt_train_step_many(mdt_engine* const* engines, const float* const* xs, int n_trials, int batch);
/* The same for whole training loops, with ONE launch per layer for all trials (SURVEY section 8f-3 as specified): a
 * group of T single-rank Adam trials of identical shape on one GPU is stepped in lock-step; every kernel of the step
 * runs once over a table of the T trials' tensor maps and arguments (grid.y resp. grid.x = trial x tiles), inside CUDA
 * graphs of 16 steps on trial 0's streams.  srcs[t] is trial t's device-resident batch source (all with the same
 * n_batches); step i of the call trains every trial on its batch (first_batch + i) % n_batches.  Each trial keeps its
 * own parameters, optimizer state, Philox stream, loss ring and mailbox; results are bit-identical to stepping the
 * trial alone with mdt_engine_set_sm_share(T).  Small-model regime only (hidden < 2048, max_batch <= 512).
 * Replaces: the per-group loop of ref:vae-hpo.py:200-202 for trials that share a GPU. */
typedef struct mdt_group mdt_group;
int mdt_group_create(mdt_engine* const* engines, int n_trials, mdt_group** out);
int mdt_group_train_steps(mdt_group* g, const mdt_batch_source* srcs, int batch, int64_t first_batch, int n_steps);
void mdt_group_destroy(mdt_group* g);

/* (new, no reference counterpart) Size this engine's kernels -- K-split counts, CTAs per launch -- for 1/n_sharing of
 * the GPU's SMs, for callers that run several trials on one GPU themselves (own threads / streams).  Default 1.
 * mdt_train_step_many applies max(n_trials, this value) for the duration of its call. */
int mdt_engine_set_sm_share(mdt_engine* e, int n_sharing);

/* Same, with HOST buffers: copies x (and eps) host->device, runs the step, copies
 * the loss back and synchronises (the `loss.item()` of ref:vae-hpo.py:73).
 * Pinned host memory makes the copies asynchronous; pageable memory also works. */
int mdt_train_step_host(mdt_engine* e, const float* x_host, int batch, const float* eps_host_or_null,
                        float* loss_host);

/* The same step in two halves, so that the host can keep the next step's input in flight while it waits for this
 * step's loss (what a prefetching DataLoader + a deferred loss.item() do for ref:vae-hpo.py:64-73):
 * mdt_train_step_host_submit enqueues the step and returns a ticket; mdt_train_step_host_wait returns that step's
 * loss.  x_host (and eps_host) must stay untouched until the matching wait has returned: when they are pinned the
 * device reads them in place.  Up to 63 steps may be in flight; tickets may be waited on in any order. */
int mdt_train_step_host_submit(mdt_engine* e, const float* x_host, int batch, const float* eps_host_or_null,
                               uint32_t* ticket);
int mdt_train_step_host_wait(mdt_engine* e, uint32_t ticket, float* loss_host);

/* A chunk of 1..16 consecutive host batches in one call: x_host is [n_steps][batch][in_dim] (pinned memory makes the
 * copy asynchronous).  The chunk is DMA'd into one half of a device ring on the copy stream -- behind the chunk that is
 * still training -- and its steps run as one CUDA graph; every step publishes its loss to the mailbox as usual.
 * *first_ticket is the ticket of the chunk's first step, the others follow consecutively (wait for each with
 * mdt_train_step_host_wait).  What a DataLoader with a prefetch depth of one chunk gives ref:vae-hpo.py:64-73.
 * x_host must stay untouched until the chunk's first ticket has been waited for.  At most two chunks in flight.
 * The first call allocates the ring (2 x 16 x max_batch x in_dim floats). */
int mdt_train_steps_host_submit(mdt_engine* e, const float* x_host, int batch, int n_steps, uint32_t* first_ticket);

/* Forward + loss only (the eval loop), ref:vae-hpo.py:101-105. */
int mdt_eval_step(mdt_engine* e, const float* x, int batch, const float* eps_or_null, float* loss_out,
                  float* recon);

/* model.module.decode(z): ref:vae-hpo.py:38-40,165. */
int mdt_decode(mdt_engine* e, const float* z, int batch, float* recon);

/* Debug/test access to an internal activation buffer by name ("h1","ml","z","h3",
 * "recon","da4","da3","dml","da1","eps"); returns its device pointer and leading
 * dimension. */
int mdt_engine_workspace(mdt_engine* e, const char* name, float** ptr, int* ld);

/* Number of kernels this engine has enqueued so far (a CUDA-graph launch counts the
 * kernels inside the graph).  Measurement aid; no reference counterpart. */
int mdt_engine_launch_count(mdt_engine* e, int64_t* kernels);

/* Gradient exchange over NVLink peer memory for a group whose ranks sit on one node (replaces the ncclAllReduce +
 * optimizer read of ref:vae-hpo.py:72-74 / torch DDP's reducer; mdt_allreduce_grads stays the NCCL form).
 * mdt_engine_exchange_handle: allocates this rank's exchange region on first use and returns its 64-byte CUDA IPC
 * handle.  The host side gathers the handles of the group (rank order) and passes all n of them to
 * mdt_engine_set_peers, which maps the peers' regions.  From then on mdt_train_step / mdt_train_step_host write the
 * gradient into the exchange region, and one kernel per rank waits for its peers' "ready" flags, sums the n partial
 * gradients in rank order (own from local memory, the others over NVLink), stores the sum into the bound gradient
 * buffer and applies Adam.  Errors: MDT_ERR_CUDA when a peer's region cannot be mapped (no peer access) -- the caller
 * then simply keeps the NCCL path.
 * One engine per GPU at a time on this path: the exchange kernel keeps all its blocks resident while it waits for the
 * peers, which leaves no room for another engine's GEMMs on the same GPU.  Two engines whose exchanges are in flight
 * together (in different orders on different ranks) therefore stall each other until the bounded wait expires; the
 * reference never does this (one trial per group at a time, ref:vae-hpo.py:200-202) -- synchronise one engine before
 * stepping the next, or give concurrent engines MDT_EXCHANGE_NCCL. */
int mdt_engine_exchange_handle(mdt_engine* e, void* handle64);
int mdt_engine_set_peers(mdt_engine* e, const void* handles64, int n);
/* *enabled: 1 once mdt_engine_set_peers succeeded; *timed_out: 1 if a bounded wait on a peer's flag ever expired (a
 * peer died or never launched its step).  The step in which that happens is skipped (parameters and moments
 * untouched, NaN loss) and every later mdt_train_step* / mdt_forward call on the engine returns MDT_ERR_STATE: the
 * kernels raise a host-visible flag, so no synchronisation is needed to notice.  This call synchronises. */
int mdt_engine_exchange_status(mdt_engine* e, int* enabled, int* timed_out);
/* Back to the NCCL exchange (e.g. when another rank of the group could not map its peers: all ranks must agree). */
int mdt_engine_disable_peers(mdt_engine* e);
/* Measurement aid for BASELINE's "subgroup allreduce bus GB/s" on the repo's own exchange: runs the group's gradient
 * exchange + optimizer kernel(s) `iters` times on the flat gradient (every rank of the group must call it together,
 * peers mapped) and returns the average seconds per exchange.  The parameters ARE updated (use a scratch engine). */
int mdt_engine_exchange_bench(mdt_engine* e, int iters, double* seconds);

/* Measurement aid: run `iters` un-graphed training steps on x with a CUDA event after
 * every kernel and write "label avg_ms\n" lines (one per kernel of the step, in launch
 * order) into `out`.  The steps DO update the parameters.  No reference counterpart
 * (the reference only prints wall time, ref:vae-hpo.py:159,173-174). */
int mdt_engine_profile_step(mdt_engine* e, const float* x, int batch, int iters, char* out, size_t out_len);

/* ---- stand-alone kernels (also used by the engine) --------------------------------- */

/* Fused flat Adam: one pass over p,g,m,v (7 x 4 bytes per element).  `step` is the
 * 1-based step being taken.  Replaces torch's _multi_tensor_adam, ref:vae-hpo.py:74. */
int mdt_adam_flat(float* p, const float* g, float* m, float* v, size_t n, const mdt_adam* h, int64_t step,
                  float grad_scale, void* stream);

/* z = mu + eps*exp(0.5*logvar), ref:vae-hpo.py:33-36.  mu/logvar/z have leading
 * dimensions ld_ml / ld_z.  eps_in NULL -> Philox(seed, offset) as torch CUDA randn;
 * eps_out (optional) receives the eps used. */
int mdt_reparam(const float* mu, const float* logvar, int ld_ml, const float* eps_in, uint64_t seed,
                uint64_t offset, int batch, int latent, float* z, int ld_z, float* eps_out, void* stream);

/* Fused BCE(sum)+KLD loss and its gradients, ref:vae-hpo.py:49-58 and its autograd.
 * d_* may be NULL.  `loss_out` is one device float.  Uses a small internally
 * allocated scratch for per-block partial sums (deterministic summation order). */
int mdt_loss_bce_kl(const float* recon, const float* x, const float* mu, const float* logvar, int batch,
                    int in_dim, int latent, float* loss_out, float* d_recon, float* d_mu, float* d_logvar,
                    void* stream);

/* Same computation and arguments as mdt_loss_bce_kl, for callers that must not trigger a
 * device allocation: code that runs while a memory budget is already fully committed (a
 * pre-sized pool owns the whole GPU), latency-critical loops where the implicit device
 * synchronisation of cudaMalloc is unacceptable, and set-ups whose allocator cannot be
 * re-entered from the calling context.  It has no counterpart in the reference.
 * The per-block partial sums and the completion ticket live in a fixed 4352-byte array that
 * is part of the library image (every GPU that loads the library holds it until exit), so the
 * call performs no cudaMalloc/cudaFree.  Limit: batch*in_dim <= 1,048,576 elements (1024
 * blocks); larger inputs are rejected with MDT_ERR_INVALID and a message.  The first call on
 * each device also looks up the array's device address (cudaGetSymbolAddress); make that
 * first call outside any stream capture.  All calls on a device share the one array: the
 * ticket is re-zeroed on `stream` before every launch, but two calls that overlap on the GPU
 * (different streams without ordering) race on it -- serialise them. */
int mdt_loss_bce_kl_noalloc(const float* recon, const float* x, const float* mu, const float* logvar, int batch,
                            int in_dim, int latent, float* loss_out, float* d_recon, float* d_mu, float* d_logvar,
                            void* stream);

/* SURVEY section 8f-1 (device-resident data path): x_out[b, :] = images[indices[b], :] / 255 as fp32 -- the values
 * datasets.MNIST + transforms.ToTensor() + the DataLoader collate hand to the step (ref:vae-hpo.py:135-150,67-68),
 * gathered on the device (e.g. straight into mdt_engine_input_buffer) from a uint8 image table that stays in HBM.
 * `indices` are device int64 sample indices (the DistributedSampler order); in_dim % 16 == 0. */
int mdt_gather_u8(const uint8_t* images, int64_t n_images, const int64_t* indices, int batch, int in_dim, float* x_out,
                  void* stream);

/* y[M,N] = act(x[M,K] . W[N,K]^T + b[N]) : nn.Linear (+F.relu / torch.sigmoid),
 * ref:vae-hpo.py:30-31,39-40.  `gemm_mode` selects the kernel family. */
int mdt_linear(const float* x, const float* w, const float* b, int M, int N, int K, int act, int gemm_mode,
               float* y, void* stream);

/* Time `iters` in-place fp32 SUM all-reduces of `bytes` bytes on `comm`; returns the
 * average seconds per all-reduce measured with CUDA events on `stream`.  Replaces
 * nothing in the reference (it only all-gathers one int64, ref:example-subgroup.py:25-33);
 * it backs BASELINE.json's "subgroup allreduce bus GB/s" metric. */
int mdt_allreduce_bench(void* nccl_comm, void* buf, size_t bytes, int iters, void* stream, double* seconds);

#ifdef __cplusplus
}
#endif
#endif /* MDT_B200_H */
