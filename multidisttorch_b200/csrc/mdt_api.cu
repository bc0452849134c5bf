// C-ABI of the B200-native VAE-HPO trial step (see include/mdt_b200.h).
// Host side: the per-trial engine (workspaces, device-resident step state, CUDA
// graph cache, NCCL subgroup all-reduce) and the stand-alone kernel entry points.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <cstdarg>
#include <map>
#include <tuple>
#include <vector>

#include "common.cuh"
#include "elementwise.cuh"
#include "gemm_simt.cuh"
#include "gemm_tcgen05.cuh"

using namespace mdt;

// ------------------------------------------------------------------ NCCL (resolved at run time)
// The library never links libnccl: it binds the instance torch already mapped
// (nvidia/nccl/lib/libnccl.so.2) so the communicator handed over by
// ProcessGroupNCCL and the functions called on it come from the same build.
namespace {
struct NcclApi {
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool tried = false, ok = false;
};
NcclApi g_nccl;

int load_nccl() {
  if (g_nccl.ok) return MDT_OK;
  if (!g_nccl.tried) {
    g_nccl.tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) {
      const char* p = getenv("MDT_NCCL_LIB");
      if (p) h = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!h) h = RTLD_DEFAULT;
    g_nccl.AllReduce = reinterpret_cast<decltype(g_nccl.AllReduce)>(dlsym(h, "ncclAllReduce"));
    g_nccl.Broadcast = reinterpret_cast<decltype(g_nccl.Broadcast)>(dlsym(h, "ncclBroadcast"));
    g_nccl.GetErrorString = reinterpret_cast<decltype(g_nccl.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    g_nccl.ok = g_nccl.AllReduce && g_nccl.Broadcast && g_nccl.GetErrorString;
  }
  if (!g_nccl.ok)
    return set_error(MDT_ERR_NCCL,
                     "NCCL symbols not found in this process: import torch (CUDA build) first, or set MDT_NCCL_LIB");
  return MDT_OK;
}

#define MDT_NCCL(expr)                                                                                   \
  do {                                                                                                   \
    ncclResult_t _r = (expr);                                                                            \
    if (_r != ncclSuccess)                                                                               \
      return set_error(MDT_ERR_NCCL, "%s failed: %s (%s:%d)", #expr, g_nccl.GetErrorString(_r), __FILE__, \
                       __LINE__);                                                                        \
  } while (0)

// ------------------------------------------------------------------ per-device info + lazy scratch
struct DeviceInfo {
  bool init = false;
  int sm_count = 0, max_threads_sm = 0, cc_major = 0, cc_minor = 0;
  float* loss_partials = nullptr;  // scratch of the stand-alone loss entry
  unsigned int* loss_ticket = nullptr;
};
DeviceInfo g_dev[64];

int device_info(DeviceInfo** out) {
  int dev = 0;
  MDT_CUDA(cudaGetDevice(&dev));
  MDT_CHECK(dev >= 0 && dev < 64, "device index %d out of range", dev);
  DeviceInfo& d = g_dev[dev];
  if (!d.init) {
    MDT_CUDA(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev));
    MDT_CUDA(cudaDeviceGetAttribute(&d.max_threads_sm, cudaDevAttrMaxThreadsPerMultiProcessor, dev));
    MDT_CUDA(cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    MDT_CUDA(cudaDeviceGetAttribute(&d.cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
    d.init = true;
  }
  *out = &d;
  return MDT_OK;
}

size_t align_up(size_t n, size_t a) { return (n + a - 1) / a * a; }
}  // namespace

// ------------------------------------------------------------------ engine
struct mdt_engine {
  mdt_cfg cfg;
  int D, H, Z, Bmax;
  Layout L;
  DeviceInfo* dev = nullptr;
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  ncclComm_t comm = nullptr;
  int gsize = 1, grank = 0;

  float *params = nullptr, *grads = nullptr, *m = nullptr, *v = nullptr;
  bool bound = false;

  char* slab = nullptr;
  float *x_stage, *eps_stage, *h1, *ml, *eps, *sd, *z, *h3, *recon, *da4, *da3, *dz, *dml_kl, *dml, *da1;
  float *loss_partials, *loss_dev;
  unsigned int* loss_ticket;
  StepState* st;
  float* loss_pinned = nullptr;
  // peer-memory gradient exchange (groups of n > 1 on one node): see adam_allreduce_kernel
  float* xch = nullptr;                 // own exchange region (cudaMalloc): L.total floats of gradient + 64 flag words
  void* xch_opened[kMaxPeers] = {};     // peers' regions mapped with cudaIpcOpenMemHandle
  PeerArgs peers{};
  bool p2p = false;
  int sm_share_pending = 1;
  // diagnostic switches, read from the environment once when the engine is created
  bool opt_fused_adam = true, opt_nccl_in_graph = true, opt_zerocopy = true, opt_h2d_prefetch = true;
  bool opt_pipeline = true, opt_graph_reduce = true, opt_graph_dbg = false;
  int opt_steps_per_graph = 16;
  bool opt_p2p_single = false;                      // MDT_P2P_SINGLE_STEPS
  bool opt_defer_shadows = true;                    // MDT_NO_DEFER_SHADOWS=1: weight shadows refreshed on the main stream
  volatile unsigned long long* mailbox = nullptr;   // ring of 64 words at loss_pinned + 64 bytes: see publish_loss()
  volatile unsigned int* err_word = nullptr;        // pinned word the exchange kernels set when a peer never answered
  float* loss_hist = nullptr;                       // device ring of the last kLossHist per-step losses
  long long cursor_expected = -1;                   // host's view of StepState.x_cursor (-1: unknown, set it)
  const void* src_sig[4] = {nullptr, nullptr, nullptr, nullptr};   // batch source the cached multi-step graphs were captured for
  bool sourced = false;                             // inside mdt_train_steps
  bool primed = false;                              // the current input set already holds batch `primed_cursor` of the source
  long long primed_cursor = -1;
  bool use_sgd = false;                             // mdt_engine_set_sgd: the flat SGD kernel replaces every Adam form
  SgdScalars sgd{};
  int exchange = 0;                                 // mdt_exchange from the cfg (0 = automatic choice)
  long long two_shot_bytes = 16ll << 20;            // automatic choice: two-shot once a rank would read more than this
  int two_shot_from = 3;                            // ... and the group has at least this many ranks
  // host-buffer steps submitted while the GPU is still busy: x is DMA'd into a small ring of device buffers on a copy
  // stream (overlapping the step in flight) instead of being read over PCIe by the step's own first kernel
  static constexpr int kPreSlots = 4;
  cudaStream_t copy_stream = nullptr;
  float* x_pre[kPreSlots] = {};
  cudaEvent_t ev_copied[kPreSlots] = {}, ev_used[kPreSlots] = {};
  bool pre_used[kPreSlots] = {};
  unsigned int pre_next = 0;
  // host-buffer steps submitted while a step is in flight: the prologue of the new step (pinned x -> input set) runs
  // on the copy stream, alternating between the two input sets of the tensor-core plan
  // mdt_train_steps_host_submit: chunks of up to kHostChunk host batches land in the two halves of a device ring
  static constexpr int kHostChunk = 16;
  float* x_ring = nullptr;
  cudaEvent_t ev_ring_copied[2] = {}, ev_ring_free[2] = {};
  bool ring_used[2] = {};
  unsigned int ring_next = 0;
  bool tail_no_prefetch = false;                    // the last step of the current mdt_train_steps call fetches nothing
  cudaEvent_t ev_set_ready[2] = {}, ev_set_free[2] = {};
  bool set_used[2] = {};
  int host_set = 0;
  float sync_losses[64] = {};
  unsigned int sync_seq = 0;                       // losses of submitted steps on configurations without the mailbox
  unsigned int pub_expected = 0;                    // training steps enqueued so far that publish to the mailbox
  int last_B = 0;  // batch of the activations currently held

  Tc05Plan* tc = nullptr;  // tcgen05 path state (tensor maps, transposed shadows); NULL in fp32 mode

  // gradient all-reduce of a multi-rank trial group: its own stream, forked from / joined to the step by events so
  // bucket 0 (fc4, ready first -- DDP's order) travels while the rest of backward still runs
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_bucket[2] = {nullptr, nullptr};
  cudaEvent_t ev_comm_done = nullptr;
  bool bucket0_sent = false;
  bool defer_buckets = false;   // compat-mode backward: the caller all-reduces later (mdt_allreduce_grads)
  bool one_bucket = false;      // small flat gradient: ONE all-reduce after backward instead of DDP's two buckets

  struct Graph {
    cudaGraphExec_t exec; long long kernels;
    // tensor-core path: the graph's prologue kernel node, re-pointed at the caller's x before a launch
    cudaGraph_t graph = nullptr; cudaGraphNode_t prep_node = nullptr; PrepArgs prep{}; dim3 prep_grid{};
  };
  std::map<std::tuple<int, int, int>, Graph> graphs;

  long long launch_count = 0;  // kernels enqueued (a graph launch adds the kernels it holds)
  bool prof_on = false;        // per-kernel event timing (mdt_engine_profile_step)
  std::vector<std::pair<const char*, cudaEvent_t>> prof_marks;

  float* seg(float* base, int s) const { return base + L.off[s]; }
};

namespace {

// Book-keeping after every kernel launch: count it and, in profiling mode, drop a timing event.
int note_launch(mdt_engine* e, const char* label) {
  MDT_CUDA(cudaGetLastError());
  e->launch_count += 1;
  if (e->prof_on) {
    cudaEvent_t ev;
    MDT_CUDA(cudaEventCreate(&ev));
    MDT_CUDA(cudaEventRecord(ev, e->stream));
    e->prof_marks.emplace_back(label, ev);
  }
  return MDT_OK;
}

__global__ void null_kernel() {}

__global__ void spin_kernel(long long ns) {
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do {
    __nanosleep(1000);
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  } while ((long long)(t1 - t0) < ns);
}

int note_cb(void* ctx, const char* label) { return note_launch(static_cast<mdt_engine*>(ctx), label); }

int carve_workspace(mdt_engine* e) {
  const size_t B = e->Bmax, D = e->D, H = e->H, Z = e->Z;
  struct Item { float** p; size_t n; };
  std::vector<Item> items = {
      {&e->x_stage, B * D}, {&e->eps_stage, B * Z}, {&e->h1, B * H}, {&e->ml, B * 2 * Z}, {&e->eps, B * Z},
      {&e->sd, B * Z}, {&e->z, B * Z}, {&e->h3, B * H}, {&e->recon, B * D}, {&e->da4, B * D}, {&e->da3, B * H},
      {&e->dz, B * Z}, {&e->dml_kl, B * 2 * Z}, {&e->dml, B * 2 * Z}, {&e->da1, B * H},
      {&e->loss_partials, (size_t)kLossMaxBlocks}, {&e->loss_dev, 64}, {&e->loss_hist, (size_t)kLossHist},
  };
  size_t bytes = 0;
  for (auto& it : items) bytes += align_up(it.n * sizeof(float), 256);
  bytes += 256 /*ticket*/ + align_up(sizeof(StepState), 256);
  MDT_CUDA(cudaMalloc(&e->slab, bytes));
  MDT_CUDA(cudaMemsetAsync(e->slab, 0, bytes, e->stream));
  char* p = e->slab;
  for (auto& it : items) {
    *it.p = reinterpret_cast<float*>(p);
    p += align_up(it.n * sizeof(float), 256);
  }
  e->loss_ticket = reinterpret_cast<unsigned int*>(p);
  p += 256;
  e->st = reinterpret_cast<StepState*>(p);
  return MDT_OK;
}

// ---- enqueue helpers (no synchronisation; safe under stream capture) --------------------------
// GEMM launch + book-keeping in one statement.
#define MDT_GEMM_SIMT(label, AK, BK, ...)                      \
  do {                                                         \
    MDT_CUDA((launch_gemm_simt<AK, BK>(__VA_ARGS__, s)));      \
    MDT_TRY(note_launch(e, label));                            \
  } while (0)

int enqueue_forward_fp32(mdt_engine* e, const float* x, int B, const float* eps_in, bool decode_only) {
  cudaStream_t s = e->stream;
  const int D = e->D, H = e->H, Z = e->Z;
  const float* P = e->params;
  if (!decode_only) {
    MDT_GEMM_SIMT("fc1_fwd", true, true, x, P + e->L.off[W1], B, H, D, D, D,
                  EpiBiasAct{e->h1, H, P + e->L.off[B1], MDT_ACT_RELU});
    MDT_GEMM_SIMT("fc2x_fwd", true, true, e->h1, P + e->L.off[W2X], B, 2 * Z, H, H, H,
                  EpiBiasAct{e->ml, 2 * Z, P + e->L.off[B2X], MDT_ACT_NONE});
    ReparamArgs a{};
    a.mu = e->ml; a.logvar = e->ml + Z; a.ld_ml = 2 * Z; a.eps_in = eps_in; a.st = e->st;
    torch_randn_policy((long long)B * Z, e->dev->sm_count, e->dev->max_threads_sm, &a.S, &a.offset_inc);
    a.B = B; a.Z = Z; a.z = e->z; a.ld_z = Z; a.eps_out = e->eps; a.std_out = e->sd;
    const int blocks = (int)(((long long)B * Z + 255) / 256);
    reparam_kernel<<<blocks < 1 ? 1 : blocks, 256, 0, s>>>(a);
    MDT_TRY(note_launch(e, "reparam"));
  }
  MDT_GEMM_SIMT("fc3_fwd", true, true, e->z, P + e->L.off[W3], B, H, Z, Z, Z,
                EpiBiasAct{e->h3, H, P + e->L.off[B3], MDT_ACT_RELU});
  MDT_GEMM_SIMT("fc4_fwd", true, true, e->h3, P + e->L.off[W4], B, D, H, H, H,
                EpiBiasAct{e->recon, D, P + e->L.off[B4], MDT_ACT_SIGMOID});
  return MDT_OK;
}

int enqueue_loss(mdt_engine* e, const float* x, int B, bool with_grads) {
  LossArgs a{};
  a.recon = e->recon; a.x = x; a.mu = e->ml; a.logvar = e->ml + e->Z; a.ld_ml = 2 * e->Z;
  a.B = B; a.D = e->D; a.Z = e->Z;
  if (with_grads) { a.d_recon = e->da4; a.d_mu = e->dml_kl; a.d_logvar = e->dml_kl + e->Z; a.ld_dml = 2 * e->Z; }
  a.partials = e->loss_partials; a.ticket = e->loss_ticket; a.loss_out = e->loss_dev;
  loss_bce_kl_kernel<true><<<loss_grid((long long)B * e->D), kLossThreads, 0, e->stream>>>(a);
  return note_launch(e, "loss_bce_kl");
}

// Backward from da4 (= dL/d logits of fc4) and dml_kl (= direct dL/d[mu|logvar]).
int enqueue_backward_fp32(mdt_engine* e, const float* x, int B) {
  cudaStream_t s = e->stream;
  const int D = e->D, H = e->H, Z = e->Z;
  const float* P = e->params;
  float* G = e->grads;
  // fc4
  MDT_GEMM_SIMT("fc4_dW", false, false, e->da4, e->h3, D, H, B, D, H, EpiStore{G + e->L.off[W4], H});
  MDT_GEMM_SIMT("fc4_dX", true, false, e->da4, P + e->L.off[W4], B, H, D, D, H, EpiReluMask{e->da3, H, e->h3, H});
  // fc3
  MDT_GEMM_SIMT("fc3_dW", false, false, e->da3, e->z, H, Z, B, H, Z, EpiStore{G + e->L.off[W3], Z});
  MDT_GEMM_SIMT("fc3_dX", true, false, e->da3, P + e->L.off[W3], B, Z, H, H, Z, EpiStore{e->dz, Z});
  {
    const int blocks = (int)(((long long)B * Z + 255) / 256);
    reparam_bwd_kernel<<<blocks < 1 ? 1 : blocks, 256, 0, s>>>(e->dz, Z, e->eps, e->sd, e->dml_kl, B, Z, e->dml);
    MDT_TRY(note_launch(e, "reparam_bwd"));
  }
  // fc21 | fc22
  MDT_GEMM_SIMT("fc2x_dW", false, false, e->dml, e->h1, 2 * Z, H, B, 2 * Z, H, EpiStore{G + e->L.off[W2X], H});
  MDT_GEMM_SIMT("fc2x_dX", true, false, e->dml, P + e->L.off[W2X], B, H, 2 * Z, 2 * Z, H,
                EpiReluMask{e->da1, H, e->h1, H});
  // fc1 (no dX)
  MDT_GEMM_SIMT("fc1_dW", false, false, e->da1, x, H, D, B, H, D, EpiStore{G + e->L.off[W1], D});
  // bias gradients, one launch
  ColsumArgs c{};
  c.B = B; c.njobs = 4;
  const float* dy[4] = {e->da4, e->da3, e->dml, e->da1};
  const int nn[4] = {D, H, 2 * Z, H};
  const int sg[4] = {B4, B3, B2X, B1};
  int blk = 0;
  for (int i = 0; i < 4; ++i) {
    c.job[i] = ColsumJob{dy[i], nn[i], nn[i], G + e->L.off[sg[i]], blk};
    blk += (nn[i] + 31) / 32;
  }
  colsum_kernel<<<blk, 256, 0, s>>>(c);
  return note_launch(e, "bias_grads");
}

int enqueue_adam(mdt_engine* e, float grad_scale, bool finalize_loss = false) {
  AdamArgs a{};
  a.p = e->params; a.g = e->grads; a.m = e->m; a.v = e->v; a.n = e->L.total; a.st = e->st; a.grad_scale = grad_scale;
  if (finalize_loss && e->tc) {   // the fused forward of the tcgen05 path leaves per-tile loss partials
    tc05_loss_partials(e->tc, e->last_B, &a.loss_part_a, &a.n_part_a, &a.loss_part_b, &a.n_part_b);
    a.loss_out = e->loss_dev;
  }
  if (e->use_sgd) {
    MDT_CUDA(launch_pdl(sgd_flat_kernel, dim3(adam_grid(a.n, e->dev->sm_count)), dim3(kAdamThreads), 0, e->stream, a, e->sgd));
    MDT_TRY(note_launch(e, "sgd_flat"));
  } else {
    MDT_CUDA(launch_pdl(adam_flat_kernel, dim3(adam_grid(a.n, e->dev->sm_count)), dim3(kAdamThreads), 0, e->stream, a));
    MDT_TRY(note_launch(e, "adam_flat"));
  }
  if (e->tc) MDT_TRY(tc05_params_changed(e->tc, e->stream));   // refresh the transposed weight shadows
  return MDT_OK;
}

// Gradient exchange over peer memory + Adam in one kernel (replaces finish_allreduce + enqueue_adam for the step).
int enqueue_adam_p2p(mdt_engine* e, float grad_scale) {
  AdamArgs a{};
  a.p = e->params; a.g = e->grads; a.m = e->m; a.v = e->v; a.n = e->L.total; a.st = e->st; a.grad_scale = grad_scale;
  tc05_loss_partials(e->tc, e->last_B, &a.loss_part_a, &a.n_part_a, &a.loss_part_b, &a.n_part_b);
  a.loss_out = e->loss_dev;
  PeerArgs pa = e->peers;
  pa.g_out = e->grads;
  // One-shot (each rank reads every peer's whole gradient: (n-1) x M bytes over NVLink) while that is small; two-shot
  // (reduce own slice, then gather the reduced slices inside the Adam pass: 2(n-1)/n x M) for 3+ ranks with more than
  // 16 MB to read.  Measured on 4 ranks: M = 2.6 MB one-shot 5.57 vs two-shot 5.27 M samples/s; M = 26.7 MB one-shot
  // 1.66 vs two-shot 1.85 M (NCCL 1.83 M).
  const long long peer_bytes = (long long)(pa.n - 1) * a.n * (long long)sizeof(float);
  const bool two_shot = e->exchange == MDT_EXCHANGE_P2P_TWO_SHOT ||
                        (e->exchange != MDT_EXCHANGE_P2P_ONE_SHOT && pa.n >= e->two_shot_from && peer_bytes > e->two_shot_bytes);
  if (two_shot) {
    const dim3 grid(adam_grid(a.n, e->dev->sm_count));
    MDT_CUDA(launch_pdl(p2p_reduce_slice_kernel, dim3(grid.x / (unsigned)pa.n + 1), dim3(kAdamThreads), 0, e->stream, pa,
                        (long long)(a.n >> 2), e->st));
    MDT_TRY(note_launch(e, "p2p_reduce_slice"));
    MDT_CUDA(launch_pdl(adam_allgather_kernel, grid, dim3(kAdamThreads), 0, e->stream, a, pa));
    MDT_TRY(note_launch(e, "adam_allgather_p2p"));
  } else {
    MDT_CUDA(launch_pdl(adam_allreduce_kernel, dim3(adam_grid(a.n, e->dev->sm_count)), dim3(kAdamThreads), 0, e->stream, a, pa));
    MDT_TRY(note_launch(e, "adam_allreduce_p2p"));
  }
  return tc05_params_changed(e->tc, e->stream);   // refresh the transposed weight shadows
}

int enqueue_forward(mdt_engine* e, const float* x, int B, const float* eps_in, bool decode_only) {
  if (e->tc) return tc05_forward(e->tc, e->stream, x, B, eps_in, decode_only, /*fuse_loss=*/false);
  return enqueue_forward_fp32(e, x, B, eps_in, decode_only);
}
int enqueue_backward(mdt_engine* e, const float* x, int B) {
  if (e->tc) return tc05_backward(e->tc, e->stream, x, B);
  return enqueue_backward_fp32(e, x, B);
}

int check_batch(mdt_engine* e, int B) {
  MDT_CHECK(e != nullptr, "engine is NULL");
  if (!e->bound) return set_error(MDT_ERR_STATE, "mdt_engine_bind_buffers has not been called");
  MDT_CHECK(B >= 0 && B <= e->Bmax, "batch %d outside [0, max_batch=%d]", B, e->Bmax);
  if (e->err_word != nullptr && *e->err_word != 0u)
    return set_error(MDT_ERR_STATE, "a peer of this trial group never delivered its gradient (peer-memory exchange timed out): "
                     "that step was skipped (NaN loss, parameters untouched) and the trial cannot continue");
  return MDT_OK;
}

// Transitive reduction of a captured step graph.  Stream capture gives every kernel an edge from the previous node of
// ITS stream: in a graph that holds several steps, a side-stream kernel of step j (a weight-gradient GEMM, the input
// prefetch) thus depends on its step j-1 twin as well as on the main-stream kernel it was forked behind.  The first
// edge is redundant (step j-1 was joined into the main stream long before), but a kernel node with two incoming edges
// is not launched early on its programmatic one: measured, those kernels started 2.5-7 us late, which is why graphs of
// several steps were no faster than single-step ones.  Removing every edge u -> v for which another path u ~> v exists
// keeps the order (each kernel of the step waits for its programmatic predecessor before it reads or triggers anything,
// so order is transitive along such paths too) and gives those kernels their single, programmatic edge back.
int reduce_graph_edges(cudaGraph_t g, int* removed) {
  *removed = 0;
  size_t nn = 0, ne = 0;
  MDT_CUDA(cudaGraphGetNodes(g, nullptr, &nn));
  MDT_CUDA(cudaGraphGetEdges_v2(g, nullptr, nullptr, nullptr, &ne));
  if (nn == 0 || ne == 0 || nn > 4096) return MDT_OK;
  std::vector<cudaGraphNode_t> nodes(nn), from(ne), to(ne);
  std::vector<cudaGraphEdgeData> data(ne);
  MDT_CUDA(cudaGraphGetNodes(g, nodes.data(), &nn));
  MDT_CUDA(cudaGraphGetEdges_v2(g, from.data(), to.data(), data.data(), &ne));
  std::map<cudaGraphNode_t, int> id;
  for (size_t i = 0; i < nn; ++i) id[nodes[i]] = (int)i;
  std::vector<std::vector<int>> succ(nn);
  for (size_t k = 0; k < ne; ++k) succ[id[from[k]]].push_back(id[to[k]]);
  std::vector<char> seen(nn);
  std::vector<int> stack;
  for (size_t k = 0; k < ne; ++k) {
    const int u = id[from[k]], v = id[to[k]];
    // is v reachable from u without the direct edge?
    std::fill(seen.begin(), seen.end(), 0);
    stack.clear();
    for (int w : succ[u]) if (w != v && !seen[w]) { seen[w] = 1; stack.push_back(w); }
    bool redundant = false;
    while (!stack.empty() && !redundant) {
      const int w = stack.back(); stack.pop_back();
      for (int y : succ[w]) {
        if (y == v) { redundant = true; break; }
        if (!seen[y]) { seen[y] = 1; stack.push_back(y); }
      }
    }
    if (redundant) {
      MDT_CUDA(cudaGraphRemoveDependencies_v2(g, &from[k], &to[k], &data[k], 1));
      ++*removed;
    }
  }
  return MDT_OK;
}

// Run `fn` through the CUDA-graph cache when enabled.
template <class F>
int run_cached(mdt_engine* e, int kind, int B, int flag, F fn) {
  if (!e->cfg.use_graph || e->prof_on) return fn();
  auto key = std::make_tuple(kind, B, flag);
  auto it = e->graphs.find(key);
  if (it == e->graphs.end()) {
    cudaGraph_t g = nullptr;
    const long long before = e->launch_count;
    if (e->tc) e->tc->prep_launched = false;
    MDT_CUDA(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeRelaxed));
    int r = fn();
    cudaError_t ce = cudaStreamEndCapture(e->stream, &g);
    const long long kernels = e->launch_count - before;
    e->launch_count = before;
    if (r != MDT_OK) {
      if (g) cudaGraphDestroy(g);
      return r;
    }
    if (ce != cudaSuccess) return set_error(MDT_ERR_CUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(ce));
    if (e->opt_graph_reduce) {
      int removed = 0;
      const int rr = reduce_graph_edges(g, &removed);
      if (rr != MDT_OK) { cudaGraphDestroy(g); return rr; }
      if (e->opt_graph_dbg) fprintf(stderr, "[mdt] graph kind %d B %d: %lld kernels, %d redundant edges removed\n", kind, B, kernels, removed);
    }
    cudaGraphExec_t ex = nullptr;
    ce = cudaGraphInstantiate(&ex, g, 0);
    if (ce != cudaSuccess) { cudaGraphDestroy(g); return set_error(MDT_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(ce)); }
    mdt_engine::Graph entry{ex, kernels};
    if (e->tc && e->tc->prep_launched && !e->sourced) {
      // remember the prologue kernel's node: its x pointer is the one launch parameter that changes between steps
      size_t nn = 0;
      cudaGraphGetNodes(g, nullptr, &nn);
      std::vector<cudaGraphNode_t> nodes(nn);
      if (nn) cudaGraphGetNodes(g, nodes.data(), &nn);
      for (size_t i = 0; i < nn; ++i) {
        cudaGraphNodeType ty;
        cudaKernelNodeParams kp{};
        if (cudaGraphNodeGetType(nodes[i], &ty) != cudaSuccess || ty != cudaGraphNodeTypeKernel) continue;
        if (cudaGraphKernelNodeGetParams(nodes[i], &kp) != cudaSuccess) continue;
        if (kp.func == reinterpret_cast<void*>(prep_kernel)) {
          entry.prep_node = nodes[i]; entry.prep = e->tc->last_prep; entry.prep_grid = kp.gridDim;
          break;
        }
      }
      cudaGetLastError();
    }
    if (entry.prep_node) entry.graph = g; else cudaGraphDestroy(g);   // node handles live as long as their graph
    it = e->graphs.emplace(key, entry).first;
  }
  mdt_engine::Graph& G = it->second;
  if (G.prep_node != nullptr && e->tc) {
    const float* src = e->tc->x_src ? e->tc->x_src : e->x_stage;
    if (src != G.prep.tr.job[0].in) {
      G.prep.tr.job[0].in = src;
      G.prep.x_copy = (src != e->x_stage) ? e->x_stage : nullptr;
      void* args[1] = {&G.prep};
      cudaKernelNodeParams kp{};
      kp.func = reinterpret_cast<void*>(prep_kernel);
      kp.gridDim = G.prep_grid; kp.blockDim = dim3(256); kp.sharedMemBytes = 0; kp.kernelParams = args; kp.extra = nullptr;
      MDT_CUDA(cudaGraphExecKernelNodeSetParams(G.exec, G.prep_node, &kp));
    }
  }
  MDT_CUDA(cudaGraphLaunch(it->second.exec, e->stream));
  e->launch_count += it->second.kernels;
  return MDT_OK;
}

int stage_inputs(mdt_engine* e, const float** x, int B, const float** eps, cudaMemcpyKind kind) {
  e->primed = false;     // whatever mdt_train_steps prefetched into the stage buffers is about to be overwritten
  if (*x != e->x_stage && B > 0)
    MDT_CUDA(cudaMemcpyAsync(e->x_stage, *x, (size_t)B * e->D * sizeof(float), kind, e->stream));
  *x = e->x_stage;
  if (*eps != nullptr) {
    if (*eps != e->eps_stage && B > 0)
      MDT_CUDA(cudaMemcpyAsync(e->eps_stage, *eps, (size_t)B * e->Z * sizeof(float), kind, e->stream));
    *eps = e->eps_stage;
  }
  return MDT_OK;
}

// The two DDP buckets are contiguous ranges of the flat gradient buffer: bucket 0 = [off(W4), total) (fc4.weight, fc4.bias:
// first to be complete in backward), bucket 1 = [0, off(W4)).
int allreduce_bucket(mdt_engine* e, int bucket, cudaStream_t on) {
  const int64_t split = e->L.off[W4];
  float* p = bucket == 0 ? e->grads + split : e->grads;
  const size_t n = bucket == 0 ? (size_t)(e->L.total - split) : (size_t)split;
  MDT_NCCL(g_nccl.AllReduce(p, p, n, ncclFloat, ncclSum, e->comm, on));
  return MDT_OK;
}

int allreduce_grads(mdt_engine* e) {
  if (e->gsize <= 1) return MDT_OK;
  if (e->comm == nullptr)
    return set_error(MDT_ERR_STATE, "group_size %d but no NCCL communicator was given to mdt_engine_create", e->gsize);
  MDT_TRY(load_nccl());
  MDT_TRY(allreduce_bucket(e, 0, e->stream));
  return allreduce_bucket(e, 1, e->stream);
}

// Called by the backward pass as soon as the kernels that produce a bucket have been enqueued on stream `on`:
// the all-reduce of that bucket goes to the communication stream behind an event.
int bucket_ready_cb(void* ctx, int bucket, cudaStream_t on) {
  mdt_engine* e = static_cast<mdt_engine*>(ctx);
  if (e->gsize <= 1 || e->comm == nullptr || e->prof_on || e->defer_buckets || e->one_bucket) return MDT_OK;
  MDT_TRY(load_nccl());
  MDT_CUDA(cudaEventRecord(e->ev_bucket[bucket], on));
  MDT_CUDA(cudaStreamWaitEvent(e->comm_stream, e->ev_bucket[bucket], 0));
  MDT_TRY(allreduce_bucket(e, bucket, e->comm_stream));
  if (bucket == 0) e->bucket0_sent = true;
  return MDT_OK;
}

// All-reduce whatever has not been sent yet and make the step's stream wait for the communication stream.
int finish_allreduce(mdt_engine* e) {
  if (e->gsize <= 1) return MDT_OK;
  if (e->comm == nullptr)
    return set_error(MDT_ERR_STATE, "group_size %d but no NCCL communicator was given to mdt_engine_create", e->gsize);
  MDT_TRY(load_nccl());
  if (e->prof_on) { e->bucket0_sent = false; return allreduce_grads(e); }
  if (e->one_bucket) {
    // Below a few MB an all-reduce is latency-bound: two buckets cost two latencies (measured on 8 ranks: 2 x 31 us
    // against 36 us for the whole 2.6 MB gradient) and only the first overlaps backward.  One call, on the
    // communication stream so that the step's graph keeps the same shape.
    MDT_CUDA(cudaEventRecord(e->ev_bucket[1], e->stream));
    MDT_CUDA(cudaStreamWaitEvent(e->comm_stream, e->ev_bucket[1], 0));
    MDT_NCCL(g_nccl.AllReduce(e->grads, e->grads, (size_t)e->L.total, ncclFloat, ncclSum, e->comm, e->comm_stream));
  } else {
    if (!e->bucket0_sent) MDT_TRY(bucket_ready_cb(e, 0, e->stream));
    MDT_TRY(bucket_ready_cb(e, 1, e->stream));
  }
  e->bucket0_sent = false;
  MDT_CUDA(cudaEventRecord(e->ev_comm_done, e->comm_stream));
  MDT_CUDA(cudaStreamWaitEvent(e->stream, e->ev_comm_done, 0));
  return MDT_OK;
}

// forward + loss + backward of one step.  On the tcgen05 path the loss and d(logits) come out of fc4's epilogue.
int enqueue_fwd_loss_bwd(mdt_engine* e, const float* x, int B, const float* eps) {
  if (e->tc) {
    MDT_TRY(tc05_forward(e->tc, e->stream, x, B, eps, false, /*fuse_loss=*/true));
    return tc05_backward(e->tc, e->stream, x, B);
  }
  MDT_TRY(enqueue_forward_fp32(e, x, B, eps, false));
  MDT_TRY(enqueue_loss(e, x, B, true));
  return enqueue_backward_fp32(e, x, B);
}

// `steps` > 1 (mdt_train_steps only): that many consecutive steps in ONE graph.
int train_step_device(mdt_engine* e, const float* x, int B, const float* eps, int steps = 1, bool tail_no_prefetch = false) {
  e->last_B = B;
  if (B == 0 && e->gsize > 1)   // the peers of a group would wait for this rank's gradient: refuse instead of hanging them
    return set_error(MDT_ERR_INVALID, "an empty batch on one rank of a %d-rank trial group would stall its peers", e->gsize);
  if (B == 0) return MDT_OK;  // empty batch: nothing to do, no optimizer step (the reference's loader never yields one)
  const float gscale = 1.0f / (float)e->gsize;
  // graphs are per (eps given, SM share, steps) and per starting parity of the two-set buffers (Tc05Plan::xpar, w3par)
  const bool piped = e->tc != nullptr && e->tc->pipelined && e->tc->simt_latent;
  if (e->tc != nullptr && !piped) e->tc->xpar = 0;
  if (!e->sourced) e->primed = false;
  const int x0 = e->tc ? e->tc->xpar : 0, w0 = e->tc ? e->tc->w3par : 0;
  const bool w3_flips = e->tc != nullptr && e->tc->simt_latent && e->gsize == 1 && e->opt_fused_adam && !e->use_sgd;
  struct ParityAfter {       // a replayed graph does not run the enqueue code that flips the parities: set them here
    mdt_engine* e; int x, w;
    ~ParityAfter() { if (e->tc) { e->tc->xpar = x; e->tc->w3par = w; e->tc->prefetch_pending = false; e->tc->shadows_pending = false; } }
  } parity_after{e, piped && !e->tc->no_prefetch ? (x0 ^ ((steps - (tail_no_prefetch ? 1 : 0)) & 1)) : (piped ? x0 : 0),
                 w3_flips ? (w0 ^ (steps & 1)) : w0};
  const int fl = (eps != nullptr ? 1 : 0) | ((e->tc ? e->tc->sm_share : 1) << 1) | (steps << 8) | (x0 << 20) | (w0 << 21) |
                 ((piped ? 1 : 0) << 22) | ((piped && e->tc->no_prefetch ? 1 : 0) << 23) | ((piped && tail_no_prefetch ? 1 : 0) << 24);
  struct TailGuard {         // single-step graphs: the step itself is the tail
    Tc05Plan* p; bool saved, on;
    ~TailGuard() { if (on) p->no_prefetch = saved; }
  } tail_guard{e->tc, e->tc ? e->tc->no_prefetch : false, piped && tail_no_prefetch && steps == 1};
  if (tail_guard.on) e->tc->no_prefetch = true;
  if (e->tc) e->pub_expected = (e->pub_expected + (unsigned)steps) & 0x7fffffffu;   // every tensor-core training step publishes
                                                                         // (sequence, loss) once: publish_loss()
  if (e->tc && !e->sourced) e->cursor_expected = -1;                     // ... and moves the device-side batch cursor
  const int kbase = e->sourced ? 20 : 0;                                 // graphs of mdt_train_steps are keyed apart
  if (steps > 1) {
    auto one = [&]() -> int {
      if (e->gsize == 1 && e->opt_fused_adam && !e->use_sgd) {
        MDT_TRY(tc05_forward(e->tc, e->stream, x, B, eps, false, /*fuse_loss=*/true));
        return tc05_backward_fused_adam(e->tc, e->stream, B);
      }
      if (e->gsize > 1 && e->p2p && !e->use_sgd && e->exchange != MDT_EXCHANGE_NCCL) {
        e->tc->grads_override = e->xch;
        e->defer_buckets = true;
        int rc = enqueue_fwd_loss_bwd(e, x, B, eps);
        e->defer_buckets = false;
        e->tc->grads_override = nullptr;
        MDT_TRY(rc);
        return enqueue_adam_p2p(e, gscale);
      }
      MDT_TRY(enqueue_fwd_loss_bwd(e, x, B, eps));
      MDT_TRY(finish_allreduce(e));
      return enqueue_adam(e, gscale, true);
    };
    return run_cached(e, kbase + 9, B, fl, [&]() -> int {
      for (int i = 0; i < steps; ++i) {
        const bool tail = piped && tail_no_prefetch && i == steps - 1;
        const bool saved = e->tc->no_prefetch;
        if (tail) e->tc->no_prefetch = true;
        e->tc->defer_shadows = e->opt_defer_shadows && i < steps - 1;   // a later step of this capture joins the refresh (tc05_join_shadows)
        const int rc = one();
        e->tc->defer_shadows = false;
        e->tc->no_prefetch = saved;
        MDT_TRY(rc);
      }
      return MDT_OK;
    });
  }
  if (e->gsize == 1 && e->tc && e->opt_fused_adam && !e->use_sgd) {
    // single-rank group on the tensor-core path: Adam runs inside the weight-gradient epilogues
    return run_cached(e, kbase + 7, B, fl, [&]() -> int {
      MDT_TRY(tc05_forward(e->tc, e->stream, x, B, eps, false, /*fuse_loss=*/true));
      return tc05_backward_fused_adam(e->tc, e->stream, B);
    });
  }
  if (e->gsize == 1) {
    return run_cached(e, kbase + 0, B, fl, [&]() -> int {
      MDT_TRY(enqueue_fwd_loss_bwd(e, x, B, eps));
      return enqueue_adam(e, gscale, true);
    });
  }
  if (e->p2p && e->tc && !e->prof_on && !e->use_sgd && e->exchange != MDT_EXCHANGE_NCCL) {
    // group on one node with peer access: gradients are exchanged through NVLink peer memory inside the Adam kernel
    return run_cached(e, kbase + 8, B, fl, [&]() -> int {
      e->tc->grads_override = e->xch;
      e->defer_buckets = true;                    // no NCCL buckets in this step
      int rc = enqueue_fwd_loss_bwd(e, x, B, eps);
      e->defer_buckets = false;
      e->tc->grads_override = nullptr;
      MDT_TRY(rc);
      return enqueue_adam_p2p(e, gscale);
    });
  }
  // NCCL outside the graphs: forward + backward as one graph, the all-reduces enqueued eagerly, the optimizer as a second
  // graph.  Taken for MDT_NCCL_GRAPH=0 and for DDP's two-bucket form in groups of more than 4 ranks: captured, the
  // 12.8 + 13.9 MB all-reduces of H = 4096 fail inside ncclAllReduce on 8 ranks ("unhandled cuda error"; the same calls
  // succeed eagerly, and captured on 4 ranks), and at that size a step is long enough not to care about two launches.
  if (!e->opt_nccl_in_graph || (!e->one_bucket && e->gsize > 4)) {
    e->defer_buckets = true;     // no all-reduce inside the captured backward
    const int rc1 = run_cached(e, kbase + 1, B, fl, [&]() -> int { return enqueue_fwd_loss_bwd(e, x, B, eps); });
    e->defer_buckets = false;
    MDT_TRY(rc1);
    MDT_TRY(finish_allreduce(e));
    return run_cached(e, kbase + 2, B, fl & ~1, [&]() -> int { return enqueue_adam(e, gscale, true); });
  }
  // one graph for the whole step, the group's NCCL all-reduces included (bucket 0 overlaps the rest of backward)
  return run_cached(e, kbase + 6, B, fl, [&]() -> int {
    MDT_TRY(enqueue_fwd_loss_bwd(e, x, B, eps));
    MDT_TRY(finish_allreduce(e));
    return enqueue_adam(e, gscale, true);
  });
}

}  // namespace

// =================================================================== C ABI
extern "C" {

int mdt_version(void) { return MDT_VERSION; }
const char* mdt_last_error(void) { return g_last_error; }

int mdt_param_layout(const mdt_cfg* cfg, int64_t offsets[MDT_NUM_SEGMENTS], int64_t numels[MDT_NUM_SEGMENTS],
                     int64_t* total) {
  MDT_CHECK(cfg && cfg->in_dim > 0 && cfg->hidden > 0 && cfg->latent > 0, "bad cfg");
  const Layout L = make_layout(cfg->in_dim, cfg->hidden, cfg->latent);
  for (int i = 0; i < kNumSeg; ++i) {
    if (offsets) offsets[i] = L.off[i];
    if (numels) numels[i] = L.numel[i];
  }
  if (total) *total = L.total;
  return MDT_OK;
}

int mdt_engine_create(const mdt_cfg* cfg, void* nccl_comm, int group_size, int group_rank, void* stream,
                      mdt_engine** out) {
  MDT_CHECK(cfg && out, "cfg/out is NULL");
  MDT_CHECK(cfg->in_dim > 0 && cfg->hidden > 0 && cfg->latent > 0 && cfg->max_batch > 0, "bad dims in cfg");
  MDT_CHECK(cfg->in_dim % 4 == 0 && cfg->hidden % 4 == 0 && cfg->latent % 4 == 0,
            "in_dim, hidden and latent must be multiples of 4 (16-byte rows)");
  MDT_CHECK(group_size >= 1 && group_rank >= 0 && group_rank < group_size, "bad group size/rank");
  MDT_CHECK(cfg->gemm_mode >= MDT_GEMM_FP32 && cfg->gemm_mode <= MDT_GEMM_BF16, "unknown gemm_mode %d", cfg->gemm_mode);
  MDT_CHECK(cfg->exchange >= MDT_EXCHANGE_AUTO && cfg->exchange <= MDT_EXCHANGE_P2P_TWO_SHOT, "unknown exchange %d", cfg->exchange);
  DeviceInfo* di = nullptr;
  MDT_TRY(device_info(&di));
  if (cfg->gemm_mode != MDT_GEMM_FP32 && di->cc_major != 10)
    return set_error(MDT_ERR_UNSUPPORTED, "tcgen05 GEMM modes need an sm_100 device (found sm_%d%d)", di->cc_major, di->cc_minor);
  mdt_engine* e = new mdt_engine();
  e->cfg = *cfg;
  e->D = cfg->in_dim; e->H = cfg->hidden; e->Z = cfg->latent; e->Bmax = cfg->max_batch;
  e->L = make_layout(e->D, e->H, e->Z);
  e->dev = di;
  cudaGetDevice(&e->device);
  e->comm = reinterpret_cast<ncclComm_t>(nccl_comm);
  e->gsize = group_size; e->grank = group_rank;
  e->opt_fused_adam = getenv("MDT_NO_FUSED_ADAM") == nullptr;
  e->opt_nccl_in_graph = !(getenv("MDT_NCCL_GRAPH") && getenv("MDT_NCCL_GRAPH")[0] == '0');
  e->opt_zerocopy = getenv("MDT_NO_ZEROCOPY") == nullptr;
  e->opt_h2d_prefetch = getenv("MDT_NO_H2D_PREFETCH") == nullptr;
  e->opt_pipeline = getenv("MDT_NO_PIPELINE") == nullptr;
  e->opt_graph_reduce = getenv("MDT_NO_GRAPH_REDUCE") == nullptr;
  e->opt_graph_dbg = getenv("MDT_GRAPH_DBG") != nullptr;
  if (const char* v = getenv("MDT_STEPS_PER_GRAPH")) e->opt_steps_per_graph = atoi(v);
  if (const char* v = getenv("MDT_P2P_SINGLE_STEPS")) e->opt_p2p_single = atoi(v) != 0;
  if (const char* v = getenv("MDT_NO_DEFER_SHADOWS")) e->opt_defer_shadows = atoi(v) == 0;
  e->exchange = cfg->exchange;
  if (const char* v = getenv("MDT_P2P_TWO_SHOT_FROM")) e->two_shot_from = atoi(v);
  if (const char* v = getenv("MDT_P2P_TWO_SHOT_BYTES")) e->two_shot_bytes = atoll(v);
  e->one_bucket = (size_t)e->L.total * sizeof(float) < ((size_t)8 << 20) && getenv("MDT_NCCL_TWO_BUCKETS") == nullptr;
  if (stream) {
    e->stream = reinterpret_cast<cudaStream_t>(stream);
  } else {
    // a *blocking* stream: implicitly ordered with the legacy default stream torch uses by default,
    // yet capturable into a CUDA graph (the legacy stream is not)
    cudaError_t ce = cudaStreamCreate(&e->stream);
    if (ce != cudaSuccess) { delete e; return set_error(MDT_ERR_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(ce)); }
    e->own_stream = true;
  }
  int r = carve_workspace(e);
  if (r == MDT_OK && group_size > 1) {
    if (cudaStreamCreateWithFlags(&e->comm_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ev_bucket[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ev_bucket[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ev_comm_done, cudaEventDisableTiming) != cudaSuccess)
      r = set_error(MDT_ERR_CUDA, "could not create the communication stream/events");
  }
  if (r == MDT_OK) {
    cudaError_t ce = cudaMallocHost(&e->loss_pinned, 1024);
    if (ce != cudaSuccess) r = set_error(MDT_ERR_CUDA, "cudaMallocHost: %s", cudaGetErrorString(ce));
  }
  if (r == MDT_OK) memset(e->loss_pinned, 0, 1024);
  if (r == MDT_OK && e->cfg.gemm_mode != MDT_GEMM_FP32) {
    // pinned page: [0] loss of synchronous steps, [8] exchange-error word, [16..144) mailbox ring of 64 words
    e->mailbox = reinterpret_cast<volatile unsigned long long*>(e->loss_pinned + 16);
    void* dptr = nullptr;
    if (cudaHostGetDevicePointer(&dptr, e->loss_pinned, 0) == cudaSuccess && dptr != nullptr) {
      unsigned long long* mb = reinterpret_cast<unsigned long long*>(static_cast<float*>(dptr) + 16);
      unsigned int* ew = reinterpret_cast<unsigned int*>(static_cast<float*>(dptr) + 8);
      if (cudaMemcpy(&e->st->host_mailbox, &mb, sizeof(mb), cudaMemcpyHostToDevice) != cudaSuccess) e->mailbox = nullptr;
      if (cudaMemcpy(&e->st->host_error, &ew, sizeof(ew), cudaMemcpyHostToDevice) == cudaSuccess)
        e->err_word = reinterpret_cast<volatile unsigned int*>(e->loss_pinned + 8);
    } else {
      e->mailbox = nullptr;
    }
    cudaGetLastError();
    if (cudaMemcpy(&e->st->loss_hist, &e->loss_hist, sizeof(float*), cudaMemcpyHostToDevice) != cudaSuccess)
      r = set_error(MDT_ERR_CUDA, "could not publish the loss-history ring");
    // host-buffer steps: copy stream, ring of device input buffers and their events -- created here, not on the first
    // submit, so that no step ever pays for a cudaMalloc
    if (r == MDT_OK && e->mailbox != nullptr) {
      if (cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking) != cudaSuccess)
        r = set_error(MDT_ERR_CUDA, "could not create the copy stream");
      for (int k = 0; k < mdt_engine::kPreSlots && r == MDT_OK; ++k) {
        if ((k < 2 && (cudaEventCreateWithFlags(&e->ev_set_ready[k], cudaEventDisableTiming) != cudaSuccess ||
                       cudaEventCreateWithFlags(&e->ev_set_free[k], cudaEventDisableTiming) != cudaSuccess)) ||
            cudaMalloc(&e->x_pre[k], (size_t)e->Bmax * e->D * sizeof(float)) != cudaSuccess ||
            cudaEventCreateWithFlags(&e->ev_copied[k], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&e->ev_used[k], cudaEventDisableTiming) != cudaSuccess)
          r = set_error(MDT_ERR_CUDA, "could not create the host-input staging ring (%zu bytes per slot)",
                        (size_t)e->Bmax * e->D * sizeof(float));
      }
    }
  }
  if (r == MDT_OK) {
    mdt_adam h{1e-3, 0.9, 0.999, 1e-8};
    r = mdt_engine_set_adam(e, &h, 0);
  }
  if (r != MDT_OK) { mdt_engine_destroy(e); return r; }
  *out = e;
  return MDT_OK;
}

void mdt_engine_destroy(mdt_engine* e) {
  if (!e) return;
  cudaStreamSynchronize(e->stream);
  for (auto& kv : e->graphs) { cudaGraphExecDestroy(kv.second.exec); if (kv.second.graph) cudaGraphDestroy(kv.second.graph); }
  if (e->tc) tc05_destroy(e->tc);
  if (e->comm_stream) { cudaStreamSynchronize(e->comm_stream); cudaStreamDestroy(e->comm_stream); }
  for (auto& ev : e->ev_bucket) if (ev) cudaEventDestroy(ev);
  if (e->ev_comm_done) cudaEventDestroy(e->ev_comm_done);
  if (e->copy_stream) { cudaStreamSynchronize(e->copy_stream); cudaStreamDestroy(e->copy_stream); }
  for (int k = 0; k < mdt_engine::kPreSlots; ++k) {
    if (e->x_pre[k]) cudaFree(e->x_pre[k]);
    if (e->ev_copied[k]) cudaEventDestroy(e->ev_copied[k]);
    if (e->ev_used[k]) cudaEventDestroy(e->ev_used[k]);
    if (k < 2 && e->ev_set_ready[k]) cudaEventDestroy(e->ev_set_ready[k]);
    if (k < 2 && e->ev_set_free[k]) cudaEventDestroy(e->ev_set_free[k]);
  }
  for (void* p : e->xch_opened) if (p) cudaIpcCloseMemHandle(p);
  if (e->x_ring) cudaFree(e->x_ring);
  for (int k = 0; k < 2; ++k) {
    if (e->ev_ring_copied[k]) cudaEventDestroy(e->ev_ring_copied[k]);
    if (e->ev_ring_free[k]) cudaEventDestroy(e->ev_ring_free[k]);
  }
  if (e->xch) cudaFree(e->xch);
  if (e->slab) cudaFree(e->slab);
  if (e->loss_pinned) cudaFreeHost(e->loss_pinned);
  if (e->own_stream) cudaStreamDestroy(e->stream);
  delete e;
}

int mdt_engine_exchange_handle(mdt_engine* e, void* handle64) {
  MDT_CHECK(e && handle64, "NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  if (e->xch == nullptr) {
    const size_t bytes = (size_t)2 * e->L.total * sizeof(float) + 64 * sizeof(unsigned int);   // gradient, reduced slices, flags
    MDT_CUDA(cudaMalloc(&e->xch, bytes));
    MDT_CUDA(cudaMemset(e->xch, 0, bytes));
    MDT_CUDA(cudaDeviceSynchronize());
  }
  cudaIpcMemHandle_t h;
  MDT_CUDA(cudaIpcGetMemHandle(&h, e->xch));
  memcpy(handle64, &h, sizeof(h));
  return MDT_OK;
}

int mdt_engine_set_peers(mdt_engine* e, const void* handles64, int n) {
  MDT_CHECK(e && handles64, "NULL argument");
  MDT_CHECK(n == e->gsize && n >= 2 && n <= kMaxPeers, "set_peers: %d handles for a group of %d (2..%d ranks supported)", n, e->gsize, kMaxPeers);
  MDT_CHECK(e->xch != nullptr, "call mdt_engine_exchange_handle first");
  MDT_CHECK(e->cfg.gemm_mode != MDT_GEMM_FP32, "the peer-memory exchange is part of the tensor-core path");
  PeerArgs pa{};
  pa.n = n; pa.me = e->grank;
  for (int r = 0; r < n; ++r) {
    float* base = e->xch;
    if (r != e->grank) {
      cudaIpcMemHandle_t h;
      memcpy(&h, static_cast<const char*>(handles64) + 64 * r, sizeof(h));
      void* p = nullptr;
      cudaError_t ce = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
      if (ce != cudaSuccess) {
        cudaGetLastError();
        return set_error(MDT_ERR_CUDA, "cudaIpcOpenMemHandle(rank %d): %s (no peer access?)", r, cudaGetErrorString(ce));
      }
      e->xch_opened[r] = p;
      base = static_cast<float*>(p);
    }
    pa.g[r] = base;
    pa.red[r] = base + e->L.total;
    pa.flags[r] = reinterpret_cast<unsigned int*>(base + 2 * e->L.total);
  }
  e->peers = pa;
  e->p2p = true;
  for (auto& kv : e->graphs) { cudaGraphExecDestroy(kv.second.exec); if (kv.second.graph) cudaGraphDestroy(kv.second.graph); }
  e->graphs.clear();
  return MDT_OK;
}

int mdt_engine_disable_peers(mdt_engine* e) {
  MDT_CHECK(e, "NULL engine");
  if (!e->p2p) return MDT_OK;
  MDT_CUDA(cudaStreamSynchronize(e->stream));
  e->p2p = false;
  for (auto& kv : e->graphs) { cudaGraphExecDestroy(kv.second.exec); if (kv.second.graph) cudaGraphDestroy(kv.second.graph); }
  e->graphs.clear();
  return MDT_OK;
}

int mdt_engine_exchange_status(mdt_engine* e, int* enabled, int* timed_out) {
  MDT_CHECK(e, "NULL engine");
  StepState h;
  MDT_CUDA(cudaMemcpyAsync(&h, e->st, sizeof(h), cudaMemcpyDeviceToHost, e->stream));
  MDT_CUDA(cudaStreamSynchronize(e->stream));
  if (enabled) *enabled = e->p2p ? 1 : 0;
  if (timed_out) *timed_out = (int)h.xch_error;
  return MDT_OK;
}

int mdt_engine_set_sm_share(mdt_engine* e, int n_sharing) {
  MDT_CHECK(e, "NULL engine");
  MDT_CHECK(n_sharing >= 1 && n_sharing <= 16, "n_sharing %d outside [1, 16]", n_sharing);
  if (e->tc) e->tc->sm_share = n_sharing;
  e->sm_share_pending = n_sharing;      // applied when the tensor-core plan is (re)built by mdt_engine_bind_buffers
  return MDT_OK;
}

int mdt_engine_stream(mdt_engine* e, void** stream) {
  MDT_CHECK(e && stream, "NULL argument");
  *stream = e->stream;
  return MDT_OK;
}

int mdt_engine_bind_buffers(mdt_engine* e, float* params, float* grads, float* exp_avg, float* exp_avg_sq,
                            size_t numel) {
  MDT_CHECK(e && params && grads && exp_avg && exp_avg_sq, "NULL buffer");
  MDT_CHECK((int64_t)numel == e->L.total, "numel %zu != layout total %lld", numel, (long long)e->L.total);
  for (const void* p : {(const void*)params, (const void*)grads, (const void*)exp_avg, (const void*)exp_avg_sq})
    MDT_CHECK(reinterpret_cast<uintptr_t>(p) % 128 == 0, "flat buffers must be 128-byte aligned");
  e->params = params; e->grads = grads; e->m = exp_avg; e->v = exp_avg_sq;
  for (auto& kv : e->graphs) { cudaGraphExecDestroy(kv.second.exec); if (kv.second.graph) cudaGraphDestroy(kv.second.graph); }
  e->graphs.clear();
  if (e->cfg.gemm_mode != MDT_GEMM_FP32) {
    if (e->tc) { tc05_destroy(e->tc); e->tc = nullptr; }
    Tc05Buffers b{};
    b.params = params; b.grads = grads; b.exp_avg = exp_avg; b.exp_avg_sq = exp_avg_sq; b.L = e->L; b.D = e->D; b.H = e->H; b.Z = e->Z; b.Bmax = e->Bmax;
    b.x_stage = e->x_stage; b.h1 = e->h1; b.ml = e->ml; b.eps = e->eps; b.sd = e->sd; b.z = e->z; b.h3 = e->h3;
    b.recon = e->recon; b.da4 = e->da4; b.da3 = e->da3; b.dz = e->dz; b.dml_kl = e->dml_kl; b.dml = e->dml; b.da1 = e->da1;
    b.st = e->st; b.sm_count = e->dev->sm_count; b.max_threads_sm = e->dev->max_threads_sm;
    b.loss_dev = e->loss_dev; b.note_ctx = e; b.note = note_cb; b.serial = &e->prof_on;
    b.bucket_ready = (e->gsize > 1) ? bucket_ready_cb : nullptr;
    b.split3 = (e->cfg.gemm_mode == MDT_GEMM_TF32X3);
    b.bf16 = (e->cfg.gemm_mode == MDT_GEMM_BF16);
    MDT_TRY(tc05_create(&b, e->stream, &e->tc));
    e->tc->sm_share = e->sm_share_pending;
  }
  e->bound = true;
  return MDT_OK;
}

int mdt_engine_broadcast_params(mdt_engine* e) {
  MDT_TRY(check_batch(e, 0));
  if (e->gsize <= 1) return MDT_OK;
  if (!e->comm) return set_error(MDT_ERR_STATE, "no NCCL communicator");
  MDT_TRY(load_nccl());
  MDT_NCCL(g_nccl.Broadcast(e->params, e->params, (size_t)e->L.total, ncclFloat, 0, e->comm, e->stream));
  if (e->tc) MDT_TRY(tc05_params_changed(e->tc, e->stream));
  return MDT_OK;
}

int mdt_engine_set_adam(mdt_engine* e, const mdt_adam* h, int64_t completed_steps) {
  MDT_CHECK(e && h, "NULL argument");
  struct { long long t; double lr, b1, b2, eps, b1_pow, b2_pow; float next[6]; } host{
      completed_steps, h->lr, h->beta1, h->beta2, h->eps, pow(h->beta1, (double)completed_steps),
      pow(h->beta2, (double)completed_steps), {0, 0, 0, 0, 0, 0}};
  adam_next_scalars(host.lr, host.b1, host.b2, host.eps, host.b1_pow, host.b2_pow, host.next);
  MDT_CUDA(cudaMemcpyAsync(&e->st->adam_t, &host, sizeof(host), cudaMemcpyHostToDevice, e->stream));
  MDT_CUDA(cudaStreamSynchronize(e->stream));  // `host` lives on this stack frame
  if (e->use_sgd) {                            // back from SGD: the captured steps hold the other optimizer's kernels
    e->use_sgd = false;
    for (auto& kv : e->graphs) { cudaGraphExecDestroy(kv.second.exec); if (kv.second.graph) cudaGraphDestroy(kv.second.graph); }
    e->graphs.clear();
  }
  return MDT_OK;
}

namespace {
void drop_graphs(mdt_engine* e) {
  for (auto& kv : e->graphs) { cudaGraphExecDestroy(kv.second.exec); if (kv.second.graph) cudaGraphDestroy(kv.second.graph); }
  e->graphs.clear();
}
}  // namespace

int mdt_engine_set_sgd(mdt_engine* e, const mdt_sgd* h, int64_t completed_steps) {
  MDT_CHECK(e && h, "NULL argument");
  MDT_CHECK(h->lr >= 0 && h->momentum >= 0 && h->weight_decay >= 0, "negative SGD hyper-parameter");
  MDT_CHECK(!(h->nesterov && (h->momentum <= 0 || h->dampening != 0)), "Nesterov momentum requires a momentum and zero dampening");
  MDT_CUDA(cudaStreamSynchronize(e->stream));
  e->use_sgd = true;
  e->sgd = SgdScalars{(float)h->lr, (float)h->momentum, (float)h->dampening, (float)h->weight_decay, h->nesterov ? 1 : 0};
  long long t = completed_steps;
  MDT_CUDA(cudaMemcpy(&e->st->adam_t, &t, sizeof(t), cudaMemcpyHostToDevice));
  drop_graphs(e);                                  // the scalars are kernel arguments of the captured steps
  return MDT_OK;
}

int mdt_engine_set_rng(mdt_engine* e, uint64_t philox_seed, uint64_t philox_offset) {
  MDT_CHECK(e, "NULL engine");
  MDT_CHECK(philox_offset % 4 == 0, "philox offset must be a multiple of 4 (torch keeps it so)");
  unsigned long long host[2] = {philox_seed, philox_offset};
  MDT_CUDA(cudaMemcpyAsync(&e->st->philox_seed, host, sizeof(host), cudaMemcpyHostToDevice, e->stream));
  MDT_CUDA(cudaStreamSynchronize(e->stream));
  return MDT_OK;
}

int mdt_engine_get_state(mdt_engine* e, int64_t* completed_steps, uint64_t* philox_offset) {
  MDT_CHECK(e, "NULL engine");
  StepState h;
  MDT_CUDA(cudaMemcpyAsync(&h, e->st, sizeof(h), cudaMemcpyDeviceToHost, e->stream));
  MDT_CUDA(cudaStreamSynchronize(e->stream));
  if (completed_steps) *completed_steps = h.adam_t;
  if (philox_offset) *philox_offset = h.philox_offset;
  return MDT_OK;
}

int mdt_engine_input_buffer(mdt_engine* e, float** x_stage) {
  MDT_CHECK(e && x_stage, "NULL argument");
  *x_stage = e->x_stage;
  return MDT_OK;
}

int mdt_forward(mdt_engine* e, const float* x, int batch, const float* eps_or_null, float* recon, float* mu,
                float* logvar) {
  MDT_TRY(check_batch(e, batch));
  MDT_CHECK(x, "x is NULL");
  e->last_B = batch;
  if (batch == 0) return MDT_OK;
  const float* eps = eps_or_null;
  MDT_TRY(stage_inputs(e, &x, batch, &eps, cudaMemcpyDeviceToDevice));
  MDT_TRY(run_cached(e, 3, batch, eps != nullptr, [&]() -> int { return enqueue_forward(e, x, batch, eps, false); }));
  const size_t zb = e->Z * sizeof(float);
  if (recon) MDT_CUDA(cudaMemcpyAsync(recon, e->recon, (size_t)batch * e->D * sizeof(float), cudaMemcpyDeviceToDevice, e->stream));
  if (mu) MDT_CUDA(cudaMemcpy2DAsync(mu, zb, e->ml, 2 * zb, zb, batch, cudaMemcpyDeviceToDevice, e->stream));
  if (logvar) MDT_CUDA(cudaMemcpy2DAsync(logvar, zb, e->ml + e->Z, 2 * zb, zb, batch, cudaMemcpyDeviceToDevice, e->stream));
  return MDT_OK;
}

int mdt_backward(mdt_engine* e, const float* d_recon, const float* d_mu, const float* d_logvar) {
  MDT_TRY(check_batch(e, 0));
  const int B = e->last_B;
  if (B == 0) return MDT_OK;
  MDT_CHECK(d_recon, "d_recon is NULL");
  const long long n = (long long)B * e->D;
  int blocks = (int)((n + 255) / 256);
  if (blocks > e->dev->sm_count * 8) blocks = e->dev->sm_count * 8;
  sigmoid_bwd_pack_kernel<<<blocks, 256, 0, e->stream>>>(d_recon, e->recon, n, e->da4, d_mu, d_logvar, B, e->Z, e->dml_kl);
  MDT_TRY(note_launch(e, "sigmoid_bwd"));
  if (e->tc) MDT_TRY(tc05_after_external_da4(e->tc, e->stream, B));
  e->defer_buckets = true;
  const int rc = run_cached(e, 4, B, 0, [&]() -> int { return enqueue_backward(e, e->x_stage, B); });
  e->defer_buckets = false;
  return rc;
}

int mdt_allreduce_grads(mdt_engine* e) {
  MDT_TRY(check_batch(e, 0));
  return finish_allreduce(e);
}

int mdt_adam_step(mdt_engine* e, float grad_scale) {
  MDT_TRY(check_batch(e, 0));
  return enqueue_adam(e, grad_scale);
}

int mdt_train_step(mdt_engine* e, const float* x, int batch, const float* eps_or_null, float* loss_out) {
  MDT_TRY(check_batch(e, batch));
  MDT_CHECK(x || batch == 0, "x is NULL");
  const float* eps = eps_or_null;
  if (e->tc) {
    // tensor-core path: the step's prologue kernel reads x where it lies and fills the stage buffer itself
    e->tc->x_src = x;
    x = e->x_stage;
  }
  MDT_TRY(stage_inputs(e, &x, batch, &eps, cudaMemcpyDeviceToDevice));
  if (batch == 0) MDT_CUDA(cudaMemsetAsync(e->loss_dev, 0, sizeof(float), e->stream));
  const int rc_step = train_step_device(e, x, batch, eps);
  if (e->tc) e->tc->x_src = nullptr;
  MDT_TRY(rc_step);
  if (loss_out) MDT_CUDA(cudaMemcpyAsync(loss_out, e->loss_dev, sizeof(float), cudaMemcpyDeviceToDevice, e->stream));
  return MDT_OK;
}

int mdt_train_steps(mdt_engine* e, const mdt_batch_source* src, int batch, int64_t first_batch, int n_steps,
                    float* losses_out) {
  MDT_TRY(check_batch(e, batch));
  MDT_CHECK(src != nullptr && n_steps >= 0 && batch > 0, "bad argument");
  MDT_CHECK(src->n_batches > 0 && first_batch >= 0, "the batch source is empty or first_batch is negative");
  MDT_CHECK(src->x_pool != nullptr || (src->images_u8 != nullptr && src->indices != nullptr && src->n_images > 0),
            "the batch source needs x_pool, or images_u8 + indices + n_images");
  MDT_CHECK(losses_out == nullptr || n_steps <= (int)kLossHist, "losses_out holds at most %u steps per call", kLossHist);
  if (n_steps == 0) return MDT_OK;
  if (e->tc == nullptr) {
    // fp32 FFMA path (no step prologue kernel): one call per batch
    MDT_CHECK(src->x_pool != nullptr, "the uint8 source needs a tensor-core gemm_mode");
    for (int i = 0; i < n_steps; ++i) {
      const float* x = src->x_pool + (size_t)((first_batch + i) % src->n_batches) * batch * e->D;
      MDT_TRY(mdt_train_step(e, x, batch, nullptr, losses_out ? losses_out + i : nullptr));
    }
    return MDT_OK;
  }
  // MDT_P2P_SINGLE_STEPS=1: groups that exchange gradients over peer memory take one mdt_train_step per batch (the
  // form every DDP fixture checks) instead of multi-step graphs.  Multi-step graphs against single steps were compared
  // bit for bit on one GPU (fused and separate optimizer) and on 2 GPUs with the NCCL exchange; the one 2-GPU
  // peer-memory comparison of round 2 came out negative for reasons it could not pin down, so
  // bench.py --gpus N repeats it before timing and falls back to this switch when it fails.
  if (e->opt_p2p_single && e->gsize > 1 && e->p2p && !e->use_sgd && e->exchange != MDT_EXCHANGE_NCCL) {
    const unsigned int seq_before = e->pub_expected;
    for (int i = 0; i < n_steps; ++i) {
      const long long idx = (first_batch + i) % src->n_batches;
      const float* x = nullptr;
      if (src->x_pool != nullptr) {
        x = src->x_pool + (size_t)idx * batch * e->D;
      } else {
        MDT_TRY(mdt_gather_u8(src->images_u8, src->n_images, src->indices + idx * batch, batch, e->D, e->x_stage, e->stream));
        x = e->x_stage;
      }
      MDT_TRY(mdt_train_step(e, x, batch, nullptr, nullptr));
    }
    if (losses_out != nullptr) {
      const unsigned int lo = seq_before & (kLossHist - 1u);
      const unsigned int n1 = (lo + (unsigned)n_steps <= kLossHist) ? (unsigned)n_steps : kLossHist - lo;
      MDT_CUDA(cudaMemcpyAsync(losses_out, e->loss_hist + lo, n1 * sizeof(float), cudaMemcpyDeviceToDevice, e->stream));
      if (n1 < (unsigned)n_steps)
        MDT_CUDA(cudaMemcpyAsync(losses_out + n1, e->loss_hist, ((unsigned)n_steps - n1) * sizeof(float),
                                 cudaMemcpyDeviceToDevice, e->stream));
    }
    return MDT_OK;
  }
  // the batch index of the next step lives on the device (StepState.x_cursor, +1 per published step): only written from
  // here when the caller jumps somewhere the device does not expect
  if (e->cursor_expected < 0 || (e->cursor_expected - first_batch) % src->n_batches != 0) {
    const long long c = first_batch;
    MDT_CUDA(cudaMemcpyAsync(&e->st->x_cursor, &c, sizeof(c), cudaMemcpyHostToDevice, e->stream));
    MDT_CUDA(cudaStreamSynchronize(e->stream));     // `c` lives on this stack frame
  }
  Tc05Plan* p = e->tc;
  p->src_pool = src->x_pool; p->src_images = src->x_pool ? nullptr : src->images_u8;
  p->src_indices = reinterpret_cast<const long long*>(src->indices);
  p->src_n_batches = src->n_batches; p->src_n_images = src->n_images;
  int U = e->opt_steps_per_graph < 1 ? 1 : (e->opt_steps_per_graph > 32 ? 32 : e->opt_steps_per_graph);
  if (U > 1 && (U & 1)) ++U;                         // a multi-step graph must leave the two-set parities where it found them
  // NCCL with DDP's two buckets (large gradients): one step per graph, the form that has run on 4 and 8 GPUs since
  // round 1 -- sixteen steps' worth of bucketed all-reduces in ONE capture failed inside ncclAllReduce ("unhandled
  // cuda error") on 8 ranks at H = 4096
  const bool nccl_step = e->gsize > 1 && !(e->p2p && !e->use_sgd && e->exchange != MDT_EXCHANGE_NCCL);
  if (nccl_step && !e->one_bucket) U = 1;
  // graphs are specific to the source (its pointers are kernel arguments): a new source drops the cached ones
  const void* sig[4] = {src->x_pool, src->x_pool ? nullptr : (const void*)src->images_u8, src->indices,
                        reinterpret_cast<const void*>((uintptr_t)src->n_batches)};
  const bool same_src = memcmp(sig, e->src_sig, sizeof(sig)) == 0;
  if (!same_src) {
    for (auto it = e->graphs.begin(); it != e->graphs.end();) {
      if (std::get<0>(it->first) >= 20) {
        cudaGraphExecDestroy(it->second.exec); if (it->second.graph) cudaGraphDestroy(it->second.graph);
        it = e->graphs.erase(it);
      } else {
        ++it;
      }
    }
    memcpy(e->src_sig, sig, sizeof(sig));
  }
  const unsigned int seq0 = e->pub_expected;        // published steps before this call
  int rc = MDT_OK;
  // Pipelined steps (CUDA-core latent path): no prologue kernel inside the step -- each step fetches its successor's
  // batch into the other input set while it runs.  Only the first step of a call that does not continue the previous
  // one needs its batch fetched up front.
  const bool piped = e->opt_pipeline && p->simt_latent && !e->prof_on;
  if (piped) {
    const bool primed = e->primed && same_src && e->primed_cursor >= 0 && (e->primed_cursor - first_batch) % src->n_batches == 0;
    if (!primed) rc = tc05_prep(p, e->stream, e->x_stage, batch, p->xpar, false, 0, "prep_xT");
    p->pipelined = true;
  }
  e->sourced = true;
  for (int done = 0; done < n_steps && rc == MDT_OK;) {
    const int u = (n_steps - done >= U) ? U : 1;
    rc = train_step_device(e, e->x_stage, batch, nullptr, u, e->tail_no_prefetch && done + u == n_steps);
    done += u;
  }
  e->sourced = false;
  p->pipelined = false;
  p->src_pool = nullptr; p->src_images = nullptr; p->src_indices = nullptr;
  if (rc != MDT_OK) e->primed = false;
  MDT_TRY(rc);
  e->cursor_expected = first_batch + n_steps;
  e->primed = piped && !e->tail_no_prefetch;
  e->primed_cursor = first_batch + n_steps;
  if (losses_out != nullptr) {
    const unsigned int lo = seq0 & (kLossHist - 1u);
    const unsigned int n1 = (lo + (unsigned)n_steps <= kLossHist) ? (unsigned)n_steps : kLossHist - lo;
    MDT_CUDA(cudaMemcpyAsync(losses_out, e->loss_hist + lo, n1 * sizeof(float), cudaMemcpyDeviceToDevice, e->stream));
    if (n1 < (unsigned)n_steps)
      MDT_CUDA(cudaMemcpyAsync(losses_out + n1, e->loss_hist, ((unsigned)n_steps - n1) * sizeof(float),
                               cudaMemcpyDeviceToDevice, e->stream));
  }
  return MDT_OK;
}

int mdt_engine_exchange_bench(mdt_engine* e, int iters, double* seconds) {
  MDT_TRY(check_batch(e, 0));
  MDT_CHECK(seconds && iters > 0, "bad argument");
  MDT_CHECK(e->p2p && e->tc, "the peer-memory exchange is not enabled on this engine (mdt_engine_set_peers)");
  cudaEvent_t e0, e1;
  MDT_CUDA(cudaEventCreate(&e0));
  MDT_CUDA(cudaEventCreate(&e1));
  const float gs = 1.0f / (float)e->gsize;
  e->last_B = e->Bmax;
  int rc = enqueue_adam_p2p(e, gs);                  // warm-up (also the first cross-rank rendezvous)
  if (rc == MDT_OK) rc = cudaEventRecord(e0, e->stream) == cudaSuccess ? MDT_OK : set_error(MDT_ERR_CUDA, "event");
  for (int i = 0; i < iters && rc == MDT_OK; ++i) rc = enqueue_adam_p2p(e, gs);
  if (rc == MDT_OK) {
    cudaEventRecord(e1, e->stream);
    if (cudaEventSynchronize(e1) != cudaSuccess) rc = set_error(MDT_ERR_CUDA, "exchange bench failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  float ms = 0.f;
  if (rc == MDT_OK) cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  e->pub_expected = (e->pub_expected + (unsigned)iters + 1u) & 0x7fffffffu;   // every exchange kernel publishes once
  MDT_TRY(rc);
  *seconds = (double)ms * 1e-3 / iters;
  return MDT_OK;
}

int mdt_train_step_many(mdt_engine* const* engines, const float* const* xs, int n_trials, int batch) {
  MDT_CHECK(engines && xs && n_trials >= 0, "bad argument");
  // the trials share the GPU: each one's kernels are sized for 1/n of the SMs (Tc05Plan::sm_share)
  const int share = n_trials < 1 ? 1 : (n_trials > 16 ? 16 : n_trials);
  int rc = MDT_OK;
  for (int i = 0; i < n_trials && rc == MDT_OK; ++i) {
    mdt_engine* e = engines[i];
    MDT_CHECK(e != nullptr, "engines[%d] is NULL", i);
    const int keep = e->tc ? e->tc->sm_share : 1;
    if (e->tc && share > keep) e->tc->sm_share = share;
    rc = mdt_train_step(e, xs[i], batch, nullptr, nullptr);
    if (e->tc) e->tc->sm_share = keep;
  }
  return rc;
}

int mdt_train_step_host_submit(mdt_engine* e, const float* x_host, int batch, const float* eps_host_or_null,
                               uint32_t* ticket) {
  MDT_TRY(check_batch(e, batch));
  MDT_CHECK(x_host || batch == 0, "x_host is NULL");
  MDT_CHECK(ticket, "ticket is NULL");
  const float* x = x_host;
  const float* eps = eps_host_or_null;
  if (e->tc && e->mailbox != nullptr && batch > 0 && e->opt_zerocopy) {
    // Tensor-core path.  Pinned x is read by the step's prologue kernel straight from host memory (no separate H2D
    // copy in front of the graph), and the loss comes back through the mailbox ring the step's last thread writes.
    cudaPointerAttributes at{};
    const bool mapped = cudaPointerGetAttributes(&at, x_host) == cudaSuccess && at.type == cudaMemoryTypeHost &&
                        at.devicePointer != nullptr;
    cudaGetLastError();
    int slot = -1;
    const bool ahead_on = e->opt_pipeline;
    const unsigned int last_pub = e->pub_expected;
    const bool in_flight = last_pub != 0 && (unsigned int)(e->mailbox[last_pub & 63u] >> 32) != last_pub;
    if (mapped && in_flight && ahead_on && e->opt_h2d_prefetch && e->copy_stream != nullptr && e->tc->simt_latent) {
      // A step is still running: this step's prologue (pinned x, read in place over PCIe -> stage copy, x^T, bf16 copy)
      // goes to the copy stream and into the input set the running step is NOT using; the step itself is then
      // enqueued without a prologue kernel.  The transfer and the prologue hide behind the step in flight.
      Tc05Plan* p = e->tc;
      const int set = (e->host_set ^= 1);
      if (e->set_used[set]) MDT_CUDA(cudaStreamWaitEvent(e->copy_stream, e->ev_set_free[set], 0));
      p->x_src = static_cast<const float*>(at.devicePointer);
      int rc = tc05_prep(p, e->copy_stream, e->x_stage, batch, set, false, 0, "prep_xT_ahead", /*programmatic=*/false);
      p->x_src = nullptr;
      MDT_TRY(rc);
      MDT_CUDA(cudaEventRecord(e->ev_set_ready[set], e->copy_stream));
      MDT_CUDA(cudaStreamWaitEvent(e->stream, e->ev_set_ready[set], 0));
      x = e->x_stage;
      rc = stage_inputs(e, &x, batch, &eps, cudaMemcpyHostToDevice);      // eps only
      if (rc == MDT_OK) {
        p->pipelined = true; p->no_prefetch = true; p->xpar = set;
        rc = train_step_device(e, x, batch, eps);
        p->pipelined = false; p->no_prefetch = false;
      }
      MDT_TRY(rc);
      MDT_CUDA(cudaEventRecord(e->ev_set_free[set], e->stream));
      e->set_used[set] = true;
      *ticket = e->pub_expected;
      return MDT_OK;
    }
    if (mapped) {
      e->tc->x_src = static_cast<const float*>(at.devicePointer);
      x = e->x_stage;                      // stage_inputs then only handles eps
      // Is the step submitted last still running?  Then there is compute to hide a DMA behind.
      const unsigned int last = e->pub_expected;
      const bool busy = last != 0 && (unsigned int)(e->mailbox[last & 63u] >> 32) != last;
      if (busy && e->opt_h2d_prefetch && e->copy_stream != nullptr) {
        slot = (int)(e->pre_next++ % mdt_engine::kPreSlots);
        if (e->pre_used[slot]) MDT_CUDA(cudaStreamWaitEvent(e->copy_stream, e->ev_used[slot], 0));
        MDT_CUDA(cudaMemcpyAsync(e->x_pre[slot], x_host, (size_t)batch * e->D * sizeof(float), cudaMemcpyHostToDevice,
                                 e->copy_stream));
        MDT_CUDA(cudaEventRecord(e->ev_copied[slot], e->copy_stream));
        MDT_CUDA(cudaStreamWaitEvent(e->stream, e->ev_copied[slot], 0));
        e->tc->x_src = e->x_pre[slot];
      }
    }
    int rc = stage_inputs(e, &x, batch, &eps, cudaMemcpyHostToDevice);
    if (rc == MDT_OK) rc = train_step_device(e, x, batch, eps);
    e->tc->x_src = nullptr;
    MDT_TRY(rc);
    if (slot >= 0) {                       // the slot is free again once this step (its prologue, really) has run
      MDT_CUDA(cudaEventRecord(e->ev_used[slot], e->stream));
      e->pre_used[slot] = true;
    }
    if (e->ev_set_free[0] != nullptr) {    // this step worked on input set 0
      MDT_CUDA(cudaEventRecord(e->ev_set_free[0], e->stream));
      e->set_used[0] = true;
      e->host_set = 0;
    }
    *ticket = e->pub_expected;
    return MDT_OK;
  }
  // no mailbox (fp32 FFMA path, empty batch, MDT_NO_ZEROCOPY): the step runs to completion here
  MDT_TRY(stage_inputs(e, &x, batch, &eps, cudaMemcpyHostToDevice));
  if (batch == 0) MDT_CUDA(cudaMemsetAsync(e->loss_dev, 0, sizeof(float), e->stream));
  MDT_TRY(train_step_device(e, x, batch, eps));
  MDT_CUDA(cudaMemcpyAsync(e->loss_pinned, e->loss_dev, sizeof(float), cudaMemcpyDeviceToHost, e->stream));
  MDT_CUDA(cudaStreamSynchronize(e->stream));
  e->sync_seq = (e->sync_seq + 1u) | 0x80000000u;      // tickets of this kind have the top bit set
  e->sync_losses[e->sync_seq & 63u] = *e->loss_pinned;
  *ticket = e->sync_seq;
  return MDT_OK;
}

int mdt_train_step_host_wait(mdt_engine* e, uint32_t ticket, float* loss_host) {
  MDT_CHECK(e, "NULL engine");
  if (ticket & 0x80000000u) {
    if (loss_host) *loss_host = e->sync_losses[ticket & 63u];
    return MDT_OK;
  }
  MDT_CHECK(e->mailbox != nullptr, "no step is in flight on this engine");
  MDT_CHECK(((e->pub_expected - ticket) & 0x7fffffffu) < 64u, "ticket %u is more than 63 steps old (latest %u): its mailbox slot has been reused",
            ticket, e->pub_expected);
  volatile unsigned long long* slot = e->mailbox + (ticket & 63u);
  unsigned long long v = 0;
  for (unsigned int spins = 0;; ++spins) {
    v = *slot;
    if ((unsigned int)(v >> 32) == ticket) break;
    if ((spins & 255u) == 255u) {
      const cudaError_t q = cudaStreamQuery(e->stream);
      if (q == cudaSuccess) {              // stream idle: the word must be there by now
        v = *slot;
        if ((unsigned int)(v >> 32) == ticket) break;
        return set_error(MDT_ERR_STATE, "step %u finished without publishing its loss (mailbox holds %u)", ticket,
                         (unsigned int)(v >> 32));
      }
      if (q != cudaErrorNotReady) return set_error(MDT_ERR_CUDA, "training step failed: %s", cudaGetErrorString(q));
    }
  }
  if (loss_host) {
    const unsigned int bits = (unsigned int)(v & 0xffffffffull);
    memcpy(loss_host, &bits, sizeof(float));
  }
  return MDT_OK;
}

int mdt_train_step_host(mdt_engine* e, const float* x_host, int batch, const float* eps_host_or_null,
                        float* loss_host) {
  uint32_t ticket = 0;
  MDT_TRY(mdt_train_step_host_submit(e, x_host, batch, eps_host_or_null, &ticket));
  return mdt_train_step_host_wait(e, ticket, loss_host);
}

int mdt_train_steps_host_submit(mdt_engine* e, const float* x_host, int batch, int n_steps, uint32_t* first_ticket) {
  MDT_TRY(check_batch(e, batch));
  MDT_CHECK(x_host && first_ticket && batch > 0, "bad argument");
  MDT_CHECK(n_steps >= 1 && n_steps <= mdt_engine::kHostChunk, "n_steps %d outside [1, %d]", n_steps, mdt_engine::kHostChunk);
  if (e->tc == nullptr || e->mailbox == nullptr || e->copy_stream == nullptr)
    return set_error(MDT_ERR_UNSUPPORTED, "mdt_train_steps_host_submit needs a tensor-core gemm_mode (its losses come back through the mailbox)");
  constexpr int C = mdt_engine::kHostChunk;
  if (e->x_ring == nullptr) {     // first use: the ring (2 x 16 batches of max_batch rows) and its events
    MDT_CUDA(cudaMalloc(&e->x_ring, (size_t)2 * C * e->Bmax * e->D * sizeof(float)));
    for (int k = 0; k < 2; ++k) {
      MDT_CUDA(cudaEventCreateWithFlags(&e->ev_ring_copied[k], cudaEventDisableTiming));
      MDT_CUDA(cudaEventCreateWithFlags(&e->ev_ring_free[k], cudaEventDisableTiming));
    }
  }
  const int half = (int)(e->ring_next++ & 1u);
  const size_t bstride = (size_t)batch * e->D;
  float* dst = e->x_ring + (size_t)half * C * bstride;
  if (e->ring_used[half]) MDT_CUDA(cudaStreamWaitEvent(e->copy_stream, e->ev_ring_free[half], 0));
  MDT_CUDA(cudaMemcpyAsync(dst, x_host, (size_t)n_steps * bstride * sizeof(float), cudaMemcpyHostToDevice, e->copy_stream));
  MDT_CUDA(cudaEventRecord(e->ev_ring_copied[half], e->copy_stream));
  MDT_CUDA(cudaStreamWaitEvent(e->stream, e->ev_ring_copied[half], 0));
  // the ring is one pool of 2 x 16 batches of this batch size; the chunk's last step fetches nothing (what follows it
  // in the ring may still be on its way over PCIe)
  mdt_batch_source src{};
  src.x_pool = e->x_ring;
  src.n_batches = 2 * C;
  const unsigned int t0 = (e->pub_expected + 1u) & 0x7fffffffu;
  e->tail_no_prefetch = true;
  const int rc = mdt_train_steps(e, &src, batch, (int64_t)half * C, n_steps, nullptr);
  e->tail_no_prefetch = false;
  MDT_TRY(rc);
  MDT_CUDA(cudaEventRecord(e->ev_ring_free[half], e->stream));
  e->ring_used[half] = true;
  *first_ticket = t0;
  return MDT_OK;
}

int mdt_eval_step(mdt_engine* e, const float* x, int batch, const float* eps_or_null, float* loss_out, float* recon) {
  MDT_TRY(check_batch(e, batch));
  MDT_CHECK(x || batch == 0, "x is NULL");
  e->last_B = batch;
  const float* eps = eps_or_null;
  MDT_TRY(stage_inputs(e, &x, batch, &eps, cudaMemcpyDeviceToDevice));
  if (batch == 0) {
    MDT_CUDA(cudaMemsetAsync(e->loss_dev, 0, sizeof(float), e->stream));
  } else {
    MDT_TRY(run_cached(e, 5, batch, eps != nullptr, [&]() -> int {
      MDT_TRY(enqueue_forward(e, x, batch, eps, false));
      return enqueue_loss(e, x, batch, false);
    }));
  }
  if (loss_out) MDT_CUDA(cudaMemcpyAsync(loss_out, e->loss_dev, sizeof(float), cudaMemcpyDeviceToDevice, e->stream));
  if (recon && batch > 0)
    MDT_CUDA(cudaMemcpyAsync(recon, e->recon, (size_t)batch * e->D * sizeof(float), cudaMemcpyDeviceToDevice, e->stream));
  return MDT_OK;
}

int mdt_decode(mdt_engine* e, const float* z, int batch, float* recon) {
  MDT_TRY(check_batch(e, batch));
  MDT_CHECK(z && recon, "NULL argument");
  if (batch == 0) return MDT_OK;
  e->last_B = 0;  // activations no longer describe a full forward
  MDT_CUDA(cudaMemcpyAsync(e->z, z, (size_t)batch * e->Z * sizeof(float), cudaMemcpyDeviceToDevice, e->stream));
  MDT_TRY(enqueue_forward(e, nullptr, batch, nullptr, true));
  MDT_CUDA(cudaMemcpyAsync(recon, e->recon, (size_t)batch * e->D * sizeof(float), cudaMemcpyDeviceToDevice, e->stream));
  return MDT_OK;
}

int mdt_engine_workspace(mdt_engine* e, const char* name, float** ptr, int* ld) {
  MDT_CHECK(e && name && ptr, "NULL argument");
  struct { const char* n; float* p; int ld; } tab[] = {
      {"x", e->x_stage, e->D}, {"h1", e->h1, e->H}, {"ml", e->ml, 2 * e->Z}, {"eps", e->eps, e->Z}, {"std", e->sd, e->Z},
      {"z", e->z, e->Z}, {"h3", e->h3, e->H}, {"recon", e->recon, e->D}, {"da4", e->da4, e->D}, {"da3", e->da3, e->H},
      {"dz", e->dz, e->Z}, {"dml_kl", e->dml_kl, 2 * e->Z}, {"dml", e->dml, 2 * e->Z}, {"da1", e->da1, e->H},
      {"loss", e->loss_dev, 1},
  };
  for (auto& t : tab)
    if (strcmp(t.n, name) == 0) {
      *ptr = t.p;
      if (ld) *ld = t.ld;
      return MDT_OK;
    }
  return set_error(MDT_ERR_INVALID, "unknown workspace '%s'", name);
}

int mdt_engine_launch_count(mdt_engine* e, int64_t* kernels) {
  MDT_CHECK(e && kernels, "NULL argument");
  *kernels = e->launch_count;
  return MDT_OK;
}

int mdt_engine_profile_step(mdt_engine* e, const float* x, int batch, int iters, char* out, size_t out_len) {
  MDT_TRY(check_batch(e, batch));
  MDT_CHECK(x && out && out_len > 0 && iters > 0 && batch > 0, "bad argument");
  MDT_CHECK(e->gsize == 1, "profile_step times the single-GPU step only");
  const float* eps = nullptr;
  MDT_TRY(stage_inputs(e, &x, batch, &eps, cudaMemcpyDeviceToDevice));
  std::vector<const char*> names;
  std::vector<double> ms;
  e->prof_on = true;
  int rc = MDT_OK;
  for (int it = 0; it < iters && rc == MDT_OK; ++it) {
    e->prof_marks.clear();
    // keep the GPU busy while the host enqueues the whole step, so intervals are device time, not launch time
    spin_kernel<<<1, 1, 0, e->stream>>>(1000000);
    rc = note_launch(e, "_start");
    if (rc == MDT_OK) {   // calibration: an empty kernel, so the caller can subtract the launch + event overhead
      null_kernel<<<1, 32, 0, e->stream>>>();
      rc = note_launch(e, "_null");
    }
    if (rc == MDT_OK) rc = train_step_device(e, x, batch, nullptr);
    if (rc == MDT_OK && cudaStreamSynchronize(e->stream) != cudaSuccess) rc = set_error(MDT_ERR_CUDA, "sync failed in profile_step");
    if (rc == MDT_OK) {
      if (names.empty()) {
        for (size_t i = 1; i < e->prof_marks.size(); ++i) names.push_back(e->prof_marks[i].first);
        ms.assign(names.size(), 0.0);
      }
      for (size_t i = 1; i < e->prof_marks.size() && i - 1 < ms.size(); ++i) {
        float t = 0.f;
        cudaEventElapsedTime(&t, e->prof_marks[i - 1].second, e->prof_marks[i].second);
        ms[i - 1] += t;
      }
    }
    for (auto& m : e->prof_marks) cudaEventDestroy(m.second);
    e->prof_marks.clear();
  }
  e->prof_on = false;
  MDT_TRY(rc);
  size_t pos = 0;
  out[0] = 0;
  for (size_t i = 0; i < names.size(); ++i) {
    int w = snprintf(out + pos, out_len - pos, "%s %.6f\n", names[i], ms[i] / iters);
    if (w < 0 || (size_t)w >= out_len - pos) return set_error(MDT_ERR_INVALID, "profile buffer too small");
    pos += (size_t)w;
  }
  return MDT_OK;
}

}  // extern "C" (re-opened below)

// =================================================================== grouped trials
// T independent trials of the same shape on ONE GPU, stepped in lock-step with ONE launch per layer: every kernel of
// the step is launched once over a table of the T trials' tensor maps / arguments (grid.y resp. grid.x = trial x
// tiles).  The step's enqueue code is reused unchanged: it is run once per trial under a launch recorder, the T records
// are checked for congruence, and one of them is replayed with grouped launches inside a CUDA graph capture.
struct mdt_group {
  std::vector<mdt_engine*> eng;
  struct Graph { cudaGraphExec_t exec = nullptr; void* tables = nullptr; long long kernels = 0; };
  std::map<std::tuple<int, int, int>, Graph> graphs;      // (batch, steps, flags)
  const void* src_sig[2] = {nullptr, nullptr};
};

namespace {

template <int kEpi>
int group_gemm_launch_e(bool split, const Tc05GroupEntry* table, int tiles_m, int T, const TcOp& op, int cluster) {
  using SS = Tc05Smem<32, false>;
  using ST = Tc05Smem<32, true>;
  const dim3 grid((op.N + 31) / 32, tiles_m * T, op.nsplit);
  const size_t smem = split ? (size_t)ST::TOTAL : (size_t)SS::TOTAL;
  const dim3 block(tc_threads(split));
  cudaError_t ce;
  if (split) {
    static bool attr = false;
    if (!attr) { MDT_CUDA(cudaFuncSetAttribute(tc05_gemm_group_kernel<32, true, kEpi>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST::TOTAL)); attr = true; }
    ce = cluster ? launch_pdl_cluster(tc05_gemm_group_kernel<32, true, kEpi>, grid, block, smem, op.stream, (unsigned)op.nsplit, table, tiles_m, op.M, op.N, op.K)
                 : launch_pdl(tc05_gemm_group_kernel<32, true, kEpi>, grid, block, smem, op.stream, table, tiles_m, op.M, op.N, op.K);
  } else {
    static bool attr = false;
    if (!attr) { MDT_CUDA(cudaFuncSetAttribute(tc05_gemm_group_kernel<32, false, kEpi>, cudaFuncAttributeMaxDynamicSharedMemorySize, SS::TOTAL)); attr = true; }
    ce = cluster ? launch_pdl_cluster(tc05_gemm_group_kernel<32, false, kEpi>, grid, block, smem, op.stream, (unsigned)op.nsplit, table, tiles_m, op.M, op.N, op.K)
                 : launch_pdl(tc05_gemm_group_kernel<32, false, kEpi>, grid, block, smem, op.stream, table, tiles_m, op.M, op.N, op.K);
  }
  MDT_CUDA(ce);
  return MDT_OK;
}

int group_gemm_launch(const Tc05GroupEntry* table, int tiles_m, int T, const TcOp& op, int cluster) {
  switch (op.epi) {
    case EPI_BIAS_ACT: return group_gemm_launch_e<EPI_BIAS_ACT>(op.split != 0, table, tiles_m, T, op, cluster);
    case EPI_SIGMOID_BCE: return group_gemm_launch_e<EPI_SIGMOID_BCE>(op.split != 0, table, tiles_m, T, op, cluster);
    case EPI_RELU_MASK: return group_gemm_launch_e<EPI_RELU_MASK>(op.split != 0, table, tiles_m, T, op, cluster);
    case EPI_GRADW: return group_gemm_launch_e<EPI_GRADW>(op.split != 0, table, tiles_m, T, op, cluster);
    default: return set_error(MDT_ERR_UNSUPPORTED, "grouped launch: epilogue %d is not part of the fused step", op.epi);
  }
}

template <class K, class A>
int group_latent_launch(K kernel, const A* table, int bpt, int T, size_t smem, cudaStream_t s) {
  MDT_CHECK(smem <= 48 * 1024, "grouped latent kernel: %zu bytes of shared memory", smem);
  MDT_CUDA(launch_pdl(kernel, dim3(bpt * T), dim3(kLatThreads), smem, s, table, bpt));
  return MDT_OK;
}

int group_lat_fwd(const LatentFwdArgs* table, const TcOp& op, int T) {
  const int bpt = latent_grid(op.lf.B, op.sc);
  const size_t smem = (size_t)op.sc * op.lf.H * sizeof(float);
  const bool z5 = op.lf.Z <= 20;
  if (op.sc == 1) return z5 ? group_latent_launch(latent_fwd_group_kernel<1, 5>, table, bpt, T, smem, op.stream) : group_latent_launch(latent_fwd_group_kernel<1, 8>, table, bpt, T, smem, op.stream);
  if (op.sc == 2) return z5 ? group_latent_launch(latent_fwd_group_kernel<2, 5>, table, bpt, T, smem, op.stream) : group_latent_launch(latent_fwd_group_kernel<2, 8>, table, bpt, T, smem, op.stream);
  return z5 ? group_latent_launch(latent_fwd_group_kernel<4, 5>, table, bpt, T, smem, op.stream) : group_latent_launch(latent_fwd_group_kernel<4, 8>, table, bpt, T, smem, op.stream);
}
int group_lat_bwd(const LatentBwdArgs* table, const TcOp& op, int T) {
  const int bpt = latent_grid(op.lb.B, op.sc);
  const size_t smem = (size_t)op.sc * op.lb.H * sizeof(float);
  const bool z5 = op.lb.Z <= 20;
  if (op.sc == 1) return z5 ? group_latent_launch(latent_bwd_group_kernel<1, 5>, table, bpt, T, smem, op.stream) : group_latent_launch(latent_bwd_group_kernel<1, 8>, table, bpt, T, smem, op.stream);
  if (op.sc == 2) return z5 ? group_latent_launch(latent_bwd_group_kernel<2, 5>, table, bpt, T, smem, op.stream) : group_latent_launch(latent_bwd_group_kernel<2, 8>, table, bpt, T, smem, op.stream);
  return z5 ? group_latent_launch(latent_bwd_group_kernel<4, 5>, table, bpt, T, smem, op.stream) : group_latent_launch(latent_bwd_group_kernel<4, 8>, table, bpt, T, smem, op.stream);
}

// One trial's record of `steps` pipelined fused-Adam steps (what train_step_device enqueues for a single-rank group).
int group_record(mdt_engine* e, int B, int steps, bool tail_no_prefetch, TcRecorder* rec) {
  Tc05Plan* p = e->tc;
  tc_recorder() = rec;
  int rc = MDT_OK;
  for (int i = 0; i < steps && rc == MDT_OK; ++i) {
    const bool saved = p->no_prefetch;
    if (tail_no_prefetch && i == steps - 1) p->no_prefetch = true;
    rc = tc05_forward(p, e->stream, e->x_stage, B, nullptr, false, /*fuse_loss=*/true);
    if (rc == MDT_OK) rc = tc05_backward_fused_adam(p, e->stream, B);
    p->no_prefetch = saved;
  }
  tc_recorder() = nullptr;
  return rc;
}

size_t group_table_bytes(const TcOp& op, int T) {
  size_t one = 0;
  switch (op.kind) {
    case TcOp::GEMM: one = sizeof(Tc05GroupEntry); break;
    case TcOp::LAT_FWD: one = sizeof(LatentFwdArgs); break;
    case TcOp::LAT_BWD: one = sizeof(LatentBwdArgs); break;
    case TcOp::PREP: one = sizeof(PrepArgs); break;
    default: return 0;
  }
  return align_up(one * T, 256);
}

}  // namespace

extern "C" {

int mdt_group_create(mdt_engine* const* engines, int n_trials, mdt_group** out) {
  MDT_CHECK(engines && out && n_trials >= 1 && n_trials <= 64, "bad argument");
  mdt_engine* e0 = engines[0];
  for (int t = 0; t < n_trials; ++t) {
    mdt_engine* e = engines[t];
    MDT_CHECK(e != nullptr && e->bound && e->tc != nullptr, "trial %d: engine missing, unbound or not on the tensor-core path", t);
    MDT_CHECK(e->gsize == 1 && !e->use_sgd && e->opt_fused_adam, "trial %d: grouped launches cover single-rank Adam trials", t);
    MDT_CHECK(e->D == e0->D && e->H == e0->H && e->Z == e0->Z && e->Bmax == e0->Bmax && e->cfg.gemm_mode == e0->cfg.gemm_mode &&
                  e->device == e0->device, "trial %d differs from trial 0 in shape, batch limit, gemm_mode or device", t);
    MDT_CHECK(e->tc->simt_latent && e->tc->bn_h == 32 && e->tc->bn_fc4 == 32 && e->tc->bn_dwa == 32,
              "grouped launches cover the small-model regime (CUDA-core latent path, 32-wide tiles: hidden < 2048, max_batch <= 512)");
  }
  mdt_group* g = new mdt_group();
  g->eng.assign(engines, engines + n_trials);
  *out = g;
  return MDT_OK;
}

void mdt_group_destroy(mdt_group* g) {
  if (!g) return;
  if (!g->eng.empty()) cudaStreamSynchronize(g->eng[0]->stream);
  for (auto& kv : g->graphs) { cudaGraphExecDestroy(kv.second.exec); if (kv.second.tables) cudaFree(kv.second.tables); }
  delete g;
}

int mdt_group_train_steps(mdt_group* g, const mdt_batch_source* srcs, int batch, int64_t first_batch, int n_steps) {
  MDT_CHECK(g && srcs && batch > 0 && n_steps >= 0 && first_batch >= 0, "bad argument");
  const int T = (int)g->eng.size();
  mdt_engine* e0 = g->eng[0];
  cudaStream_t s0 = e0->stream;
  if (n_steps == 0) return MDT_OK;
  for (int t = 0; t < T; ++t) {
    MDT_TRY(check_batch(g->eng[t], batch));
    MDT_CHECK(srcs[t].n_batches > 0 && srcs[t].n_batches == srcs[0].n_batches, "every trial's source must hold the same number of batches");
    MDT_CHECK(srcs[t].x_pool != nullptr || (srcs[t].images_u8 != nullptr && srcs[t].indices != nullptr && srcs[t].n_images > 0),
              "trial %d: the batch source needs x_pool, or images_u8 + indices + n_images", t);
  }
  // the trials' other streams must not be mid-flight: everything of the group runs on trial 0's streams
  for (int t = 1; t < T; ++t) MDT_CUDA(cudaStreamSynchronize(g->eng[t]->stream));
  // sources are baked into the captured tables: a new set of sources drops the cached graphs
  const void* sig[2] = {srcs[0].x_pool ? (const void*)srcs[0].x_pool : (const void*)srcs[0].images_u8,
                        srcs[T - 1].x_pool ? (const void*)srcs[T - 1].x_pool : (const void*)srcs[T - 1].images_u8};
  if (memcmp(sig, g->src_sig, sizeof(sig)) != 0) {
    MDT_CUDA(cudaStreamSynchronize(s0));
    for (auto& kv : g->graphs) { cudaGraphExecDestroy(kv.second.exec); if (kv.second.tables) cudaFree(kv.second.tables); }
    g->graphs.clear();
    memcpy(g->src_sig, sig, sizeof(sig));
    for (mdt_engine* e : g->eng) e->primed = false;
  }
  // per-trial device state: batch cursor where the call starts, first batch staged
  for (int t = 0; t < T; ++t) {
    mdt_engine* e = g->eng[t];
    Tc05Plan* p = e->tc;
    p->sm_share = T;
    p->src_pool = srcs[t].x_pool; p->src_images = srcs[t].x_pool ? nullptr : srcs[t].images_u8;
    p->src_indices = reinterpret_cast<const long long*>(srcs[t].indices);
    p->src_n_batches = srcs[t].n_batches; p->src_n_images = srcs[t].n_images;
    if (e->cursor_expected < 0 || (e->cursor_expected - first_batch) % srcs[t].n_batches != 0) {
      const long long c = first_batch;
      MDT_CUDA(cudaMemcpy(&e->st->x_cursor, &c, sizeof(c), cudaMemcpyHostToDevice));
    }
    const bool primed = e->primed && e->primed_cursor >= 0 && (e->primed_cursor - first_batch) % srcs[t].n_batches == 0 &&
                        p->xpar == g->eng[0]->tc->xpar;
    if (!primed) {
      p->xpar = g->eng[0]->tc->xpar;
      MDT_TRY(tc05_prep(p, s0, e->x_stage, batch, p->xpar, false, 0, "prep_xT"));      // on trial 0's stream: ordered before the graphs
    }
    p->w3par = (t == 0) ? p->w3par : p->w3par;
    e->last_B = batch;
  }
  // the W3^T parity must agree across the trials (they are stepped in lock-step from here on); bring stragglers in line
  for (int t = 1; t < T; ++t)
    if (g->eng[t]->tc->w3par != e0->tc->w3par) {
      g->eng[t]->tc->w3par = e0->tc->w3par;
      MDT_TRY(tc05_weight_shadows(g->eng[t]->tc, s0));
    }
  int rc = MDT_OK;
  const int U = 16;
  for (int done = 0; done < n_steps && rc == MDT_OK;) {
    const int u = (n_steps - done >= U) ? U : 1;
    const int x0 = e0->tc->xpar, w0 = e0->tc->w3par;
    const auto key = std::make_tuple(batch, u, x0 | (w0 << 1));
    auto it = g->graphs.find(key);
    if (it == g->graphs.end()) {
      // 1. record the u steps once per trial
      std::vector<TcRecorder> recs(T);
      for (int t = 0; t < T && rc == MDT_OK; ++t) {
        Tc05Plan* p = g->eng[t]->tc;
        p->xpar = x0; p->w3par = w0; p->pipelined = true; p->prefetch_pending = false;
        rc = group_record(g->eng[t], batch, u, false, &recs[t]);
        p->pipelined = false;
      }
      if (rc != MDT_OK) break;
      const std::vector<TcOp>& ops = recs[0].ops;
      size_t bytes = 0;
      for (int t = 1; t < T; ++t)
        if (recs[t].ops.size() != ops.size()) { rc = set_error(MDT_ERR_STATE, "grouped launch: trial %d recorded %zu operations, trial 0 %zu", t, recs[t].ops.size(), ops.size()); break; }
      if (rc != MDT_OK) break;
      for (const TcOp& op : ops) bytes += group_table_bytes(op, T);
      // 2. build the tables (host), upload once
      std::vector<char> host(bytes);
      void* dev = nullptr;
      MDT_CUDA(cudaMalloc(&dev, bytes ? bytes : 256));
      std::vector<size_t> off(ops.size(), 0);
      size_t cur = 0;
      for (size_t i = 0; i < ops.size() && rc == MDT_OK; ++i) {
        const TcOp& op = ops[i];
        off[i] = cur;
        for (int t = 0; t < T; ++t) {
          const TcOp& ot = recs[t].ops[i];
          const bool same = ot.kind == op.kind && ot.bn == op.bn && ot.split == op.split && ot.epi == op.epi && ot.M == op.M && ot.N == op.N &&
                            ot.K == op.K && ot.nsplit == op.nsplit && ot.sc == op.sc && ot.blocks == op.blocks;
          if (!same) { rc = set_error(MDT_ERR_STATE, "grouped launch: operation %zu of trial %d is not congruent with trial 0's", i, t); break; }
          char* dst = host.data() + cur;
          if (op.kind == TcOp::GEMM) {
            if (op.bn != 32) { rc = set_error(MDT_ERR_UNSUPPORTED, "grouped launch: BLOCK_N %d", op.bn); break; }
            Tc05GroupEntry en{};
            en.a = ot.a; en.b = ot.b; en.ep = ot.ep; en.ep.dbg = nullptr;
            memcpy(dst + (size_t)t * sizeof(Tc05GroupEntry), &en, sizeof(en));
          } else if (op.kind == TcOp::LAT_FWD) {
            memcpy(dst + (size_t)t * sizeof(LatentFwdArgs), &ot.lf, sizeof(LatentFwdArgs));
          } else if (op.kind == TcOp::LAT_BWD) {
            memcpy(dst + (size_t)t * sizeof(LatentBwdArgs), &ot.lb, sizeof(LatentBwdArgs));
          } else if (op.kind == TcOp::PREP) {
            memcpy(dst + (size_t)t * sizeof(PrepArgs), &ot.pa, sizeof(PrepArgs));
          }
        }
        cur += group_table_bytes(op, T);
      }
      // cluster form of a split GEMM: decided for the whole (T x tiles) grid, written into every trial's entry
      std::vector<int> cluster(ops.size(), 0);
      for (size_t i = 0; i < ops.size() && rc == MDT_OK; ++i) {
        const TcOp& op = ops[i];
        if (op.kind != TcOp::GEMM) continue;
        cluster[i] = (op.nsplit > 1 && op.nsplit <= 8 && op.use_clusters) ? 1 : 0;
        const int tiles = ((op.N + 31) / 32) * ((op.M + kTcBM - 1) / kTcBM) * T;
        if (cluster[i] && tiles * op.nsplit > e0->dev->sm_count) cluster[i] = 0;      // clusters must all be co-resident
        for (int t = 0; t < T; ++t)
          reinterpret_cast<Tc05GroupEntry*>(host.data() + off[i])[t].ep.cluster = cluster[i];
      }
      if (rc == MDT_OK && bytes) {
        cudaError_t ce = cudaMemcpy(dev, host.data(), bytes, cudaMemcpyHostToDevice);
        if (ce != cudaSuccess) rc = set_error(MDT_ERR_CUDA, "grouped launch: table upload: %s", cudaGetErrorString(ce));
      }
      if (rc != MDT_OK) { cudaFree(dev); break; }
      // 3. replay trial 0's record with grouped launches inside a capture
      cudaGraph_t graph = nullptr;
      long long kernels = 0;
      MDT_CUDA(cudaStreamBeginCapture(s0, cudaStreamCaptureModeRelaxed));
      for (size_t i = 0; i < ops.size() && rc == MDT_OK; ++i) {
        const TcOp& op = ops[i];
        char* tab = static_cast<char*>(dev) + off[i];
        cudaError_t ce = cudaSuccess;
        switch (op.kind) {
          case TcOp::EV_RECORD: ce = cudaEventRecord(op.ev, op.stream); break;
          case TcOp::STREAM_WAIT: ce = cudaStreamWaitEvent(op.stream, op.ev, 0); break;
          case TcOp::GEMM:
            rc = group_gemm_launch(reinterpret_cast<const Tc05GroupEntry*>(tab), (op.M + kTcBM - 1) / kTcBM, T, op, cluster[i]);
            ++kernels;
            break;
          case TcOp::LAT_FWD: rc = group_lat_fwd(reinterpret_cast<const LatentFwdArgs*>(tab), op, T); ++kernels; break;
          case TcOp::LAT_BWD: rc = group_lat_bwd(reinterpret_cast<const LatentBwdArgs*>(tab), op, T); ++kernels; break;
          case TcOp::PREP:
            if (op.programmatic) ce = launch_pdl(prep_group_kernel, dim3(op.blocks * T), dim3(256), 0, op.stream, reinterpret_cast<const PrepArgs*>(tab), op.blocks);
            else prep_group_kernel<<<dim3(op.blocks * T), dim3(256), 0, op.stream>>>(reinterpret_cast<const PrepArgs*>(tab), op.blocks);
            if (ce == cudaSuccess) ce = cudaGetLastError();
            ++kernels;
            break;
        }
        if (ce != cudaSuccess) rc = set_error(MDT_ERR_CUDA, "grouped launch: %s", cudaGetErrorString(ce));
      }
      cudaError_t ce = cudaStreamEndCapture(s0, &graph);
      if (rc == MDT_OK && ce != cudaSuccess) rc = set_error(MDT_ERR_CUDA, "grouped launch: cudaStreamEndCapture: %s", cudaGetErrorString(ce));
      if (rc == MDT_OK) {
        int removed = 0;
        rc = reduce_graph_edges(graph, &removed);
      }
      mdt_group::Graph entry;
      if (rc == MDT_OK && cudaGraphInstantiate(&entry.exec, graph, 0) != cudaSuccess) rc = set_error(MDT_ERR_CUDA, "grouped launch: cudaGraphInstantiate failed");
      if (graph) cudaGraphDestroy(graph);
      if (rc != MDT_OK) { cudaFree(dev); break; }
      entry.tables = dev; entry.kernels = kernels;
      it = g->graphs.emplace(key, entry).first;
    }
    MDT_CUDA(cudaGraphLaunch(it->second.exec, s0));
    // host-side bookkeeping of every trial, as if it had been stepped alone
    for (int t = 0; t < T; ++t) {
      mdt_engine* e = g->eng[t];
      e->launch_count += it->second.kernels;
      e->pub_expected = (e->pub_expected + (unsigned)u) & 0x7fffffffu;
      e->tc->xpar = x0 ^ (u & 1); e->tc->w3par = w0 ^ (u & 1); e->tc->prefetch_pending = false;
    }
    done += u;
  }
  for (int t = 0; t < T; ++t) {
    mdt_engine* e = g->eng[t];
    Tc05Plan* p = e->tc;
    p->sm_share = e->sm_share_pending;
    p->src_pool = nullptr; p->src_images = nullptr; p->src_indices = nullptr;
    e->cursor_expected = rc == MDT_OK ? first_batch + n_steps : -1;
    e->primed = rc == MDT_OK;
    e->primed_cursor = first_batch + n_steps;
    // the step graphs of the engine itself were captured for other SM shares / streams: nothing to invalidate, they are keyed
  }
  // the trials' own streams must see the group's work as done before they are used again
  if (rc == MDT_OK && T > 1) {
    cudaEvent_t ev = e0->ev_set_free[0] ? e0->ev_set_free[0] : nullptr;
    if (ev != nullptr) {
      MDT_CUDA(cudaEventRecord(ev, s0));
      for (int t = 1; t < T; ++t) MDT_CUDA(cudaStreamWaitEvent(g->eng[t]->stream, ev, 0));
    } else {
      MDT_CUDA(cudaStreamSynchronize(s0));
    }
  }
  return rc;
}

}  // extern "C"

extern "C" {

// ------------------------------------------------------------------ stand-alone kernels

int mdt_adam_flat(float* p, const float* g, float* m, float* v, size_t n, const mdt_adam* h, int64_t step,
                  float grad_scale, void* stream) {
  MDT_CHECK(p && g && m && v && h, "NULL argument");
  MDT_CHECK(step >= 1, "step is 1-based");
  for (const void* q : {(const void*)p, (const void*)g, (const void*)m, (const void*)v})
    MDT_CHECK(reinterpret_cast<uintptr_t>(q) % 16 == 0, "buffers must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  MDT_TRY(device_info(&di));
  if (n == 0) return MDT_OK;
  AdamArgs a{};
  a.p = p; a.g = g; a.m = m; a.v = v; a.n = (long long)n; a.st = nullptr; a.grad_scale = grad_scale;
  const double bc1 = 1.0 - pow(h->beta1, (double)step), bc2 = 1.0 - pow(h->beta2, (double)step);
  a.step_size = (float)(h->lr / bc1);
  a.bc2_sqrt = (float)sqrt(bc2);
  a.one_minus_b1 = (float)(1.0 - h->beta1);
  a.b2 = (float)h->beta2;
  a.one_minus_b2 = (float)(1.0 - h->beta2);
  a.eps = (float)h->eps;
  adam_flat_kernel<<<adam_grid(a.n, di->sm_count), kAdamThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
  MDT_CUDA(cudaGetLastError());
  return MDT_OK;
}

int mdt_reparam(const float* mu, const float* logvar, int ld_ml, const float* eps_in, uint64_t seed, uint64_t offset,
                int batch, int latent, float* z, int ld_z, float* eps_out, void* stream) {
  MDT_CHECK(mu && logvar && z, "NULL argument");
  MDT_CHECK(batch >= 0 && latent > 0 && ld_ml >= latent && ld_z >= latent, "bad shape");
  MDT_CHECK(offset % 4 == 0, "philox offset must be a multiple of 4");
  DeviceInfo* di = nullptr;
  MDT_TRY(device_info(&di));
  if (batch == 0) return MDT_OK;
  ReparamArgs a{};
  a.mu = mu; a.logvar = logvar; a.ld_ml = ld_ml; a.eps_in = eps_in; a.st = nullptr; a.seed = seed; a.offset = offset;
  torch_randn_policy((long long)batch * latent, di->sm_count, di->max_threads_sm, &a.S, &a.offset_inc);
  a.B = batch; a.Z = latent; a.z = z; a.ld_z = ld_z; a.eps_out = eps_out; a.std_out = nullptr;
  long long blocks = ((long long)batch * latent + 255) / 256;
  if (blocks > di->sm_count * 8) blocks = di->sm_count * 8;
  reparam_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a);
  MDT_CUDA(cudaGetLastError());
  return MDT_OK;
}

int mdt_loss_bce_kl(const float* recon, const float* x, const float* mu, const float* logvar, int batch, int in_dim,
                    int latent, float* loss_out, float* d_recon, float* d_mu, float* d_logvar, void* stream) {
  MDT_CHECK(recon && x && mu && logvar && loss_out, "NULL argument");
  MDT_CHECK(batch >= 0 && in_dim > 0 && latent > 0, "bad shape");
  for (const void* q : {(const void*)recon, (const void*)x, (const void*)d_recon})
    MDT_CHECK(reinterpret_cast<uintptr_t>(q) % 16 == 0, "recon/x/d_recon must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  MDT_TRY(device_info(&di));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (di->loss_partials == nullptr) {
    MDT_CUDA(cudaMalloc(&di->loss_partials, kLossMaxBlocks * sizeof(float)));
    MDT_CUDA(cudaMalloc(&di->loss_ticket, 256));
    MDT_CUDA(cudaMemset(di->loss_ticket, 0, 256));
  }
  LossArgs a{};
  a.recon = recon; a.x = x; a.mu = mu; a.logvar = logvar; a.ld_ml = latent; a.B = batch; a.D = in_dim; a.Z = latent;
  a.d_recon = d_recon; a.d_mu = d_mu; a.d_logvar = d_logvar; a.ld_dml = latent;
  a.partials = di->loss_partials; a.ticket = di->loss_ticket; a.loss_out = loss_out;
  loss_bce_kl_kernel<false><<<loss_grid((long long)batch * in_dim), kLossThreads, 0, s>>>(a);
  MDT_CUDA(cudaGetLastError());
  return MDT_OK;
}

}  // extern "C" (re-opened below)

// ---- allocation-free variant of the loss kernel entry ------------------------------------------
// Its scratch (one partial sum per block + the "last block" ticket) is this fixed array in the
// library image instead of a cudaMalloc'ed buffer.  Every GPU that loads the library carries these
// 4352 bytes until the process exits.  In HOST code the name `g_loss_static_scratch` is only a
// stand-in for the symbol, not a device address: the address is looked up once per device with
// cudaGetSymbolAddress and that pointer is what the kernel receives.
constexpr int kNoAllocMaxBlocks = 1024;                                     // partial sums available
constexpr long long kNoAllocMaxElems = (long long)kNoAllocMaxBlocks * kLossThreads * 4;  // 1,048,576 recon elements
constexpr int kNoAllocTicketFloats = 64;                                    // ticket word lives in its own 256 bytes
static __device__ __align__(256) float g_loss_static_scratch[kNoAllocTicketFloats + kNoAllocMaxBlocks];
static float* g_loss_static_ptr[64] = {nullptr};                            // per device, filled at first use

extern "C" {

int mdt_loss_bce_kl_noalloc(const float* recon, const float* x, const float* mu, const float* logvar, int batch,
                            int in_dim, int latent, float* loss_out, float* d_recon, float* d_mu, float* d_logvar,
                            void* stream) {
  MDT_CHECK(recon && x && mu && logvar && loss_out, "NULL argument");
  MDT_CHECK(batch >= 0 && in_dim > 0 && latent > 0, "bad shape");
  for (const void* q : {(const void*)recon, (const void*)x, (const void*)d_recon})
    MDT_CHECK(reinterpret_cast<uintptr_t>(q) % 16 == 0, "recon/x/d_recon must be 16-byte aligned");
  const long long elems = (long long)batch * in_dim;
  MDT_CHECK(elems <= kNoAllocMaxElems,
            "mdt_loss_bce_kl_noalloc: batch*in_dim = %lld exceeds the fixed scratch capacity of %lld elements "
            "(%d blocks); use mdt_loss_bce_kl for larger inputs",
            elems, kNoAllocMaxElems, kNoAllocMaxBlocks);
  int dev = 0;
  MDT_CUDA(cudaGetDevice(&dev));
  MDT_CHECK(dev >= 0 && dev < 64, "device index %d out of range", dev);
  if (g_loss_static_ptr[dev] == nullptr) {
    // one-time set-up for this device (it is the current one): resolve the symbol to a device address
    float* p = nullptr;
    MDT_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_loss_static_scratch));
    g_loss_static_ptr[dev] = p;
  }
  float* scratch = g_loss_static_ptr[dev];
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  // the array keeps whatever the previous call left; only the ticket has to start from zero
  MDT_CUDA(cudaMemsetAsync(scratch, 0, sizeof(unsigned int), s));
  LossArgs a{};
  a.recon = recon; a.x = x; a.mu = mu; a.logvar = logvar; a.ld_ml = latent; a.B = batch; a.D = in_dim; a.Z = latent;
  a.d_recon = d_recon; a.d_mu = d_mu; a.d_logvar = d_logvar; a.ld_dml = latent;
  a.partials = scratch + kNoAllocTicketFloats;
  a.ticket = reinterpret_cast<unsigned int*>(scratch);
  a.loss_out = loss_out;
  const int grid = loss_grid(elems);   // <= kNoAllocMaxBlocks because of the size check above
  loss_bce_kl_kernel<false><<<grid, kLossThreads, 0, s>>>(a);
  MDT_CUDA(cudaGetLastError());
  return MDT_OK;
}

// Debug aid (not part of the public header): copies the 16 phase-timing stamps of the latest tcgen05 launch.
int mdt_debug_tc05_stamps(long long* out512, int reset_index) {
  long long* buf = tc05_dbg_buffer();
  if (!buf || !out512) return set_error(MDT_ERR_STATE, "set MDT_TC05_DBG=1 before the first tcgen05 launch");
  MDT_CUDA(cudaDeviceSynchronize());
  MDT_CUDA(cudaMemcpy(out512, buf, 512 * sizeof(long long), cudaMemcpyDeviceToHost));
  if (reset_index) tc05_dbg_launch_index() = 0;
  // re-arm: zeros everywhere, +inf in the atomicMin slot (13) of every launch slot
  static long long init[512];
  for (int i = 0; i < 512; ++i) init[i] = (i % 16 == 13) ? 0x7fffffffffffffffLL : 0;
  MDT_CUDA(cudaMemcpy(buf, init, sizeof(init), cudaMemcpyHostToDevice));
  return MDT_OK;
}

int mdt_gather_u8(const uint8_t* images, int64_t n_images, const int64_t* indices, int batch, int in_dim, float* x_out,
                  void* stream) {
  MDT_CHECK(images && indices && x_out, "NULL argument");
  MDT_CHECK(batch >= 0 && in_dim > 0 && n_images > 0, "bad shape");
  MDT_CHECK(in_dim % 16 == 0, "in_dim must be a multiple of 16 (16-byte pixel rows), got %d", in_dim);
  MDT_CHECK(reinterpret_cast<uintptr_t>(images) % 16 == 0 && reinterpret_cast<uintptr_t>(x_out) % 16 == 0,
            "images and x_out must be 16-byte aligned");
  DeviceInfo* di = nullptr;
  MDT_TRY(device_info(&di));
  if (batch == 0) return MDT_OK;
  long long blocks = ((long long)batch * (in_dim / 16) + 255) / 256;
  if (blocks > di->sm_count * 8) blocks = di->sm_count * 8;
  gather_u8_kernel<<<(int)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      images, reinterpret_cast<const long long*>(indices), batch, in_dim, (long long)n_images, x_out);
  MDT_CUDA(cudaGetLastError());
  return MDT_OK;
}

int mdt_linear(const float* x, const float* w, const float* b, int M, int N, int K, int act, int gemm_mode, float* y,
               void* stream) {
  MDT_CHECK(x && w && y, "NULL argument");
  MDT_CHECK(M >= 0 && N > 0 && K > 0, "bad shape");
  MDT_CHECK(act >= MDT_ACT_NONE && act <= MDT_ACT_SIGMOID, "bad act");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (M == 0) return MDT_OK;
  if (gemm_mode == MDT_GEMM_FP32) {
    MDT_CUDA((launch_gemm_simt<true, true>(x, w, M, N, K, K, K, EpiBiasAct{y, N, b, act}, s)));
    return MDT_OK;
  }
  if (gemm_mode == MDT_GEMM_BF16)
    return set_error(MDT_ERR_UNSUPPORTED, "mdt_linear takes fp32 operands: MDT_GEMM_BF16 exists inside the engine only (bf16 operand mirrors)");
  MDT_CHECK(gemm_mode == MDT_GEMM_TF32 || gemm_mode == MDT_GEMM_TF32X3, "unknown gemm_mode %d", gemm_mode);
  DeviceInfo* di = nullptr;
  MDT_TRY(device_info(&di));
  if (di->cc_major != 10) return set_error(MDT_ERR_UNSUPPORTED, "tcgen05 GEMM needs an sm_100 device");
  return tc05_linear(x, w, b, M, N, K, act, gemm_mode == MDT_GEMM_TF32X3, y, s);
}

int mdt_allreduce_bench(void* nccl_comm, void* buf, size_t bytes, int iters, void* stream, double* seconds) {
  MDT_CHECK(nccl_comm && buf && seconds && iters > 0 && bytes % 4 == 0, "bad argument");
  MDT_TRY(load_nccl());
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ncclComm_t comm = reinterpret_cast<ncclComm_t>(nccl_comm);
  cudaEvent_t e0, e1;
  MDT_CUDA(cudaEventCreate(&e0));
  MDT_CUDA(cudaEventCreate(&e1));
  MDT_NCCL(g_nccl.AllReduce(buf, buf, bytes / 4, ncclFloat, ncclSum, comm, s));  // warm-up
  MDT_CUDA(cudaEventRecord(e0, s));
  for (int i = 0; i < iters; ++i) MDT_NCCL(g_nccl.AllReduce(buf, buf, bytes / 4, ncclFloat, ncclSum, comm, s));
  MDT_CUDA(cudaEventRecord(e1, s));
  MDT_CUDA(cudaEventSynchronize(e1));
  float ms = 0.f;
  MDT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *seconds = (double)ms * 1e-3 / iters;
  return MDT_OK;
}

}  // extern "C"
