// Shared device/host helpers for the B200 VAE-HPO step kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/mdt_b200.h"

namespace mdt {

// ---------------------------------------------------------------- errors
inline thread_local char g_last_error[1024] = "";

inline int set_error(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
inline int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return code;
}

#define MDT_CUDA(expr)                                                                                       \
  do {                                                                                                       \
    cudaError_t _e = (expr);                                                                                 \
    if (_e != cudaSuccess)                                                                                   \
      return ::mdt::set_error(MDT_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                              __LINE__);                                                                     \
  } while (0)

#define MDT_CHECK(cond, ...)                                         \
  do {                                                               \
    if (!(cond)) return ::mdt::set_error(MDT_ERR_INVALID, __VA_ARGS__); \
  } while (0)

#define MDT_TRY(expr)          \
  do {                         \
    int _r = (expr);           \
    if (_r != MDT_OK) return _r; \
  } while (0)

// ---------------------------------------------------------------- layout
constexpr int kSegAlign = 32;  // floats
enum Seg { W1 = 0, B1, W2X, B2X, W3, B3, W4, B4, kNumSeg };

struct Layout {
  int64_t off[kNumSeg];
  int64_t numel[kNumSeg];
  int64_t total;
};

inline int64_t round_up(int64_t n, int64_t a) { return (n + a - 1) / a * a; }

inline Layout make_layout(int D, int H, int Z) {
  Layout L;
  const int64_t n[kNumSeg] = {(int64_t)H * D, H, (int64_t)2 * Z * H, 2 * Z, (int64_t)H * Z, H, (int64_t)D * H, D};
  int64_t o = 0;
  for (int i = 0; i < kNumSeg; ++i) {
    L.off[i] = o;
    L.numel[i] = n[i];
    o += round_up(n[i], kSegAlign);
  }
  L.total = o;
  return L;
}

// ---------------------------------------------------------------- device state
// Per-trial scalars the kernels read from device memory, so a captured CUDA graph
// of the step never needs host-side arguments refreshed.
struct StepState {
  unsigned long long philox_seed;
  unsigned long long philox_offset;  // advanced by the reparameterisation kernel
  long long adam_t;                  // completed Adam steps; advanced by the Adam kernel
  double lr, beta1, beta2, eps;
  double b1_pow, b2_pow;             // beta1^adam_t, beta2^adam_t (kept incrementally: no pow() in the kernel)
  float next[6];                     // fp32 scalars of the NEXT Adam step (step_size, sqrt(bc2), 1-b1, b2, 1-b2, eps),
  float pad2[2];                     // refreshed by the last block of every Adam launch and by mdt_engine_set_adam
  unsigned int ticket_reparam;       // "last block done" counters (self-resetting)
  unsigned int ticket_adam;
  unsigned int ticket_loss;
  unsigned int host_seq;             // training steps published to host_mailbox so far
  unsigned long long* host_mailbox;  // pinned, device-mapped ring of 64 words: slot seq % 64 = (seq << 32) | bits(loss)
  unsigned int xch_seq;              // peer-memory gradient exchanges completed (never reset: the flags are monotonic)
  unsigned int xch_error;            // set when a bounded wait on a peer's flag expired
  unsigned int* host_error;          // pinned, device-mapped word the host checks before every step (or NULL): receives
                                     // xch_error, so a failed exchange stops the trial at the next call without a sync
  long long x_cursor;                // mdt_train_steps: index of the batch the NEXT step reads from the caller's pool /
                                     // index table; advanced once per published step
  float* loss_hist;                  // ring of kLossHist per-step losses (slot = (published step count - 1) % kLossHist)
  long long x_cursor_pf;             // index of the next batch the staging kernel fetches AHEAD (pipelined steps).  Only
                                     // staging kernels write it, so its value does not depend on whether the running
                                     // step has already published (and advanced x_cursor) when the fetch executes
  unsigned int ticket_prep;          // "last block done" counter of the staging kernel (self-resetting)
};
constexpr unsigned int kLossHist = 4096;

// A bounded wait on a peer expired: remember it on the device and tell the host (one posted store over PCIe).
__device__ __forceinline__ void flag_exchange_error(StepState* st) {
  st->xch_error = 1u;
  if (st->host_error != nullptr) *reinterpret_cast<volatile unsigned int*>(st->host_error) = 1u;
}

// Called by the one thread that finalises a training step's loss: hands (sequence number, loss) to the host in a
// single 8-byte store over PCIe, so mdt_train_step_host can poll for it instead of issuing a D2H copy + stream sync.
__device__ __forceinline__ void publish_loss(StepState* st, float loss) {
  unsigned long long* mb = st->host_mailbox;
  const unsigned int seq = (st->host_seq + 1u) & 0x7fffffffu;   // 31 bits: the host keeps the top bit for tickets of
                                                                  // steps that did not go through the mailbox
  st->host_seq = seq;
  st->x_cursor += 1;
  if (st->loss_hist != nullptr) st->loss_hist[(seq - 1u) & (kLossHist - 1u)] = loss;
  if (mb == nullptr) return;
  // one posted 8-byte store, nothing ordered behind it: no system-scope fence (it would hold the step's last CTA for
  // a PCIe round trip)
  // the mailbox is a ring of 64 words indexed by the sequence number, so the host may have several steps in flight
  *reinterpret_cast<volatile unsigned long long*>(mb + (seq & 63u)) =
      ((unsigned long long)seq << 32) | (unsigned long long)__float_as_uint(loss);
}

// ---------------------------------------------------------------- programmatic dependent launch
// Kernels of a step are chained with PDL: a kernel may be scheduled while its predecessor in the stream is still
// running, sets up (barriers, TMEM, descriptor prefetch) and then blocks in pdl_wait() until the predecessor has
// completed and flushed its memory; pdl_trigger() lets the successor start its own set-up.  Both are no-ops for
// launches without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline bool pdl_enabled() {
  static int on = -1;
  if (on < 0) on = getenv("MDT_NO_PDL") ? 0 : 1;
  return on == 1;
}

template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// Same, with the CTAs along z grouped into thread-block clusters of `cluster_z` (<= 8: portable cluster size).
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                      unsigned cluster_z, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = cluster_z;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---------------------------------------------------------------- device helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over a block of up to 1024 threads; result valid in thread 0.  `red` needs 32 floats.
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) red[w] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
  if (w == 0) v = warp_sum(v);
  __syncthreads();
  return v;
}

__device__ __forceinline__ float sigmoidf_ref(float a) { return 1.0f / (1.0f + expf(-a)); }

// Element of F.binary_cross_entropy(reduction='sum'): each log clamped at -100.
__device__ __forceinline__ float bce_term(float r, float x) {
  const float l1 = fmaxf(log1pf(-r), -100.0f);
  const float l0 = fmaxf(logf(r), -100.0f);
  return (x - 1.0f) * l1 - x * l0;
}
// binary_cross_entropy_backward (grad_output == 1).
__device__ __forceinline__ float bce_grad(float r, float x) { return (r - x) / fmaxf((1.0f - r) * r, 1e-12f); }
// sigmoid_backward
__device__ __forceinline__ float sigmoid_bwd(float g, float r) { return g * (1.0f - r) * r; }

}  // namespace mdt
