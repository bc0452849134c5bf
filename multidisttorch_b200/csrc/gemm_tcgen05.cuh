// tcgen05 / TMA / TMEM GEMM path (MDT_GEMM_TF32, MDT_GEMM_TF32X3) for sm_100a.
//
//   C[M,N] = epilogue( A[M,K] . B[N,K]^T ),   A and B row-major fp32 with K contiguous ("K-major").
//
// * operands are staged by TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B, out-of-bounds rows/columns
//   zero-filled) into a 4-stage shared-memory ring: A tile 128 x 32 fp32 (16 KB), B tile BN x 32 fp32;
// * one elected thread issues tcgen05.mma.cta_group::1.kind::tf32 (UMMA 128 x BN x 8), accumulating
//   in TMEM (BN fp32 columns x 128 lanes); tcgen05.commit releases ring slots / signals the epilogue;
// * TF32X3: four warps split every landed stage in place into hi = tf32(a) and lo = tf32(a - hi)
//   (elementwise, so the swizzled placement is untouched) and the issuer runs three MMAs per k-step
//   (lo.hi + hi.lo + hi.hi): fp32-grade products on the tensor cores;
// * the same four warps then drain TMEM (tcgen05.ld 32x32b.x16) into a padded shared-memory tile and
//   run the fused epilogue with coalesced row-major AND transposed global stores.
//
// Because every operand must be K-major, each activation that is later contracted over the batch
// (dW = dY^T X) is also written transposed by the epilogue that produces it, and the three weights
// that back-propagate (W4, W3, W2x) keep a transposed shadow refreshed after every Adam step.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>

#include <map>
#include <vector>

#include "common.cuh"
#include "latent.cuh"
#include "elementwise.cuh"

namespace mdt {

// ------------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------------
namespace ptx {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Spin on the phase parity.  A watchdog (about 2 s of wall clock) traps instead of hanging the GPU
// if a descriptor or byte count is wrong.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  unsigned long long t0 = 0;
  for (uint32_t spins = 0;; ++spins) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) return;
    if ((spins & 0x3ff) == 0x3ff) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (t0 == 0) t0 = t;
      else if (t - t0 > 2000000000ull) __trap();
    }
  }
}

__device__ __forceinline__ void prefetch_tmap(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}
// 2-D tiled TMA load: coordinates are {inner (K, elements), outer (row)}.
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t slot_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] . B[smem]^T, tf32 inputs, fp32 accumulate.
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]^T, bf16 (or fp16) inputs as the instruction descriptor says, fp32 accumulate.
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Same with the A operand in tensor memory (TS form): A[m][k] = lane m, column (a_tmem & 0xffff) + k, one tf32 per column.
__device__ __forceinline__ void mma_tf32_ts(uint32_t tmem_d, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 lanes x 16 consecutive 32-bit columns: thread t of the warp writes lane (base_lane + t).
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
        "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])),
        "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])),
        "r"(__float_as_uint(v[11])), "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])),
        "r"(__float_as_uint(v[15]))
      : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// mbarrier arrive once every tcgen05 op issued so far by this thread has completed.
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 16 consecutive fp32 columns: thread t of the warp receives lane (base_lane + t).
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
// Upper part of the hi/lo split: keep sign, exponent and the 10 mantissa bits tf32 has.
__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }
__device__ __forceinline__ void bar_sync_epi() { asm volatile("bar.sync 1, 256;" ::: "memory"); }  // the 8 epilogue warps

// thread-block cluster: barrier over every thread of the cluster, and access to a peer CTA's shared memory
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_map(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ float4 cluster_ld4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void cluster_st4(uint32_t addr, float4 v) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

}  // namespace ptx

// ------------------------------------------------------------------------------------------------
// descriptors
// ------------------------------------------------------------------------------------------------
// Shared-memory matrix descriptor, K-major tile of 128-byte rows, SWIZZLE_128B (8-row x 128 B atoms,
// 1024 B apart): start address >> 4 in bits [0,14), LBO (unused here) in [16,30), SBO >> 4 in [32,46),
// descriptor version 1 in [46,48), layout type 2 (SWIZZLE_128B) in [61,64).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)(1024u >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// Instruction descriptor, kind::tf32: D format F32 (bits [4,6) = 1), A and B format TF32 (bits [7,10) and
// [10,13) = 2), both operands K-major (bits 15, 16 = 0), N >> 3 in [17,23), M >> 4 in [24,29).
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// Instruction descriptor, kind::f16 with BF16 operands: D format F32 (bits [4,6) = 1), A and B format BF16 (bits [7,10)
// and [10,13) = 1), both K-major, N >> 3 in [17,23), M >> 4 in [24,29).  UMMA_K = 16 (32 bytes, like tf32's 8 x 4).
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ------------------------------------------------------------------------------------------------
// bf16 operand mirrors (MDT_GEMM_BF16): every buffer a GEMM reads through TMA exists as bf16; the kernel that produces
// the fp32 value writes the mirror too.  Buffers that are ONLY GEMM operands (the transposed activations) hold bf16
// instead of fp32 in that mode -- `b16` selects the element type of such a store.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_operand(void* base, size_t idx, float v, bool b16) {
  if (b16) reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16_rn(v);
  else reinterpret_cast<float*>(base)[idx] = v;
}
__device__ __forceinline__ void st_bf16x4(__nv_bfloat16* dst, float4 v) {      // dst 8-byte aligned
  const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<const uint32_t*>(&lo);
  u.y = *reinterpret_cast<const uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(dst) = u;
}

// ------------------------------------------------------------------------------------------------
// epilogues
// ------------------------------------------------------------------------------------------------
enum Tc05EpiKind {
  EPI_BIAS_ACT = 0,     // C = act(acc + bias[n]); optional transposed copy
  EPI_REPARAM = 1,      // (fc21|fc22)^T tile -> ml, eps, std, z, z^T, direct KLD gradients, KLD partial sum
  EPI_SIGMOID_BCE = 2,  // fc4 tile -> recon, BCE partial, d(logits) and its transpose; last CTA writes the loss
  EPI_RELU_MASK = 3,    // dX tile masked by the forward activation; optional row-major and transposed stores
  EPI_GRADW = 4,        // weight-gradient tile + bias gradient (row sums of the A operand)
  EPI_REPARAM_BWD = 5,  // dz^T tile -> d[mu|logvar] and its transpose
};

struct Tc05Epi {
  // generic
  float* C; int ldc;
  void* CT; int ldct;          // transposed copy: fp32, or bf16 when `bf16` is set
  const float* bias; int act;
  const float* mask; int ldmask;
  // reparameterisation (forward)
  const float* eps_in; StepState* st; long long philox_S; unsigned long long philox_inc;
  float *z; void* zT; float *eps_out, *sd_out, *dml_kl; int Z; int ldzT;
  float* kl_partials;
  // BCE
  const float* x; int ldx; float* da4; void* da4T; int ld_da4T;
  float* bce_partials; unsigned int* ticket; float* loss_out; int n_kl_partials;
  // weight gradient
  const float* a_src; int lda_src; float* db;
  // reparameterisation (backward)
  const float *eps, *sd; float *dml; void* dmlT; int ld_dmlT;
  // weight-gradient GEMM with the Adam update fused in (single-rank trial groups): parameter / moment segments of
  // this weight (same [M][ldc] layout as the gradient), its transposed shadow (or NULL), the bias segments, the
  // device-resident step state and, for the CTA that takes the step's last ticket, the loss partials to add up
  float *adam_p, *adam_m, *adam_v, *adam_pT; int adam_ldpT;
  float *adam_pb, *adam_mb, *adam_vb;
  StepState* adam_st; int adam_total_ctas;
  const float* loss_part_a; int n_part_a; const float* loss_part_b; int n_part_b; float* adam_loss_out;
  // split-K workspace: [tile][split][128*BN] partial tiles + one ticket counter per tile
  float* ws; unsigned int* ws_counters;
  int a_tmem;    // TF32X3: the split warps put hi / lo of the A tile into tensor memory (MMA TS form) instead of
                 // writing lo next to it in shared memory -- the MMAs then read only B from shared memory
  int cluster;   // split-K CTAs of a tile form a thread-block cluster (1,1,splits): partials are exchanged through DSMEM
  // MDT_GEMM_BF16: operands are bf16 (kind::f16, 64 K-values per k-block); transposed outputs (CT, da4T, zT, dmlT,
  // adam_pT16) hold bf16; Cb / da4b / zb / dmlb are row-major bf16 mirrors of C / da4 / z / dml (or NULL)
  int bf16;
  __nv_bfloat16* Cb; int ldcb;
  __nv_bfloat16* da4b; __nv_bfloat16* zb; int ldzb; __nv_bfloat16* dmlb;
  __nv_bfloat16* adam_p16; int adam_ldp16;      // row-major bf16 copy of the updated weight
  __nv_bfloat16* adam_pT16;                     // transposed bf16 shadow of the updated weight (leading dim adam_ldpT)
  int pre_a, pre_b;  // the A / B operand is not written by any kernel of this step before this one (a weight or its
                     // transposed shadow: last written a whole step ago): its TMA loads for the first ring pass are issued
                     // BEFORE griddepcontrol.wait, so they land while the predecessor kernel is still running
  // optional phase timing (clock64 stamps of CTA (0,0)); NULL in production
  long long* dbg;
};

constexpr int kTcEpiThreads = 256;                 // warps 2-9: hi/lo split (TF32X3) + epilogue
constexpr int kTcThreads = 64 + kTcEpiThreads;     // warp 0: TMA, warp 1: MMA + TMEM alloc
constexpr int kTcDrainThreads = 128;               // TF32X3 only: warps 10-13 drain the per-k-block partial accumulators
constexpr int tc_threads(bool split) { return split ? kTcThreads + kTcDrainThreads : kTcThreads; }
constexpr int kTcBM = 128;
constexpr int kTcBK = 32;        // fp32 elements = one 128-byte swizzle row

template <int BN, bool kSplit>
struct Tc05Smem {
  static constexpr int A_BYTES = kTcBM * 128;
  static constexpr int B_BYTES = BN * 128;
  static constexpr int HB = A_BYTES + B_BYTES;          // bytes TMA delivers per stage
  static constexpr int STAGE = kSplit ? 2 * HB : HB;    // + the "lo" halves in split mode
  static constexpr int STAGES = kSplit ? 4 : 8;         // ring depth: about 190 KB either way at BN = 64
  static constexpr int RING = STAGES * STAGE;
  static constexpr int TILE_LD = BN + 4;                // padded epilogue tile (16-byte rows), reuses the ring
  static constexpr int BAR_OFF = RING;
  static constexpr int TOTAL = RING + 256 + 1024;       // barriers/tmem slot/reduction scratch + alignment slack
  static_assert((kTcBM * TILE_LD + BN * 128) * 4 <= RING, "epilogue tile + transposed-store scratch must fit in the ring");
};

// ------------------------------------------------------------------------------------------------
// fused epilogues: run by the 128 epilogue threads on the accumulator tile T[128][ldt] in shared memory
// ------------------------------------------------------------------------------------------------
// Epilogue math of the tensor-core path.  These run on a handful of warps per SM, where the length of the dependent
// instruction chain per element (not throughput) sets the time, so the short-latency SFU forms are used:
// ex2/lg2/rcp approximations are accurate to ~2 ulp, far inside the tolerance of this path, and the clamp
// semantics of F.binary_cross_entropy are kept exactly (log(0) = -inf -> -100; denominators below 1e-12).
__device__ __forceinline__ float fast_sigmoid(float a) { return __frcp_rn(1.0f + __expf(-a)); }
__device__ __forceinline__ float fast_bce_term(float r, float x) {
  const float l1 = fmaxf(__logf(1.0f - r), -100.0f);
  const float l0 = fmaxf(__logf(r), -100.0f);
  return (x - 1.0f) * l1 - x * l0;
}
// d loss / d logit = bce_grad * r(1-r): equals r - x whenever r(1-r) >= 1e-12 (the clamp is inactive)
__device__ __forceinline__ float fast_dlogit(float r, float x) {
  const float p = (1.0f - r) * r;
  return p >= 1e-12f ? (r - x) : (r - x) * 1e12f * p;
}

// Transposed store of the tile: thread = (tile row, column half); reads float4 from its row, writes four scalars,
// each of which is coalesced across the warp (consecutive lanes = consecutive m).
template <int BN>
__device__ __forceinline__ void tc05_store_transposed(const float* T, int ldt, void* CT, int ldct, int m0, int n0, int M,
                                                      int N, int et, bool b16 = false) {
  const int r = et % kTcBM, ch = et / kTcBM;     // 256 threads: two column halves
  const int m = m0 + r;
  if (m >= M) return;
  constexpr int HALF = BN / 2;
  const float* trow = T + r * ldt + ch * HALF;
  const size_t base = (size_t)(n0 + ch * HALF) * ldct + m;
  const int cvalid = N - n0 - ch * HALF;         // columns of this half that exist
#pragma unroll
  for (int c = 0; c < HALF; c += 4) {
    const float4 t = *reinterpret_cast<const float4*>(trow + c);
    if (c + 0 < cvalid) st_operand(CT, base + (size_t)(c + 0) * ldct, t.x, b16);
    if (c + 1 < cvalid) st_operand(CT, base + (size_t)(c + 1) * ldct, t.y, b16);
    if (c + 2 < cvalid) st_operand(CT, base + (size_t)(c + 2) * ldct, t.z, b16);
    if (c + 3 < cvalid) st_operand(CT, base + (size_t)(c + 3) * ldct, t.w, b16);
  }
}

// The same for the RPP = 1024 / BN tile rows of one epilogue pass (row_lo .. row_lo + RPP): thread = (row, group of 4
// columns), one float4 read and 4 coalesced scalar stores.  Used when the passes of a tile are shared out among the
// CTAs of a split-K cluster.
template <int BN>
__device__ __forceinline__ void tc05_store_transposed_rows(const float* T, int ldt, void* CT, int ldct, int m0, int n0,
                                                           int M, int N, int et, int row_lo, bool b16 = false) {
  constexpr int RPP = kTcEpiThreads / (BN / 4);
  const int r = row_lo + et % RPP, c0 = (et / RPP) * 4;
  const int m = m0 + r;
  if (m >= M) return;
  const float4 t = *reinterpret_cast<const float4*>(T + r * ldt + c0);
  const size_t base = (size_t)(n0 + c0) * ldct + m;
  const int cvalid = N - n0 - c0;
  if (0 < cvalid) st_operand(CT, base, t.x, b16);
  if (1 < cvalid) st_operand(CT, base + (size_t)ldct, t.y, b16);
  if (2 < cvalid) st_operand(CT, base + (size_t)2 * ldct, t.z, b16);
  if (3 < cvalid) st_operand(CT, base + (size_t)3 * ldct, t.w, b16);
}

// Global inputs of the element-wise epilogues (bias, ReLU mask / BCE target) that do not depend on this kernel's own
// result: fetched into registers right after the dependency wait, so their L2 latency hides behind the main loop
// instead of sitting at the head of the epilogue.
template <int BN>
struct Tc05Pre {
  static constexpr int ITERS = kTcBM / (kTcEpiThreads / (BN / 4));
  float4 bias;
  float4 aux[ITERS];
  // fused Adam (EPI_GRADW): the step's scalars, and -- for narrow tiles, where they fit in registers -- this thread's
  // slice of the weight and both moments, fetched before the main loop so the epilogue starts with them in hand
  static constexpr bool kAdamRegs = ITERS <= 4;
  float as[6];
  float4 ap[kAdamRegs ? ITERS : 1], am[kAdamRegs ? ITERS : 1], av[kAdamRegs ? ITERS : 1];
  bool adam_loaded;
};

template <int BN, int kEpi>
__device__ __forceinline__ void tc05_prefetch(const Tc05Epi& ep, int m0, int n0, int M, int N, int et, Tc05Pre<BN>& pre,
                                              int it_rank, int it_ns) {
  constexpr int F4R = BN / 4;
  constexpr int RPP = kTcEpiThreads / F4R;
  constexpr int ITERS = kTcBM / RPP;
  const int c4 = et % F4R, rbase = et / F4R;
  const int n = n0 + 4 * c4;
  const bool nvalid = n < N;
  const int rows_valid = nvalid ? (M - m0 - rbase + RPP - 1) / RPP : 0;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  pre.bias = zero4;
  if ((kEpi == EPI_BIAS_ACT || kEpi == EPI_SIGMOID_BCE) && ep.bias != nullptr && nvalid)
    pre.bias = __ldg(reinterpret_cast<const float4*>(ep.bias + n));
  if (kEpi == EPI_RELU_MASK || kEpi == EPI_SIGMOID_BCE) {
    const float* const src = (kEpi == EPI_RELU_MASK) ? ep.mask : ep.x;
    const int ld = (kEpi == EPI_RELU_MASK) ? ep.ldmask : ep.ldx;
    const float* const sp = src + (size_t)(m0 + rbase) * ld + n;
    const size_t step = (size_t)RPP * ld;
    const float fill = (kEpi == EPI_RELU_MASK) ? 0.f : 0.5f;
#pragma unroll
    for (int i = 0; i < ITERS; ++i)
      pre.aux[i] = (i < rows_valid && (i % it_ns) == it_rank) ? __ldg(reinterpret_cast<const float4*>(sp + i * step))
                                                             : make_float4(fill, fill, fill, fill);
  }
}

// Fused Adam (EPI_GRADW): the step's scalars and this thread's slice of the weight and both moments.  Called BEFORE the
// dependency wait: the scalars change only after every CTA of the step's weight-gradient GEMMs has taken its ticket
// (kernel tail of the PREVIOUS step by the time any kernel of this step runs), and nobody but this tile's CTA writes
// this tile of p / m / v, last in the previous step.  Plain loads (the data is rewritten in the epilogue).
template <int BN, int kEpi>
__device__ __forceinline__ void tc05_prefetch_adam(const Tc05Epi& ep, int m0, int n0, int M, int N, int et, Tc05Pre<BN>& pre) {
  constexpr int F4R = BN / 4;
  constexpr int RPP = kTcEpiThreads / F4R;
  constexpr int ITERS = kTcBM / RPP;
  const int c4 = et % F4R, rbase = et / F4R;
  const int n = n0 + 4 * c4;
  const bool nvalid = n < N;
  const int rows_valid = nvalid ? (M - m0 - rbase + RPP - 1) / RPP : 0;
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  pre.adam_loaded = false;
  if (kEpi == EPI_GRADW && ep.adam_p != nullptr) {
#pragma unroll
    for (int k = 0; k < 6; ++k) pre.as[k] = ep.adam_st->next[k];
    if (Tc05Pre<BN>::kAdamRegs && gridDim.z == 1) {
      const size_t off0 = (size_t)(m0 + rbase) * ep.ldc + n, ostep = (size_t)RPP * ep.ldc;
#pragma unroll
      for (int i = 0; i < (Tc05Pre<BN>::kAdamRegs ? ITERS : 1); ++i) {
        const bool ok = i < rows_valid;
        const size_t o = off0 + (size_t)i * ostep;
        pre.ap[i] = ok ? *reinterpret_cast<const float4*>(ep.adam_p + o) : zero4;
        pre.am[i] = ok ? *reinterpret_cast<const float4*>(ep.adam_m + o) : zero4;
        pre.av[i] = ok ? *reinterpret_cast<const float4*>(ep.adam_v + o) : zero4;
      }
      pre.adam_loaded = true;
    }
  }
}

template <int BN, int kEpi>
__device__ __forceinline__ void tc05_epilogue(const Tc05Epi& ep, float* T, int ldt, int m0, int n0, int M, int N, int et,
                                              float* red, const Tc05Pre<BN>& pre, int it_rank, int it_ns, int by) {
  // it_rank / it_ns: in a split-K cluster the tile's epilogue passes (RPP rows each) are shared out among the CTAs --
  // this CTA holds the summed rows of, and runs, the passes i with i % it_ns == it_rank.  (0, 1) = the whole tile.
  const int lane = et & 31, ew = et >> 5;
  // Element-wise epilogues: 256 threads sweep the tile row by row in float4 units.  A thread always owns the same
  // four columns and rows rbase, rbase+RPP, ...; everything it needs (tile values from shared memory, mask / x
  // from global) is loaded up-front, the math is straight-line code over all its elements, only stores are
  // predicated.  N, every leading dimension and every base pointer are multiples of 4 floats (checked on the host).
  constexpr int F4R = BN / 4;                    // float4 per tile row
  constexpr int RPP = kTcEpiThreads / F4R;       // tile rows covered per pass
  constexpr int ITERS = kTcBM / RPP;
  const int c4 = et % F4R, rbase = et / F4R;
  const int n = n0 + 4 * c4;
  const bool nvalid = n < N;
  const int rows_valid = nvalid ? (M - m0 - rbase + RPP - 1) / RPP : 0;   // element i is valid iff i < rows_valid
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (kEpi == EPI_BIAS_ACT || kEpi == EPI_RELU_MASK || kEpi == EPI_GRADW) {
    float* const Cp = ep.C;
    void* const CTp0 = (kEpi != EPI_GRADW) ? ep.CT : nullptr;
    const int ldc = ep.ldc, act = ep.act;
    float4 v[ITERS];
#pragma unroll
    for (int i = 0; i < ITERS; ++i) v[i] = *reinterpret_cast<const float4*>(T + (rbase + i * RPP) * ldt + 4 * c4);
    if (kEpi == EPI_BIAS_ACT) {
      const float4 b = pre.bias;
#pragma unroll
      for (int i = 0; i < ITERS; ++i) { v[i].x += b.x; v[i].y += b.y; v[i].z += b.z; v[i].w += b.w; }
      if (act == MDT_ACT_RELU) {
#pragma unroll
        for (int i = 0; i < ITERS; ++i) {
          v[i].x = fmaxf(v[i].x, 0.f); v[i].y = fmaxf(v[i].y, 0.f); v[i].z = fmaxf(v[i].z, 0.f); v[i].w = fmaxf(v[i].w, 0.f);
        }
      } else if (act == MDT_ACT_SIGMOID) {
#pragma unroll
        for (int i = 0; i < ITERS; ++i) {
          v[i].x = fast_sigmoid(v[i].x); v[i].y = fast_sigmoid(v[i].y); v[i].z = fast_sigmoid(v[i].z); v[i].w = fast_sigmoid(v[i].w);
        }
      }
    } else if (kEpi == EPI_RELU_MASK) {
#pragma unroll
      for (int i = 0; i < ITERS; ++i) {
        const float4 mk = pre.aux[i];
        v[i].x = mk.x > 0.f ? v[i].x : 0.f; v[i].y = mk.y > 0.f ? v[i].y : 0.f;
        v[i].z = mk.z > 0.f ? v[i].z : 0.f; v[i].w = mk.w > 0.f ? v[i].w : 0.f;
      }
    }
    if (Cp != nullptr) {
      float* const cp = Cp + (size_t)(m0 + rbase) * ldc + n;
      const size_t cstep = (size_t)RPP * ldc;
#pragma unroll
      for (int i = 0; i < ITERS; ++i)
        if (i < rows_valid && (i % it_ns) == it_rank) *reinterpret_cast<float4*>(cp + i * cstep) = v[i];
    }
    if (kEpi != EPI_GRADW && ep.Cb != nullptr) {          // bf16 mirror of the activation (operand of a later GEMM)
      __nv_bfloat16* const cb = ep.Cb + (size_t)(m0 + rbase) * ep.ldcb + n;
      const size_t cstep = (size_t)RPP * ep.ldcb;
#pragma unroll
      for (int i = 0; i < ITERS; ++i)
        if (i < rows_valid && (i % it_ns) == it_rank) st_bf16x4(cb + i * cstep, v[i]);
    }
    void* CTp = CTp0;
    bool ct16 = ep.bf16 != 0;                              // element type of the transposed store below
    if (kEpi == EPI_GRADW && ep.adam_p != nullptr) {
      // fused Adam on this tile of the weight (adam_elem_fast: same formula as adam_flat_kernel, SFU sqrt/quotient)
      const float step_size = pre.as[0], bc2_sqrt = __frcp_rn(pre.as[1]), omb1 = pre.as[2], b2 = pre.as[3], omb2 = pre.as[4],
                  aeps = pre.as[5];
      const size_t off0 = (size_t)(m0 + rbase) * ldc + n, ostep = (size_t)RPP * ldc;
      constexpr int CHK = ITERS < 4 ? ITERS : 4;           // rows per batch of loads (register budget)
#pragma unroll
      for (int i0 = 0; i0 < ITERS; i0 += CHK) {
        float4 pp[CHK], mm[CHK], vv[CHK];
#pragma unroll
        for (int c = 0; c < CHK; ++c) {
          const bool ok = i0 + c < rows_valid && ((i0 + c) % it_ns) == it_rank;
          const size_t o = off0 + (size_t)(i0 + c) * ostep;
          if (Tc05Pre<BN>::kAdamRegs && pre.adam_loaded) {
            pp[c] = pre.ap[Tc05Pre<BN>::kAdamRegs ? i0 + c : 0];
            mm[c] = pre.am[Tc05Pre<BN>::kAdamRegs ? i0 + c : 0];
            vv[c] = pre.av[Tc05Pre<BN>::kAdamRegs ? i0 + c : 0];
          } else {
            pp[c] = ok ? *reinterpret_cast<const float4*>(ep.adam_p + o) : zero4;
            mm[c] = ok ? *reinterpret_cast<const float4*>(ep.adam_m + o) : zero4;
            vv[c] = ok ? *reinterpret_cast<const float4*>(ep.adam_v + o) : zero4;
          }
        }
#pragma unroll
        for (int c = 0; c < CHK; ++c) {
          const float4 g = v[i0 + c];
          adam_elem_fast(pp[c].x, g.x, mm[c].x, vv[c].x, step_size, bc2_sqrt, omb1, b2, omb2, aeps);
          adam_elem_fast(pp[c].y, g.y, mm[c].y, vv[c].y, step_size, bc2_sqrt, omb1, b2, omb2, aeps);
          adam_elem_fast(pp[c].z, g.z, mm[c].z, vv[c].z, step_size, bc2_sqrt, omb1, b2, omb2, aeps);
          adam_elem_fast(pp[c].w, g.w, mm[c].w, vv[c].w, step_size, bc2_sqrt, omb1, b2, omb2, aeps);
          if (i0 + c < rows_valid && ((i0 + c) % it_ns) == it_rank) {
            const size_t o = off0 + (size_t)(i0 + c) * ostep;
            *reinterpret_cast<float4*>(ep.adam_p + o) = pp[c];
            *reinterpret_cast<float4*>(ep.adam_m + o) = mm[c];
            *reinterpret_cast<float4*>(ep.adam_v + o) = vv[c];
          }
          v[i0 + c] = pp[c];                               // the updated weight feeds the transposed shadow below
        }
      }
      if (ep.adam_p16 != nullptr) {                        // bf16 mode: the GEMMs of the next step read this copy
        __nv_bfloat16* const wb = ep.adam_p16 + (size_t)(m0 + rbase) * ep.adam_ldp16 + n;
        const size_t wstep = (size_t)RPP * ep.adam_ldp16;
#pragma unroll
        for (int i = 0; i < ITERS; ++i)
          if (i < rows_valid && (i % it_ns) == it_rank) st_bf16x4(wb + i * wstep, v[i]);
      }
      CTp = ep.adam_pT;
      ct16 = false;                                        // the fp32 shadow (read by the CUDA-core latent kernels)
    }
    const int ldct = (kEpi == EPI_GRADW) ? ep.adam_ldpT : ep.ldct;
    void* const CT16 = (kEpi == EPI_GRADW) ? (void*)ep.adam_pT16 : nullptr;   // second, bf16 shadow of an updated weight
    if (CTp != nullptr || CT16 != nullptr) {
#pragma unroll
      for (int i = 0; i < ITERS; ++i) *reinterpret_cast<float4*>(T + (rbase + i * RPP) * ldt + 4 * c4) = v[i];
      ptx::bar_sync_epi();
#pragma unroll
      for (int pass = 0; pass < 2; ++pass) {
        void* const dst = pass == 0 ? CTp : CT16;
        const bool b16 = pass == 0 ? ct16 : true;
        if (dst == nullptr) continue;
        if (it_ns == 1) {
          tc05_store_transposed<BN>(T, ldt, dst, ldct, m0, n0, M, N, et, b16);
        } else {
          for (int i = it_rank; i < ITERS; i += it_ns) tc05_store_transposed_rows<BN>(T, ldt, dst, ldct, m0, n0, M, N, et, i * RPP, b16);
        }
      }
    }
  } else if (kEpi == EPI_SIGMOID_BCE) {
    const int ldc = ep.ldc;
    const float4 b = pre.bias;
    float4 v[ITERS], xv[ITERS];
#pragma unroll
    for (int i = 0; i < ITERS; ++i) xv[i] = pre.aux[i];
#pragma unroll
    for (int i = 0; i < ITERS; ++i) v[i] = *reinterpret_cast<const float4*>(T + (rbase + i * RPP) * ldt + 4 * c4);
    float part = 0.f;
    float* const rp = ep.C + (size_t)(m0 + rbase) * ldc + n;       // recon
    float* const gp = ep.da4 + (size_t)(m0 + rbase) * ldc + n;     // dL/d(logits)
    const size_t cstep = (size_t)RPP * ldc;
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
      if ((i % it_ns) != it_rank) continue;        // CTA-uniform
      float4 r, g;
      r.x = fast_sigmoid(v[i].x + b.x); r.y = fast_sigmoid(v[i].y + b.y);
      r.z = fast_sigmoid(v[i].z + b.z); r.w = fast_sigmoid(v[i].w + b.w);
      const float term = (fast_bce_term(r.x, xv[i].x) + fast_bce_term(r.y, xv[i].y)) +
                         (fast_bce_term(r.z, xv[i].z) + fast_bce_term(r.w, xv[i].w));
      g.x = fast_dlogit(r.x, xv[i].x); g.y = fast_dlogit(r.y, xv[i].y);
      g.z = fast_dlogit(r.z, xv[i].z); g.w = fast_dlogit(r.w, xv[i].w);
      const bool ok = i < rows_valid;
      part += ok ? term : 0.f;
      if (ok) {
        *reinterpret_cast<float4*>(rp + i * cstep) = r;
        *reinterpret_cast<float4*>(gp + i * cstep) = g;
        if (ep.da4b != nullptr) st_bf16x4(ep.da4b + (size_t)(m0 + rbase) * ldc + n + i * cstep, g);
      }
      v[i] = g;
    }
#pragma unroll
    for (int i = 0; i < ITERS; ++i) *reinterpret_cast<float4*>(T + (rbase + i * RPP) * ldt + 4 * c4) = v[i];
    part = warp_sum(part);
    if (lane == 0) red[ew] = part;
    ptx::bar_sync_epi();
    if (it_ns == 1) {
      tc05_store_transposed<BN>(T, ldt, ep.da4T, ep.ld_da4T, m0, n0, M, N, et, ep.bf16 != 0);
    } else {
      for (int i = it_rank; i < ITERS; i += it_ns)
        tc05_store_transposed_rows<BN>(T, ldt, ep.da4T, ep.ld_da4T, m0, n0, M, N, et, i * RPP, ep.bf16 != 0);
    }
    // per-CTA partial only: the loss itself is not needed by any later kernel of the step, so the fixed-order sum
    // of the partials is done by the Adam kernel (finalize_loss), after a kernel boundary -- no fence / ticket here.
    // A tile owns gridDim.z consecutive slots (one per split); without the cluster exchange only slot 0 is used.
    if (et == 0) {
      float* slot = ep.bce_partials + (size_t)(by * gridDim.x + blockIdx.x) * gridDim.z;
      slot[it_ns == 1 ? 0 : it_rank] = ((red[0] + red[1]) + (red[2] + red[3])) + ((red[4] + red[5]) + (red[6] + red[7]));
      if (it_ns == 1)
        for (unsigned z = 1; z < gridDim.z; ++z) slot[z] = 0.f;
    }
  } else if (kEpi == EPI_REPARAM || kEpi == EPI_REPARAM_BWD) {
    // These two GEMMs have a skinny output (2Z or Z features), so they run as the TRANSPOSED problem
    // C^T[feature, sample] = W[feature, K] . X[sample, K]^T: the batch becomes the tiled N dimension and the work
    // spreads over B/BN CTAs instead of one.  T[f][n] holds feature f of sample n0 + n.  The per-sample inputs of
    // the tile (eps / std / dml_kl rows n0..n0+nb) are contiguous in global memory and are staged through shared
    // memory with coalesced float4 loads, so the element loop has no global-load latency in it.
    const int Z = ep.Z;
    const int nb = min(BN, N - n0);              // valid samples of this tile (N = batch)
    const int ne = nb * Z;
    float* const S0 = T + kTcBM * ldt;
    if (kEpi == EPI_REPARAM) {
      float* const sE = S0;                      // eps [nb][Z]
      float* const sB = S0 + BN * 32;            // bias [2Z]
      {
        const float4* src = reinterpret_cast<const float4*>(ep.eps_in + (size_t)n0 * Z);
        float4* dst = reinterpret_cast<float4*>(sE);
#pragma unroll
        for (int q = 0; q < (BN * 32 / 4 + kTcEpiThreads - 1) / kTcEpiThreads; ++q) {
          const int i = et + q * kTcEpiThreads;
          if (i < ne / 4) dst[i] = __ldg(src + i);
        }
        if (et < 2 * Z) sB[et] = __ldg(ep.bias + et);
      }
      ptx::bar_sync_epi();
      float kl = 0.f;
      for (int idx = et; idx < ne; idx += kTcEpiThreads) {
        const int n = idx / Z, j = idx - n * Z;
        const size_t m = (size_t)(n0 + n);
        const float mu = T[j * ldt + n] + sB[j];
        const float lv = T[(Z + j) * ldt + n] + sB[Z + j];
        const float e = sE[idx];
        const float sd = __expf(0.5f * lv);
        const float ev = __expf(lv);
        const float zv = mu + e * sd;
        ep.C[m * 2 * Z + j] = mu;
        ep.C[m * 2 * Z + Z + j] = lv;
        ep.z[m * Z + j] = zv;
        st_operand(ep.zT, (size_t)j * ep.ldzT + m, zv, ep.bf16 != 0);
        if (ep.zb != nullptr) ep.zb[m * ep.ldzb + j] = __float2bfloat16_rn(zv);
        if (ep.eps_out != ep.eps_in) ep.eps_out[m * Z + j] = e;
        ep.sd_out[m * Z + j] = sd;
        ep.dml_kl[m * 2 * Z + j] = mu;
        ep.dml_kl[m * 2 * Z + Z + j] = 0.5f * (ev - 1.0f);
        kl += -0.5f * (1.0f + lv - mu * mu - ev);
      }
      kl = warp_sum(kl);
      if (lane == 0) red[ew] = kl;
      ptx::bar_sync_epi();
      if (et == 0)
        ep.kl_partials[blockIdx.x] = ((red[0] + red[1]) + (red[2] + red[3])) + ((red[4] + red[5]) + (red[6] + red[7]));
    } else {
      float* const sE = S0;                      // eps [nb][Z]
      float* const sS = S0 + BN * 32;            // std [nb][Z]
      float* const sK = S0 + BN * 64;            // dml_kl [nb][2Z]
      {
        const float4* s0 = reinterpret_cast<const float4*>(ep.eps + (size_t)n0 * Z);
        const float4* s1 = reinterpret_cast<const float4*>(ep.sd + (size_t)n0 * Z);
        const float4* s2 = reinterpret_cast<const float4*>(ep.dml_kl + (size_t)n0 * 2 * Z);
        float4* d0 = reinterpret_cast<float4*>(sE);
        float4* d1 = reinterpret_cast<float4*>(sS);
        float4* d2 = reinterpret_cast<float4*>(sK);
        constexpr int Q = (BN * 64 / 4 + kTcEpiThreads - 1) / kTcEpiThreads;
        float4 t0[Q], t1[Q], t2[Q];
#pragma unroll
        for (int q = 0; q < Q; ++q) {              // fixed trip count: every load is in flight before the first store
          const int i = et + q * kTcEpiThreads;
          if (i < ne / 4) { t0[q] = __ldg(s0 + i); t1[q] = __ldg(s1 + i); }
          if (i < ne / 2) t2[q] = __ldg(s2 + i);
        }
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          const int i = et + q * kTcEpiThreads;
          if (i < ne / 4) { d0[i] = t0[q]; d1[i] = t1[q]; }
          if (i < ne / 2) d2[i] = t2[q];
        }
      }
      ptx::bar_sync_epi();
      for (int idx = et; idx < ne; idx += kTcEpiThreads) {
        const int n = idx / Z, j = idx - n * Z;
        const size_t m = (size_t)(n0 + n);
        const float g = T[j * ldt + n];
        const float dmu = g + sK[n * 2 * Z + j];
        const float dlv = g * sE[idx] * sS[idx] * 0.5f + sK[n * 2 * Z + Z + j];
        ep.dml[m * 2 * Z + j] = dmu;
        ep.dml[m * 2 * Z + Z + j] = dlv;
        st_operand(ep.dmlT, (size_t)j * ep.ld_dmlT + m, dmu, ep.bf16 != 0);
        st_operand(ep.dmlT, (size_t)(Z + j) * ep.ld_dmlT + m, dlv, ep.bf16 != 0);
        if (ep.dmlb != nullptr) {
          ep.dmlb[m * 2 * Z + j] = __float2bfloat16_rn(dmu);
          ep.dmlb[m * 2 * Z + Z + j] = __float2bfloat16_rn(dlv);
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------------
// `by`: index of the 128-row tile (blockIdx.y, or blockIdx.y modulo the tiles of one trial in the grouped launch)
template <int BN, bool kSplit, int kEpi>
__device__ __forceinline__ void tc05_gemm_body(const CUtensorMap* const tmap_a_p, const CUtensorMap* const tmap_b_p, int M, int N,
                                               int K, const Tc05Epi& ep, const int by) {
  using S = Tc05Smem<BN, kSplit>;
  // SWIZZLE_128B needs 1024-byte aligned tiles.  The kernel has no static shared memory, so the dynamic window
  // starts at offset 0; keeping `smem` a plain shared-space array (no integer re-alignment) lets the compiler
  // emit LDS/STS instead of generic loads/stores for every access below.
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t smem_base = ptx::smem_u32(smem);
  if ((smem_base & 1023u) != 0) __trap();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN, m0 = by * kTcBM;
  // split-K: gridDim.z CTAs share one output tile, each streaming a contiguous range of k-blocks
  // elements of K per k-block: one 128-byte swizzle row = 32 fp32 (tf32 kinds) or 64 bf16 (kind::f16)
  const int kbk = (!kSplit && ep.bf16) ? 2 * kTcBK : kTcBK;
  const int total_kb = (K + kbk - 1) / kbk;
  const int kb_per_split = (total_kb + (int)gridDim.z - 1) / (int)gridDim.z;
  const int kb0 = (int)blockIdx.z * kb_per_split;
  const int num_kb = min(kb_per_split, total_kb - kb0);      // >= 1 (the host picks gridDim.z accordingly)
  const bool dbg = ep.dbg != nullptr && blockIdx.x == 0 && by == 0 && blockIdx.z == (ep.cluster ? 0 : gridDim.z - 1);
#define TC_STAMP(i) do { if (dbg) ep.dbg[i] = clock64(); } while (0)
  if (threadIdx.x == 0) {
    TC_STAMP(0);
    if (dbg) { unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); ep.dbg[14] = (long long)g; }
  }

  const uint32_t bar_full = smem_base + S::BAR_OFF;           // [S::STAGES] TMA bytes landed
  const uint32_t bar_split = bar_full + 8 * S::STAGES;        // [S::STAGES] hi/lo split done (split mode)
  const uint32_t bar_empty = bar_split + 8 * S::STAGES;       // [S::STAGES] MMAs that read the slot retired
  const uint32_t bar_accum = bar_empty + 8 * S::STAGES;       // accumulator complete (single-pass mode)
  const uint32_t bar_afull = bar_accum + 8;                   // [2] partial accumulator of a k-block ready (TF32X3)
  const uint32_t bar_aempty = bar_afull + 16;                 // [2] partial accumulator drained (TF32X3)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::BAR_OFF + 8 * (3 * S::STAGES + 5));
  float* red = reinterpret_cast<float*>(smem + S::BAR_OFF + 8 * (3 * S::STAGES + 5) + 16);  // 8 floats (one per epilogue warp)
  constexpr int kDrainWarps = kTcDrainThreads / 32;                 // TF32X3: one drain warp per TMEM lane quadrant
  constexpr int kTmemAccCols = kSplit ? (2 * BN < 32 ? 32 : 2 * BN) : (BN < 32 ? 32 : BN);
  constexpr int kTmemABase = kTmemAccCols;              // TS form: A stages follow the accumulators
  const int kTmemCols = (kSplit && ep.a_tmem) ? 512 : kTmemAccCols;   // power of two >= 2 BN + 4 * 64

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(tmap_a_p);
    ptx::prefetch_tmap(tmap_b_p);
    for (int s = 0; s < S::STAGES; ++s) {
      ptx::mbar_init(bar_full + 8 * s, 1);
      ptx::mbar_init(bar_split + 8 * s, kTcEpiThreads / 32);   // one arrival per splitter warp
      ptx::mbar_init(bar_empty + 8 * s, 1);
    }
    ptx::mbar_init(bar_accum, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(bar_afull + 8 * i, 1);
      ptx::mbar_init(bar_aempty + 8 * i, kDrainWarps);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), kTmemCols);   // power of two, at least 32 columns
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // everything above touched only shared memory, TMEM and kernel parameters.  Still ahead of the dependency wait:
  // whatever this kernel reads that NO kernel of this step has written before it -- by the time this kernel can be
  // scheduled its predecessor has passed its own wait, so everything older than the predecessor is complete.
  int pre_kb = 0;                                  // ring stages whose static operand is already in flight (producer only)
  if ((ep.pre_a | ep.pre_b) != 0 && warp == 0 && lane == 0) {
    pre_kb = min(num_kb, S::STAGES);
    for (int kb = 0; kb < pre_kb; ++kb) {
      const uint32_t sa = smem_base + kb * S::STAGE;
      ptx::mbar_expect_tx(bar_full + 8 * kb, S::HB);
      if (ep.pre_a) ptx::tma_load_2d(sa, tmap_a_p, (kb0 + kb) * kbk, m0, bar_full + 8 * kb);
      if (ep.pre_b) ptx::tma_load_2d(sa + S::A_BYTES, tmap_b_p, (kb0 + kb) * kbk, n0, bar_full + 8 * kb);
    }
  }
  Tc05Pre<BN> pre;
  const bool is_epi_warp = warp >= 2 && warp < 2 + kTcEpiThreads / 32;
  if (kEpi == EPI_GRADW && is_epi_warp) tc05_prefetch_adam<BN, kEpi>(ep, m0, n0, M, N, (int)threadIdx.x - 64, pre);
  pdl_wait();
  pdl_trigger();
  if (threadIdx.x == 0) TC_STAMP(1);
  if (ep.dbg != nullptr && threadIdx.x == 0) {   // grid-wide: earliest moment any CTA got past the dependency wait
    unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
    atomicMin(reinterpret_cast<unsigned long long*>(ep.dbg + 13), g);
  }

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % S::STAGES;
        const uint32_t ph = (kb / S::STAGES) & 1;
        const uint32_t sa = smem_base + s * S::STAGE;
        const bool early = kb < pre_kb;            // expect_tx and the static operand(s) went out before the wait
        if (!early) {
          ptx::mbar_wait(bar_empty + 8 * s, ph ^ 1);
          ptx::mbar_expect_tx(bar_full + 8 * s, S::HB);
        }
        if (!(early && ep.pre_a)) ptx::tma_load_2d(sa, tmap_a_p, (kb0 + kb) * kbk, m0, bar_full + 8 * s);
        if (!(early && ep.pre_b)) ptx::tma_load_2d(sa + S::A_BYTES, tmap_b_p, (kb0 + kb) * kbk, n0, bar_full + 8 * s);
      }
      TC_STAMP(2);
    }
    if (ep.cluster) { __syncwarp(); ptx::cluster_sync_all(); ptx::cluster_sync_all(); }   // the two exchange barriers (below)
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_tf32(kTcBM, BN);
      constexpr uint32_t idesc16 = umma_idesc_bf16(kTcBM, BN);
      const bool b16 = !kSplit && ep.bf16 != 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % S::STAGES;
        const uint32_t ph = (kb / S::STAGES) & 1;
        ptx::mbar_wait((kSplit ? bar_split : bar_full) + 8 * s, ph);
        // TF32X3: every k-block accumulates into a fresh TMEM buffer (two, alternating) that the epilogue warps
        // drain and add in fp32 registers.  The tensor core's own fp32 accumulation rounds toward zero, so a long
        // chain in TMEM loses ~K * 2^-24; keeping each chain to 12 MMAs and summing the partials with round-to-nearest
        // FADDs restores fp32-grade accuracy.
        const int ab = kb & 1;
        if (kSplit) ptx::mbar_wait(bar_aempty + 8 * ab, ((kb >> 1) & 1) ^ 1);
        ptx::tc_fence_after();
        if (kb == 0) TC_STAMP(3);
        const uint32_t sa = smem_base + s * S::STAGE;
        const uint32_t td = tmem_base + (kSplit ? (uint32_t)(ab * BN) : 0u);
        const uint64_t a_hi = umma_desc_sw128(sa), b_hi = umma_desc_sw128(sa + S::A_BYTES);
#pragma unroll
        for (int k = 0; k < kTcBK / 8; ++k) {
          const uint64_t koff = (uint64_t)((k * 8 * 4) >> 4);   // 32 bytes per UMMA_K=8 step, in 16-byte units
          if (kSplit && ep.a_tmem) {
            const uint64_t b_lo = umma_desc_sw128(sa + S::HB + S::A_BYTES);
            const uint32_t ta_hi = tmem_base + (uint32_t)(kTmemABase + s * 64 + k * 8), ta_lo = ta_hi + 32u;
            ptx::mma_tf32_ts(td, ta_lo, b_hi + koff, idesc, k ? 1u : 0u);
            ptx::mma_tf32_ts(td, ta_hi, b_lo + koff, idesc, 1u);
            ptx::mma_tf32_ts(td, ta_hi, b_hi + koff, idesc, 1u);
          } else if (kSplit) {
            const uint64_t a_lo = umma_desc_sw128(sa + S::HB), b_lo = umma_desc_sw128(sa + S::HB + S::A_BYTES);
            ptx::mma_tf32(td, a_lo + koff, b_hi + koff, idesc, k ? 1u : 0u);
            ptx::mma_tf32(td, a_hi + koff, b_lo + koff, idesc, 1u);
            ptx::mma_tf32(td, a_hi + koff, b_hi + koff, idesc, 1u);
          } else if (b16) {                         // 16 bf16 K-values per instruction: the same 32-byte step
            ptx::mma_f16(td, a_hi + koff, b_hi + koff, idesc16, (kb | k) ? 1u : 0u);
          } else {
            ptx::mma_tf32(td, a_hi + koff, b_hi + koff, idesc, (kb | k) ? 1u : 0u);
          }
        }
        ptx::mma_commit(bar_empty + 8 * s);
        if (kSplit) ptx::mma_commit(bar_afull + 8 * ab);
      }
      if (!kSplit) ptx::mma_commit(bar_accum);
      TC_STAMP(4);
    }
    if (ep.cluster) { __syncwarp(); ptx::cluster_sync_all(); ptx::cluster_sync_all(); }
  } else if (kSplit && warp >= 2 + kTcEpiThreads / 32) {
    // ===== TF32X3 drain warps: sum the per-k-block partial accumulators in fp32 registers =====
    const int q = warp & 3;                       // TMEM lane quadrant of this warp
    const int row = q * 32 + lane;
    float pacc[BN];
#pragma unroll
    for (int i = 0; i < BN; ++i) pacc[i] = 0.f;
    for (int j = 0; j < num_kb; ++j) {
      const int ab = j & 1;
      ptx::mbar_wait(bar_afull + 8 * ab, (j >> 1) & 1);
      ptx::tc_fence_after();
#pragma unroll
      for (int c = 0; c < BN / 16; ++c) {
        float v[16];
        ptx::tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * BN + c * 16), v);
#pragma unroll
        for (int i = 0; i < 16; ++i) pacc[c * 16 + i] += v[i];
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar_aempty + 8 * ab);
    }
    // every MMA has retired (the last afull fired), and it was issued only after all split warps finished their last
    // stage: the ring is free -> park the tile where the epilogue warps expect it and release them
    float* T = reinterpret_cast<float*>(smem);
#pragma unroll
    for (int c = 0; c < BN / 4; ++c)
      *reinterpret_cast<float4*>(T + row * S::TILE_LD + 4 * c) = make_float4(pacc[4 * c], pacc[4 * c + 1], pacc[4 * c + 2], pacc[4 * c + 3]);
    asm volatile("bar.sync 2, 384;" ::: "memory");
    if (ep.cluster) { ptx::cluster_sync_all(); ptx::cluster_sync_all(); }
  } else {
    // ===== split (TF32X3) + epilogue warps =====
    const int et = threadIdx.x - 64;              // 0..255
    const int q = warp & 3;                       // TMEM lane quadrant this warp may read
    const int chalf = (warp - 2) >> 2;            // which half of the accumulator columns this warp drains
    const int row = q * 32 + lane;                // accumulator row held by this thread
    // split-K cluster: the row-wise epilogues are shared out among the cluster's CTAs by epilogue pass (see below)
    constexpr bool kDist = kEpi != EPI_REPARAM && kEpi != EPI_REPARAM_BWD;
    const bool dist = kDist && ep.cluster && gridDim.z > 1;
    const int it_rank = dist ? (int)blockIdx.z : 0, it_ns = dist ? (int)gridDim.z : 1;
    tc05_prefetch<BN, kEpi>(ep, m0, n0, M, N, et, pre, it_rank, it_ns);
    // bias gradient of the weight-gradient GEMMs: row sums of the A operand (fixed order), done first: it overlaps the latency of the first TMA loads
    if (kEpi == EPI_GRADW && blockIdx.x == 0 && blockIdx.z == 0 && ep.db != nullptr) {
      // one warp per row, lanes stride the batch dimension (coalesced), fixed-shape shuffle tree: deterministic
      constexpr int RW = kTcBM / (kTcEpiThreads / 32);      // rows per warp (16)
      const int ew = et >> 5;
      const float* const abase = ep.a_src + (size_t)(m0 + ew * RW) * ep.lda_src;
      const size_t lda = ep.lda_src;
      const int rows_here = min(RW, M - m0 - ew * RW);
      float acc[RW];
#pragma unroll
      for (int i = 0; i < RW; ++i) acc[i] = 0.f;
      if (!kSplit && ep.bf16) {
        // the A operand is bf16 here: 8 values per 16-byte load (rows are 16-byte aligned: lda % 8 == 0)
        const __nv_bfloat16* const ab = reinterpret_cast<const __nv_bfloat16*>(ep.a_src) + (size_t)(m0 + ew * RW) * ep.lda_src;
        const int K8 = K & ~7;
        for (int k = 8 * lane; k < K8; k += 256) {
          uint4 t[RW];
#pragma unroll
          for (int i = 0; i < RW; ++i)
            t[i] = (i < rows_here) ? __ldg(reinterpret_cast<const uint4*>(ab + i * lda + k)) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
          for (int i = 0; i < RW; ++i) {
            const uint32_t w[4] = {t[i].x, t[i].y, t[i].z, t[i].w};
            float s8 = 0.f;
#pragma unroll
            for (int q = 0; q < 4; ++q) s8 += __uint_as_float(w[q] << 16) + __uint_as_float(w[q] & 0xffff0000u);
            acc[i] += s8;
          }
        }
        for (int k = K8 + lane; k < K; k += 32) {
#pragma unroll
          for (int i = 0; i < RW; ++i) acc[i] += (i < rows_here) ? __bfloat162float(ab[i * lda + k]) : 0.f;
        }
      } else {
      const int K4 = K & ~3;                               // lda is a multiple of 4 floats: rows are 16-byte aligned
      for (int k = 4 * lane; k < K4; k += 128) {           // RW independent 16-byte loads in flight per step
        float4 t[RW];
#pragma unroll
        for (int i = 0; i < RW; ++i)
          t[i] = (i < rows_here) ? __ldg(reinterpret_cast<const float4*>(abase + i * lda + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int i = 0; i < RW; ++i) acc[i] += (t[i].x + t[i].y) + (t[i].z + t[i].w);
      }
      for (int k = K4 + lane; k < K; k += 32) {            // ragged batch tail
#pragma unroll
        for (int i = 0; i < RW; ++i) acc[i] += (i < rows_here) ? __ldg(abase + i * lda + k) : 0.f;
      }
      }
      float mine = 0.f;                                   // lane i keeps the sum of row i
#pragma unroll
      for (int i = 0; i < RW; ++i) {
        const float t = warp_sum(acc[i]);
        if (lane == i) mine = t;
      }
      if (lane < rows_here) {
        const int m = m0 + ew * RW + lane;
        ep.db[m] = mine;
        if (ep.adam_pb != nullptr) {                        // fused Adam on the bias element
          float pb = ep.adam_pb[m], mb = ep.adam_mb[m], vb = ep.adam_vb[m];
          adam_elem_fast(pb, mine, mb, vb, ep.adam_st->next[0], __frcp_rn(ep.adam_st->next[1]), ep.adam_st->next[2],
                         ep.adam_st->next[3], ep.adam_st->next[4], ep.adam_st->next[5]);
          ep.adam_pb[m] = pb; ep.adam_mb[m] = mb; ep.adam_vb[m] = vb;
        }
      }
    }

    constexpr int CH = BN / 2 < 16 ? 16 : BN / 2; // accumulator columns per warp (BN = 16: the second warp of a quadrant idles)
    const int colw = chalf * CH;                  // first column of this warp
    const bool drains = colw < BN;
    if (kSplit) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % S::STAGES;
        const uint32_t ph = (kb / S::STAGES) & 1;
        ptx::mbar_wait(bar_full + 8 * s, ph);
        float4* hi = reinterpret_cast<float4*>(smem + s * S::STAGE);
        float4* lo = reinterpret_cast<float4*>(smem + s * S::STAGE + S::HB);
        if (ep.a_tmem) {
          // A tile -> tensor memory.  Thread = accumulator-style (lane quadrant q, row q*32+lane); the two warps of a
          // quadrant take 16 of the 32 K-columns each.  Row r of the SWIZZLE_128B tile: 16-byte chunk j sits at
          // r*128 + ((j ^ (r & 7)) << 4).  hi is written explicitly truncated, lo = a - hi (exact).
          const uint8_t* arow = smem + s * S::STAGE + row * 128;
          float vh[16], vl[16];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int chunk = 4 * chalf + j;
            const float4 v = *reinterpret_cast<const float4*>(arow + ((chunk ^ (row & 7)) << 4));
            vh[4 * j + 0] = ptx::tf32_hi(v.x); vh[4 * j + 1] = ptx::tf32_hi(v.y);
            vh[4 * j + 2] = ptx::tf32_hi(v.z); vh[4 * j + 3] = ptx::tf32_hi(v.w);
            vl[4 * j + 0] = v.x - vh[4 * j + 0]; vl[4 * j + 1] = v.y - vh[4 * j + 1];
            vl[4 * j + 2] = v.z - vh[4 * j + 2]; vl[4 * j + 3] = v.w - vh[4 * j + 3];
          }
          const uint32_t ta = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(kTmemABase + s * 64 + 16 * chalf);
          ptx::tmem_st16(ta, vh);
          ptx::tmem_st16(ta + 32u, vl);
          // B tile: lo next to hi in shared memory, as before
          for (int c = S::A_BYTES / 16 + et; c < S::HB / 16; c += kTcEpiThreads) {
            const float4 v = hi[c];
            float4 l;
            l.x = v.x - ptx::tf32_hi(v.x); l.y = v.y - ptx::tf32_hi(v.y);
            l.z = v.z - ptx::tf32_hi(v.z); l.w = v.w - ptx::tf32_hi(v.w);
            lo[c] = l;
          }
          ptx::tmem_wait_st();
          ptx::tc_fence_before();
        } else
#pragma unroll 4
        for (int c = et; c < S::HB / 16; c += kTcEpiThreads) {
          // The tensor core reads only sign, exponent and the top 10 mantissa bits of an fp32 word (it truncates),
          // so the raw tile already IS the "hi" operand; only lo = a - trunc_tf32(a) (exact in fp32) is written.
          const float4 v = hi[c];
          float4 l;
          l.x = v.x - ptx::tf32_hi(v.x); l.y = v.y - ptx::tf32_hi(v.y);
          l.z = v.z - ptx::tf32_hi(v.z); l.w = v.w - ptx::tf32_hi(v.w);
          lo[c] = l;
        }
        ptx::fence_proxy_async();                 // generic-proxy writes -> visible to the tensor core's async proxy
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(bar_split + 8 * s);
        if (gridDim.z == 1 && kb < 4 && et == 0) TC_STAMP(9 + kb);
      }
    }
    if (!kSplit) {
      ptx::mbar_wait(bar_accum, 0);
      ptx::tc_fence_after();
    }
    if (et == 0) TC_STAMP(5);
    // accumulator -> padded smem tile (the ring is free: every MMA has retired and, in TF32X3 mode, every warp has
    // passed its last split; the barrier below orders the tile writes against stragglers' lo-writes of the last stage)
    float* T = reinterpret_cast<float*>(smem);
    if (kSplit) {
      asm volatile("bar.sync 2, 384;" ::: "memory");   // the drain warps have parked the summed tile in T
    } else {
      if (drains) {
#pragma unroll
        for (int c = 0; c < CH / 16; ++c) {
          float v[16];
          const int col0 = colw + c * 16;
          ptx::tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)col0, v);
          float4* dst = reinterpret_cast<float4*>(T + row * S::TILE_LD + col0);
#pragma unroll
          for (int j = 0; j < 4; ++j) dst[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
      }
      ptx::bar_sync_epi();
    }
    if (et == 0) TC_STAMP(6);
    bool run_epilogue = true;
    if (gridDim.z > 1 && ep.cluster) {
      // Split-K inside a thread-block cluster: the splits of a tile are the CTAs of one cluster.  Every CTA has parked
      // its partial tile at the same shared-memory offset.  After a cluster barrier the partials are summed through
      // distributed shared memory in fixed rank order (deterministic); a second barrier keeps every CTA's tile alive
      // until its peers have read it.  No workspace round trip through L2, no fence + ticket chain.
      //  * row-wise epilogues: CTA r sums, into its own tile, the rows of the epilogue passes i = r, r + splits, ...
      //    (a pass = the 256 float4 that the 256 epilogue threads handle together) and then runs the epilogue for just
      //    those passes -- the reduction AND the epilogue (sigmoid/BCE math, stores) are spread over the cluster;
      //  * reparameterisation epilogues (whole-tile structure): every CTA sums a 1/splits slice into rank 0's tile,
      //    rank 0 runs the epilogue.
      constexpr int F4 = kTcBM * BN / 4;
      constexpr int ITERS_E = F4 / kTcEpiThreads;
      const uint32_t ns = gridDim.z, rank = blockIdx.z;
      ptx::cluster_sync_all();
      if (ep.dbg && et == 0 && blockIdx.x == 0 && by == 0 && rank == 0) { unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); ep.dbg[9] = (long long)g; }
      if (kDist) {
        for (int i = (int)rank; i < ITERS_E; i += (int)ns) {
          const int f = i * kTcEpiThreads + et;
          float* const mine = T + (f / (BN / 4)) * S::TILE_LD + (f % (BN / 4)) * 4;
          const uint32_t off = ptx::smem_u32(mine);
          float4 v[8];
#pragma unroll
          for (uint32_t r = 0; r < 8; ++r)
            v[r] = (r < ns) ? ptx::cluster_ld4(ptx::cluster_map(off, r)) : make_float4(0.f, 0.f, 0.f, 0.f);
          float4 acc = v[0];
#pragma unroll
          for (int r = 1; r < 8; ++r) { acc.x += v[r].x; acc.y += v[r].y; acc.z += v[r].z; acc.w += v[r].w; }
          *reinterpret_cast<float4*>(mine) = acc;    // only this thread reads this element again before the next barrier
        }
        // NB the peers read this CTA's ORIGINAL partial of the rows they own, never the rows it just overwrote
      } else {
        const int chunk = (F4 + (int)ns - 1) / (int)ns;
        const int f_hi = min(F4, ((int)rank + 1) * chunk);
        for (int f = (int)rank * chunk + et; f < f_hi; f += kTcEpiThreads) {
          const uint32_t off = ptx::smem_u32(T + (f / (BN / 4)) * S::TILE_LD + (f % (BN / 4)) * 4);
          float4 v[8];
#pragma unroll
          for (uint32_t r = 0; r < 8; ++r)
            v[r] = (r < ns) ? ptx::cluster_ld4(ptx::cluster_map(off, r)) : make_float4(0.f, 0.f, 0.f, 0.f);
          float4 acc = v[0];
#pragma unroll
          for (int r = 1; r < 8; ++r) { acc.x += v[r].x; acc.y += v[r].y; acc.z += v[r].z; acc.w += v[r].w; }
          ptx::cluster_st4(ptx::cluster_map(off, 0), acc);
        }
      }
      ptx::cluster_sync_all();
      if (ep.dbg && et == 0 && blockIdx.x == 0 && by == 0 && rank == 0) { unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); ep.dbg[10] = (long long)g; }
      run_epilogue = kDist || rank == 0;
    } else if (gridDim.z > 1) {
      // Deterministic split-K reduction: every split parks its partial tile in the workspace; the CTA that takes the
      // last ticket of the tile adds the partials in split order (independent of arrival order) and runs the epilogue.
      constexpr int F4 = kTcBM * BN / 4;                       // float4 per tile
      constexpr int PER = (F4 + kTcEpiThreads - 1) / kTcEpiThreads;
      const int tile_id = by * gridDim.x + blockIdx.x;
      const int nsplit = gridDim.z;
      float4* const ws = reinterpret_cast<float4*>(ep.ws) + (size_t)tile_id * nsplit * F4;
      float4* const mine = ws + (size_t)blockIdx.z * F4;
#pragma unroll
      for (int i = 0; i < PER; ++i) {
        const int f = et + i * kTcEpiThreads;
        if (f < F4) mine[f] = *reinterpret_cast<const float4*>(T + (f / (BN / 4)) * S::TILE_LD + (f % (BN / 4)) * 4);
      }
      __threadfence();
      ptx::bar_sync_epi();
      if (et == 0) {
        const unsigned int t = atomicAdd(ep.ws_counters + tile_id, 1u);
        const bool last = (t == (unsigned)nsplit - 1);
        if (last) ep.ws_counters[tile_id] = 0;                 // ready for the next launch
        __threadfence();
        red[8] = last ? 1.f : 0.f;
      }
      ptx::bar_sync_epi();
      run_epilogue = red[8] != 0.f;
      if (run_epilogue) {
        if (ep.dbg && et == 0 && blockIdx.x == 0 && by == 0) { unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); ep.dbg[9] = (long long)g; }
        float4 acc[PER];
#pragma unroll
        for (int i = 0; i < PER; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        constexpr int ZG = PER <= 2 ? 8 : (PER <= 4 ? 4 : 2);   // splits fetched together: their L2 latencies overlap
        for (int z0 = 0; z0 < nsplit; z0 += ZG) {
          float4 v[ZG][PER];
#pragma unroll
          for (int zz = 0; zz < ZG; ++zz)
#pragma unroll
            for (int i = 0; i < PER; ++i) {
              const int f = et + i * kTcEpiThreads;
              v[zz][i] = (z0 + zz < nsplit && f < F4) ? __ldcg(ws + (size_t)(z0 + zz) * F4 + f) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
          for (int zz = 0; zz < ZG; ++zz)        // still added strictly in split order: deterministic
#pragma unroll
            for (int i = 0; i < PER; ++i) {
              acc[i].x += v[zz][i].x; acc[i].y += v[zz][i].y; acc[i].z += v[zz][i].z; acc[i].w += v[zz][i].w;
            }
        }
#pragma unroll
        for (int i = 0; i < PER; ++i) {
          const int f = et + i * kTcEpiThreads;
          if (f < F4) *reinterpret_cast<float4*>(T + (f / (BN / 4)) * S::TILE_LD + (f % (BN / 4)) * 4) = acc[i];
        }
        ptx::bar_sync_epi();
        if (ep.dbg && et == 0 && blockIdx.x == 0 && by == 0) { unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); ep.dbg[10] = (long long)g; }
      }
    }
    if (run_epilogue) {
      tc05_epilogue<BN, kEpi>(ep, T, S::TILE_LD, m0, n0, M, N, et, red, pre, it_rank, it_ns, by);
      if (ep.dbg && et == 0 && blockIdx.x == 0 && by == 0) { unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); ep.dbg[11] = (long long)g; }
    }
    if (et == 0) TC_STAMP(7);
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
  if (kEpi == EPI_GRADW && ep.adam_total_ctas > 0)
    adam_step_ticket(ep.adam_st, ep.adam_total_ctas, ep.loss_part_a, ep.n_part_a, ep.loss_part_b, ep.n_part_b, ep.adam_loss_out,
                     reinterpret_cast<float*>(smem));
  if (threadIdx.x == 0) {
    TC_STAMP(8);
    if (ep.dbg != nullptr) {                     // grid-wide: the moment the last CTA left
      unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
      atomicMax(reinterpret_cast<unsigned long long*>(ep.dbg + 15), g);
    }
  }
#undef TC_STAMP
}

template <int BN, bool kSplit, int kEpi>
__global__ void __launch_bounds__(tc_threads(kSplit), 1)
tc05_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, int M, int N,
                 int K, Tc05Epi ep) {
  tc05_gemm_body<BN, kSplit, kEpi>(&tmap_a, &tmap_b, M, N, K, ep, (int)blockIdx.y);
}

// Grouped launch: the same GEMM of T independent trials in ONE launch.  blockIdx.y = trial * tiles_m + tile;
// tensor maps and epilogue arguments of trial t come from table[t] in global memory (written by the host when the
// group's graph is built); the epilogue arguments are copied into shared memory once per CTA.
struct alignas(64) Tc05GroupEntry {
  CUtensorMap a, b;
  Tc05Epi ep;
};
constexpr int kTcGroupEpiOff = 512;     // byte offset of the Tc05Epi copy behind the barrier block (1280 bytes are reserved there)
static_assert(sizeof(Tc05Epi) % 4 == 0 && kTcGroupEpiOff + sizeof(Tc05Epi) <= 256 + 1024, "Tc05Epi copy must fit behind the barriers");

template <int BN, bool kSplit, int kEpi>
__global__ void __launch_bounds__(tc_threads(kSplit), 1)
tc05_gemm_group_kernel(const Tc05GroupEntry* __restrict__ table, int tiles_m, int M, int N, int K) {
  using S = Tc05Smem<BN, kSplit>;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int trial = (int)blockIdx.y / tiles_m, by = (int)blockIdx.y - trial * tiles_m;
  const Tc05GroupEntry* const en = table + trial;
  uint32_t* const dst = reinterpret_cast<uint32_t*>(smem + S::BAR_OFF + kTcGroupEpiOff);
  const uint32_t* const src = reinterpret_cast<const uint32_t*>(&en->ep);
  for (int i = threadIdx.x; i < (int)(sizeof(Tc05Epi) / 4); i += blockDim.x) dst[i] = src[i];
  __syncthreads();
  tc05_gemm_body<BN, kSplit, kEpi>(&en->a, &en->b, M, N, K, *reinterpret_cast<const Tc05Epi*>(dst), by);
}

// ------------------------------------------------------------------------------------------------
// tiled transpose (x -> x^T per step, weight shadows after Adam, da4 -> da4^T in compat mode)
// ------------------------------------------------------------------------------------------------
struct TrJob {
  const float* in;   // [R][ld_in]
  void* out;         // [C][ld_out] transposed copy, fp32 (or bf16 when out_is16), may be NULL
  int R, C, ld_in, ld_out;
  int tile_begin, tiles_c;
  // MDT_GEMM_BF16 operand mirrors written in the same pass (NULL when not wanted)
  int out_is16;
  __nv_bfloat16* out_t16;              // second transposed copy, bf16, leading dimension ld_out
  __nv_bfloat16* copy16; int ld_copy16;   // row-major bf16 copy [R][ld_copy16]
};
struct TrArgs {
  TrJob job[4];
  int njobs;
};

// One 32 x 32 tile of a job: t[][] holds the tile; writes whatever outputs the job asks for.
__device__ __forceinline__ void tr_store_tile(const TrJob& job, float (*t)[33], int r0, int c0, int tx, int ty) {
#pragma unroll
  for (int i = ty; i < 32; i += 8) {
    const int c = c0 + i, r = r0 + tx;
    if (c < job.C && r < job.R) {
      const float v = t[tx][i];
      if (job.out != nullptr) st_operand(job.out, (size_t)c * job.ld_out + r, v, job.out_is16 != 0);
      if (job.out_t16 != nullptr) job.out_t16[(size_t)c * job.ld_out + r] = __float2bfloat16_rn(v);
    }
  }
  if (job.copy16 != nullptr) {
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
      const int r = r0 + i, c = c0 + tx;
      if (r < job.R && c < job.C) job.copy16[(size_t)r * job.ld_copy16 + c] = __float2bfloat16_rn(t[i][tx]);
    }
  }
}

__global__ void __launch_bounds__(256) transpose_kernel(TrArgs a) {
  __shared__ float t[32][33];
  pdl_wait();
  pdl_trigger();
  int j = 0;
#pragma unroll
  for (int q = 1; q < 4; ++q)
    if (q < a.njobs && (int)blockIdx.x >= a.job[q].tile_begin) j = q;
  const TrJob job = a.job[j];
  const int tile = (int)blockIdx.x - job.tile_begin;
  const int r0 = (tile / job.tiles_c) * 32, c0 = (tile % job.tiles_c) * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int i = ty; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + tx;
    t[i][tx] = (r < job.R && c < job.C) ? job.in[(size_t)r * job.ld_in + c] : 0.f;
  }
  __syncthreads();
  tr_store_tile(job, t, r0, c0, tx, ty);
}

// rows of an fp32 matrix -> bf16 (leading dimensions may differ): mdt_decode's z in bf16 mode
__global__ void __launch_bounds__(256) rows_to_bf16_kernel(const float* in, int ld_in, __nv_bfloat16* out, int ld_out, int R, int C) {
  pdl_wait();
  pdl_trigger();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R * C; i += gridDim.x * blockDim.x) {
    const int r = i / C, c = i - r * C;
    out[(size_t)r * ld_out + c] = __float2bfloat16_rn(in[(size_t)r * ld_in + c]);
  }
}

// Step prologue: x -> x^T (tiles) and, in the blocks after the tiles, the step's eps drawn from Philox exactly
// as torch.randn_like would on CUDA (ref:vae-hpo.py:35); the last eps block advances the device-resident offset.
struct PrepArgs {
  TrArgs tr;
  int tr_blocks;
  float* x_copy;       // non-NULL: x lives in the caller's buffer; also copy it (row-major, ld = C) into the stage
                       // buffer the later kernels read.  The engine re-points job[0].in / x_copy per launch of a
                       // cached graph (kernel-node parameter update), so no separate staging copy is needed.
  float* eps_out;      // NULL: caller supplied eps, nothing to generate
  long long n_eps;
  StepState* st;
  long long S;
  unsigned long long offset_inc;
  // mdt_train_steps: x is batch (StepState.x_cursor % n_batches) of a device-resident source -- either a pool of fp32
  // batches or a uint8 image table gathered through the sampler's index order (value / 255, as ToTensor()).  The
  // index is read from device memory, so the graphs that hold this kernel are launched unmodified step after step.
  const float* pool;
  const uint8_t* images;
  const long long* indices;
  long long n_batches, n_images;
  int cursor_add;      // 1: this launch fetches the NEXT step's batch while the current step runs (see Tc05Plan::pipelined)
};

__device__ __forceinline__ void prep_body(const PrepArgs& a, const int bx, const int gx) {
  pdl_wait();
  pdl_trigger();
  if (bx < a.tr_blocks) {
    __shared__ float t[32][33];
    TrJob job = a.tr.job[0];
    const int tile = bx;
    const int r0 = (tile / job.tiles_c) * 32, c0 = (tile % job.tiles_c) * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const bool sourced = a.pool != nullptr || a.images != nullptr;
    long long bidx = 0, cur_next = 0;
    if (sourced) {
      // this step's batch (x_cursor), or -- fetched ahead while a step runs -- the one after the last staged batch
      // (x_cursor_pf): the running step's last CTA advances x_cursor at an unknown time relative to this kernel
      const long long cur = a.cursor_add ? a.st->x_cursor_pf : a.st->x_cursor;
      bidx = cur % a.n_batches;
      if (a.pool != nullptr) job.in = a.pool + (size_t)bidx * job.R * job.C;
      cur_next = cur + 1;
    }
#pragma unroll
    for (int i = ty; i < 32; i += 8) {
      const int r = r0 + i, c = c0 + tx;
      const bool ok = r < job.R && c < job.C;
      float v = 0.f;
      if (ok) {
        if (a.images != nullptr && a.pool == nullptr) {
          long long src = a.indices[bidx * job.R + r];
          if (src < 0 || src >= a.n_images) src = 0;                    // never read out of bounds (host validates)
          v = __fdiv_rn((float)a.images[(size_t)src * job.C + c], 255.0f);   // IEEE division: bit-identical to ToTensor()
        } else {
          v = job.in[(size_t)r * job.ld_in + c];
        }
      }
      t[i][tx] = v;
      if (ok && a.x_copy != nullptr) a.x_copy[(size_t)r * job.C + c] = v;
    }
    __syncthreads();
    tr_store_tile(job, t, r0, c0, tx, ty);
    if (sourced && threadIdx.x == 0) {       // every block has read its cursor by the time the last ticket is taken
      __threadfence();
      if (atomicAdd(&a.st->ticket_prep, 1u) == (unsigned)a.tr_blocks - 1u) {
        a.st->x_cursor_pf = cur_next;
        a.st->ticket_prep = 0;
        __threadfence();
      }
    }
    return;
  }
  const unsigned long long seed = a.st->philox_seed, offset = a.st->philox_offset;
  const int nb = gx - a.tr_blocks;
  for (long long i = (long long)(bx - a.tr_blocks) * blockDim.x + threadIdx.x; i < a.n_eps;
       i += (long long)nb * blockDim.x)
    a.eps_out[i] = torch_cuda_randn_elem(seed, offset, i, a.S);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&a.st->ticket_reparam, 1u) == (unsigned)nb - 1) {
      a.st->philox_offset = offset + a.offset_inc;
      a.st->ticket_reparam = 0;
      __threadfence();
    }
  }
}

__global__ void __launch_bounds__(256) prep_kernel(PrepArgs a) { prep_body(a, (int)blockIdx.x, (int)gridDim.x); }
// Grouped launch: blockIdx.x = trial * blocks_per_trial + block; arguments of trial t in table[t].
__global__ void __launch_bounds__(256) prep_group_kernel(const PrepArgs* __restrict__ table, int blocks_per_trial) {
  const int trial = (int)blockIdx.x / blocks_per_trial;
  const PrepArgs a = table[trial];
  prep_body(a, (int)blockIdx.x - trial * blocks_per_trial, blocks_per_trial);
}

inline int tr_add(TrArgs& a, int& blocks, const float* in, void* out, int R, int C, int ld_in, int ld_out, bool out_is16 = false,
                  __nv_bfloat16* out_t16 = nullptr, __nv_bfloat16* copy16 = nullptr, int ld_copy16 = 0) {
  TrJob& j = a.job[a.njobs++];
  j.in = in; j.out = out; j.R = R; j.C = C; j.ld_in = ld_in; j.ld_out = ld_out;
  j.out_is16 = out_is16 ? 1 : 0; j.out_t16 = out_t16; j.copy16 = copy16; j.ld_copy16 = ld_copy16;
  j.tile_begin = blocks; j.tiles_c = (C + 31) / 32;
  blocks += ((R + 31) / 32) * j.tiles_c;
  return 0;
}

// ------------------------------------------------------------------------------------------------
// host side: tensor maps, launchers, the per-engine plan
// ------------------------------------------------------------------------------------------------
// ------------------------------------------------------------------------------------------------
// Diagnostic switches (MDT_* environment variables).  They are read ONCE, when an engine's tensor-core plan is created
// (tc05_create), into the plan; the enqueue entry points install the plan's copy for the launch helpers below, which
// have no plan argument.  Nothing on the launch path calls getenv.
// ------------------------------------------------------------------------------------------------
struct TcKnobs {
  bool splitk = true;          // MDT_NO_SPLITK=1 clears
  bool cluster_splitk = true;  // MDT_NO_CLUSTER_SPLITK=1 clears: split-K partials through a workspace instead of DSMEM
  bool a_tmem = false;         // MDT_A_TMEM=1: TF32X3 with the A operand in tensor memory
  int min_kb = 0;              // MDT_MIN_KB_PER_SPLIT (0: 3 with clusters, 4 without)
  int max_splits = 0;          // MDT_MAX_SPLITS (0: no extra cap)
  bool w4_split = false;       // MDT_W4_SPLIT=1: fc4's Adam update as a separate tile kernel
  int bn_dw4 = 0, bn_dw1 = 0;  // MDT_BN_DW4 / MDT_BN_DW1: tile width of fc4's / fc1's fused-Adam GEMM
  bool dw4_late = false;       // MDT_DW4_LATE=1: all four updates start together after latent_bwd
  bool dw3_early = false;      // MDT_DW3_EARLY=1: fc3's update starts with fc4's
  bool dw1_side = false;       // MDT_DW1_SIDE=1: fc1's update on a side stream
  int latent_sc = 0;           // MDT_LATENT_SC: samples per CTA of the CUDA-core latent kernels (1 / 2 / 4; 0: choose)
  static TcKnobs from_env() {
    TcKnobs k;
    k.splitk = getenv("MDT_NO_SPLITK") == nullptr;
    k.cluster_splitk = getenv("MDT_NO_CLUSTER_SPLITK") == nullptr;
    if (const char* v = getenv("MDT_A_TMEM")) k.a_tmem = v[0] == '1';
    if (const char* v = getenv("MDT_MIN_KB_PER_SPLIT")) k.min_kb = atoi(v);
    if (const char* v = getenv("MDT_MAX_SPLITS")) k.max_splits = atoi(v);
    k.w4_split = getenv("MDT_W4_SPLIT") != nullptr;
    if (const char* v = getenv("MDT_BN_DW4")) k.bn_dw4 = atoi(v);
    if (const char* v = getenv("MDT_BN_DW1")) k.bn_dw1 = atoi(v);
    k.dw4_late = getenv("MDT_DW4_LATE") != nullptr;
    k.dw3_early = getenv("MDT_DW3_EARLY") != nullptr;
    k.dw1_side = getenv("MDT_DW1_SIDE") != nullptr;
    if (const char* v = getenv("MDT_LATENT_SC")) k.latent_sc = atoi(v);
    return k;
  }
};
inline const TcKnobs*& tc_knobs_ptr() {
  static thread_local const TcKnobs* k = nullptr;
  return k;
}
inline const TcKnobs& tc_knobs() {
  static const TcKnobs defaults{};       // launches outside any plan (none today) get the built-in defaults
  return tc_knobs_ptr() ? *tc_knobs_ptr() : defaults;
}
struct TcKnobScope {
  const TcKnobs* prev;
  explicit TcKnobScope(const TcKnobs* k) : prev(tc_knobs_ptr()) { tc_knobs_ptr() = k; }
  ~TcKnobScope() { tc_knobs_ptr() = prev; }
};

// ------------------------------------------------------------------------------------------------
// Launch recorder (grouped launches of several trials).  While a recorder is installed, the step's enqueue
// code launches nothing: every kernel launch and every event record / stream wait is appended to the recorder instead.
// The group (mdt_api.cu) records the same step for each of its trials, checks that the T records are congruent, and
// replays ONE of them with every launch replaced by the grouped kernel over a table of the T trials' arguments.
// ------------------------------------------------------------------------------------------------
struct TcOp {
  enum Kind { GEMM, LAT_FWD, LAT_BWD, PREP, EV_RECORD, STREAM_WAIT } kind;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev = nullptr;
  // GEMM
  int bn = 0, split = 0, epi = 0, M = 0, N = 0, K = 0, nsplit = 1, use_clusters = 0;
  CUtensorMap a{}, b{};
  Tc05Epi ep{};
  // latent kernels / prologue
  LatentFwdArgs lf{};
  LatentBwdArgs lb{};
  int sc = 0;
  PrepArgs pa{};
  int blocks = 0, programmatic = 1;
};
struct TcRecorder {
  std::vector<TcOp> ops;
};
inline TcRecorder*& tc_recorder() {
  static thread_local TcRecorder* r = nullptr;
  return r;
}
inline cudaError_t tc_event_record(cudaEvent_t ev, cudaStream_t s) {
  if (TcRecorder* r = tc_recorder()) {
    TcOp op; op.kind = TcOp::EV_RECORD; op.ev = ev; op.stream = s;
    r->ops.push_back(op);
    return cudaSuccess;
  }
  return cudaEventRecord(ev, s);
}
inline cudaError_t tc_stream_wait(cudaStream_t s, cudaEvent_t ev) {
  if (TcRecorder* r = tc_recorder()) {
    TcOp op; op.kind = TcOp::STREAM_WAIT; op.ev = ev; op.stream = s;
    r->ops.push_back(op);
    return cudaSuccess;
  }
  return cudaStreamWaitEvent(s, ev, 0);
}

typedef CUresult (*PFN_mdt_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline PFN_mdt_encodeTiled tc05_encode_fn() {
  static PFN_mdt_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_mdt_encodeTiled>(p);
  }
  return fn;
}

// K-major matrix [rows][K] with leading dimension ld (elements), fp32 or bf16; box = box_rows x one 128-byte swizzle row
// (32 fp32 / 64 bf16), SWIZZLE_128B.
inline int tc05_encode(CUtensorMap* out, const void* ptr, int rows, int K, int ld, int box_rows, bool bf16 = false) {
  PFN_mdt_encodeTiled fn = tc05_encode_fn();
  if (!fn) return set_error(MDT_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled is not available from this driver");
  const int esz = bf16 ? 2 : 4;
  MDT_CHECK(rows > 0 && K > 0 && ld >= K, "tensor map: bad shape rows=%d K=%d ld=%d", rows, K, ld);
  MDT_CHECK(reinterpret_cast<uintptr_t>(ptr) % 16 == 0 && ((size_t)ld * esz) % 16 == 0, "tensor map: base and row pitch must be 16-byte aligned");
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * esz};
  cuuint32_t box[2] = {(cuuint32_t)(128 / esz), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), dims,
                  strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(MDT_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return MDT_OK;
}

// Phase-timing buffer (MDT_TC05_DBG=1): 32 slots x 16 clock64 stamps; launch i writes slot i % 32 (CTA (0,0) only).
inline int& tc05_dbg_launch_index() { static int idx = 0; return idx; }
inline long long* tc05_dbg_buffer() {
  static long long* buf = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    const char* e = getenv("MDT_TC05_DBG");
    if (e && e[0] == '1' && cudaMalloc(&buf, 512 * sizeof(long long)) == cudaSuccess) cudaMemset(buf, 0, 512 * sizeof(long long));
  }
  return buf;
}

template <int BN, bool kSplit, int kEpi>
inline int tc05_launch_t(const CUtensorMap& a, const CUtensorMap& b, int M, int N, int K, const Tc05Epi& ep, cudaStream_t s,
                         int nsplit) {
  using S = Tc05Smem<BN, kSplit>;
  static bool attr = false;
  if (!attr) {
    MDT_CUDA(cudaFuncSetAttribute(tc05_gemm_kernel<BN, kSplit, kEpi>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
    attr = true;
  }
  dim3 grid((N + BN - 1) / BN, (M + kTcBM - 1) / kTcBM, nsplit);
  Tc05Epi epd = ep;
  // split-K: the splits of a tile exchange their partials through distributed shared memory when they fit one cluster
  const bool use_clusters = tc_knobs().cluster_splitk;
  if (TcRecorder* r = tc_recorder()) {
    TcOp op; op.kind = TcOp::GEMM; op.stream = s; op.bn = BN; op.split = kSplit ? 1 : 0; op.epi = kEpi;
    op.M = M; op.N = N; op.K = K; op.nsplit = nsplit; op.use_clusters = use_clusters ? 1 : 0; op.a = a; op.b = b; op.ep = epd;
    r->ops.push_back(op);
    return MDT_OK;
  }
  epd.dbg = tc05_dbg_buffer();
  if (epd.dbg) epd.dbg += 16 * (tc05_dbg_launch_index()++ % 32);
  // MDT_A_TMEM=1: TF32X3 with the A operand in tensor memory (see Tc05Epi::a_tmem).  Correct, and measured equal in
  // speed to the shared-memory form (57.4 vs 57.5 us/step): the k-block is not bound by the MMAs' shared-memory reads of A.
  epd.a_tmem = (kSplit && tc_knobs().a_tmem) ? 1 : 0;
  epd.cluster = (nsplit > 1 && nsplit <= 8 && use_clusters) ? 1 : 0;
  if (epd.cluster) {
    // a cluster needs its CTAs on SMs of one GPC at the same time: when the tiles do not all fit as co-resident
    // clusters the launch runs in two waves (measured: fc4 forward, 49 tiles x 3 splits, 8.4 -> 14 us) -- keep the
    // workspace path then
    static int max_clusters[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (max_clusters[nsplit] == 0) {
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(1, 1, nsplit); cfg.blockDim = dim3(tc_threads(kSplit)); cfg.dynamicSmemBytes = S::TOTAL;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = nsplit;
      cfg.attrs = at; cfg.numAttrs = 1;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, tc05_gemm_kernel<BN, kSplit, kEpi>, &cfg) != cudaSuccess) { n = -1; cudaGetLastError(); }
      max_clusters[nsplit] = n > 0 ? n : -1;
    }
    if (max_clusters[nsplit] < (int)(grid.x * grid.y)) epd.cluster = 0;
  }
  if (epd.cluster)
    MDT_CUDA(launch_pdl_cluster(tc05_gemm_kernel<BN, kSplit, kEpi>, grid, dim3(tc_threads(kSplit)), (size_t)S::TOTAL, s,
                                (unsigned)nsplit, a, b, M, N, K, epd));
  else
    MDT_CUDA(launch_pdl(tc05_gemm_kernel<BN, kSplit, kEpi>, grid, dim3(tc_threads(kSplit)), (size_t)S::TOTAL, s, a, b, M, N, K, epd));
  return MDT_OK;
}

// Number of K-splits for a GEMM: only when a workspace is given, at least min_kb k-blocks (of 32 K-values) per split, and
// no more CTAs than SMs (one wave).
inline int tc05_pick_splits(int M, int N, int K, int bn, int sm_count, const Tc05Epi& ep, size_t ws_floats) {
  if (ep.ws == nullptr || ep.ws_counters == nullptr) return 1;
  const int tiles = ((N + bn - 1) / bn) * ((M + kTcBM - 1) / kTcBM);
  const int kbk = ep.bf16 ? 2 * kTcBK : kTcBK;
  const int total_kb = (K + kbk - 1) / kbk;
  // at least 3 k-blocks per split (4 without the cluster exchange, whose reduction costs ~1.5 us less: measured)
  const TcKnobs& kn = tc_knobs();
  const int min_kb = kn.min_kb > 0 ? kn.min_kb : (kn.cluster_splitk ? 3 : 4);
  int want = total_kb / (min_kb > 0 ? min_kb : 4);
  const int cap = sm_count / tiles;
  if (want > cap) want = cap;
  if (want > 8) want = 8;
  if (kn.max_splits > 0 && want > kn.max_splits) want = kn.max_splits;
  // Prefer a split count whose clusters are all co-resident (a cluster lives inside one GPC: ~sm_count/8 SMs), as long
  // as that costs at most one split level -- otherwise the launch falls back to the workspace exchange.
  if (kn.cluster_splitk && want >= 2) {
    const int gpc = sm_count / 8;
    int w = want;
    while (w > 1 && tiles > 8 * (gpc / w)) --w;
    if (w >= 2 && w + 1 >= want) want = w;
  }
  if (want < 2) return 1;
  const int per = (total_kb + want - 1) / want;
  const int ns = (total_kb + per - 1) / per;                 // every split gets at least one k-block
  if ((size_t)tiles * ns * kTcBM * bn > ws_floats || tiles > 4096) return 1;
  return ns;
}

template <int kEpi>
inline int tc05_launch(int bn, bool split, const CUtensorMap& a, const CUtensorMap& b, int M, int N, int K, const Tc05Epi& ep,
                       cudaStream_t s, int nsplit = 1) {
  if (bn == 16) return split ? tc05_launch_t<16, true, kEpi>(a, b, M, N, K, ep, s, nsplit) : tc05_launch_t<16, false, kEpi>(a, b, M, N, K, ep, s, nsplit);
  if (bn == 32) return split ? tc05_launch_t<32, true, kEpi>(a, b, M, N, K, ep, s, nsplit) : tc05_launch_t<32, false, kEpi>(a, b, M, N, K, ep, s, nsplit);
  if (bn == 64) return split ? tc05_launch_t<64, true, kEpi>(a, b, M, N, K, ep, s, nsplit) : tc05_launch_t<64, false, kEpi>(a, b, M, N, K, ep, s, nsplit);
  return set_error(MDT_ERR_INVALID, "unsupported BLOCK_N %d", bn);
}

// What the engine hands over: its flat buffers and activation workspaces (row-major, owned by the engine).
struct Tc05Buffers {
  float *params, *grads, *exp_avg, *exp_avg_sq;
  Layout L;
  int D, H, Z, Bmax;
  float *x_stage, *h1, *ml, *eps, *sd, *z, *h3, *recon, *da4, *da3, *dz, *dml_kl, *dml, *da1;
  float* loss_dev;
  StepState* st;
  int sm_count, max_threads_sm;
  bool split3;
  bool bf16;          // MDT_GEMM_BF16: bf16 operand mirrors, kind::f16 MMAs (fp32 accumulate, fp32 master weights and Adam)
  void* note_ctx;
  int (*note)(void* ctx, const char* label);   // engine book-keeping after every kernel launch
  const bool* serial;                          // when *serial is true (per-kernel profiling) nothing is forked
  int (*bucket_ready)(void* ctx, int bucket, cudaStream_t on);   // gradient bucket complete on stream `on` (or NULL)
};

struct Tc05ActMaps {   // tensor maps whose extents depend on the batch size B
  CUtensorMap x, h1, z, h3, da4, da3, dml;             // A operands [B, *]
  CUtensorMap da4T, da3T, dmlT, da1T;                  // A operands [*, B]
  CUtensorMap h3T, zT, h1T, xT;                        // B operands [*, B]
  CUtensorMap h1B, da3B;                               // B operands [B, *] of the transposed (skinny-output) GEMMs
  CUtensorMap h3Ta, h1Ta, xTa;                         // B operands [*, B] of the Adam-fused weight-gradient GEMMs (bn_dwa rows)
  CUtensorMap x_1, xT_1, xTa_1;                        // the same three x maps over the second input set (Tc05Plan::xpar == 1)
};

struct Tc05Plan {
  Tc05Buffers b;
  int ldB;            // leading dimension of every transposed activation (floats)
  int bn_h, bn_d;     // BLOCK_N used for N = H and N = D
  int bn_skinny;      // BLOCK_N (samples per tile) of the transposed fc21|fc22 and dz GEMMs
  int bn_dw;          // BLOCK_N of the weight-gradient GEMMs (wide tiles: fewer CTAs next to the dX GEMM they overlap)
  int bn_fc4;         // BLOCK_N of the fc4 forward GEMM (narrow tiles spread the BCE epilogue over more SMs)
  bool simt_latent;   // the latent-side layers (fc21|fc22 + reparam + fc3, and their backward) run as fused CUDA-core kernels
  int bn_dwa;         // BLOCK_N of the Adam-fused weight-gradient GEMMs (narrow: the update's traffic is bound per SM)
  char* slab = nullptr;
  float *xT, *h1T, *zT, *h3T, *da4T, *da3T, *dmlT, *da1T, *W4T, *W3T, *W2xT, *kl_partials, *bce_partials, *ws, *ws_dw;
  // MDT_GEMM_BF16 operand mirrors (row-major activations, weights and weight shadows); in that mode the transposed
  // activations above hold bf16 as well
  __nv_bfloat16 *x16 = nullptr, *h1_16 = nullptr, *z16 = nullptr, *h3_16 = nullptr, *da4_16 = nullptr, *da3_16 = nullptr, *dml16 = nullptr;
  __nv_bfloat16 *W1_16 = nullptr, *W2x_16 = nullptr, *W3_16 = nullptr, *W4_16 = nullptr, *W4T_16 = nullptr, *W3T_16 = nullptr, *W2xT_16 = nullptr;
  int ldz16 = 0;      // leading dimension of z16 / W3_16 (Z rounded up to 8: 16-byte rows)
  unsigned int* ticket;
  unsigned int* ws_counters;
  static constexpr size_t kWsFloats = 2u << 20;   // 8 MB split-K workspace
  CUtensorMap mW1, mW2x, mW3, mW4, mW4T, mW3T, mW2xT;  // B operands (weights); mW2x / mW3T are A operands (transposed problems)
  std::map<int, Tc05ActMaps> maps;
  // weight-gradient GEMMs run on a side stream, off the dX critical path (fork/join by events; capturable)
  // Where the next forward's x lives (NULL: already in the stage buffer).  The prologue kernel reads it from there and
  // fills the stage buffer itself; last_prep / prep_launched let the engine find and re-point that kernel in a graph.
  const float* x_src = nullptr;
  // mdt_train_steps: set while its graphs are captured (see PrepArgs)
  const float* src_pool = nullptr; const uint8_t* src_images = nullptr; const long long* src_indices = nullptr;
  long long src_n_batches = 0, src_n_images = 0;
  // Non-NULL while the engine enqueues a peer-memory exchange step: the weight-gradient GEMMs of tc05_backward write
  // the gradient there (the buffer the group's peers have mapped) instead of into the caller's gradient buffer.
  float* grads_override = nullptr;
  // How many trials share this GPU right now (mdt_train_step_many): split counts and CTA budgets are sized for
  // sm_count / sm_share SMs, so that the kernels of different trials run side by side instead of queueing for SMs
  // (measured, 8 trials: 3.5 -> 5.1 M samples/s).
  int sm_share = 1;
  // Two-set pipelining of the step (CUDA-core latent path).  The step's prologue (x -> stage copy, x^T, bf16 copy) does
  // not depend on the previous step, and fc3's update only conflicts with latent_bwd through W3^T.  So both exist twice:
  //  * input sets: while step j computes on set `xpar`, a side stream fills the other set with batch j+1 (mdt_train_steps
  //    knows it: the batch index lives on the device), so the next step starts with its first GEMM, not with a prologue
  //    kernel behind four full dependencies;
  //  * W3^T sets: latent_bwd reads W3T_set[w3par] while fc3's weight-gradient + Adam kernel, started as early as fc4's,
  //    writes the updated transpose into the other one -- it leaves the step's tail.
  float* x_stage_set[2] = {nullptr, nullptr};
  float* xT_set[2] = {nullptr, nullptr};
  __nv_bfloat16* x16_set[2] = {nullptr, nullptr};
  float* W3T_set[2] = {nullptr, nullptr};
  int xpar = 0, w3par = 0;
  bool pipelined = false;        // set by mdt_train_steps around its enqueues: no prologue in the step, prefetch the next batch
  bool no_prefetch = false;      // pipelined step whose successor's batch is fetched by the caller (host-buffer steps)
  bool prefetch_pending = false; // a prefetch was forked in this step and has not been joined yet
  cudaStream_t side_pf = nullptr;
  cudaEvent_t ev_pf_fork = nullptr, ev_pf_join = nullptr;
  // Steps whose optimizer is a separate kernel (DDP groups, SGD): the transposed weight shadows it has to refresh are
  // first read by the NEXT step's fc4 input-gradient GEMM, four kernels into that step.  Inside a multi-step graph the
  // refresh therefore runs on a side stream and is joined right before that GEMM instead of sitting between the
  // optimizer and the next step's first GEMM (fp32-operand modes only: bf16 also refreshes the operand copies the very
  // next GEMM reads).  The last step of a capture refreshes in line.
  cudaStream_t side_wt = nullptr;
  cudaEvent_t ev_wt_fork = nullptr, ev_wt_join = nullptr;
  bool defer_shadows = false, shadows_pending = false;
  TcKnobs knobs;       // the MDT_* diagnostic switches as they stood when the plan was created
  int preload = 1;     // issue weight-operand TMA loads before the dependency wait (MDT_NO_PRELOAD=1 turns it off)
  PrepArgs last_prep{};
  bool prep_launched = false;
  cudaStream_t side = nullptr, side2 = nullptr, side3 = nullptr, side4 = nullptr;
  cudaEvent_t ev_fork[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_join = nullptr, ev_join2 = nullptr, ev_join3 = nullptr, ev_join4 = nullptr;
};

#define TC_NOTE(p, label)                                                       \
  do {                                                                          \
    if (tc_recorder() == nullptr) MDT_TRY((p)->b.note((p)->b.note_ctx, label)); \
  } while (0)

inline int tc05_weight_shadows(Tc05Plan* p, cudaStream_t s) {
  const Tc05Buffers& b = p->b;
  TrArgs a{};
  int blocks = 0;
  const float* P = b.params;
  if (b.bf16) {
    // bf16 mode: W4^T is only a GEMM operand (bf16 shadow); W3^T / W2x^T are also read in fp32 by the CUDA-core latent
    // kernels; every weight gets a row-major bf16 copy
    tr_add(a, blocks, P + b.L.off[W4], nullptr, b.D, b.H, b.H, b.D, false, p->W4T_16, p->W4_16, b.H);
    tr_add(a, blocks, P + b.L.off[W3], p->W3T_set[p->w3par], b.H, b.Z, b.Z, b.H, false, p->W3T_16, p->W3_16, p->ldz16);
    tr_add(a, blocks, P + b.L.off[W2X], p->W2xT, 2 * b.Z, b.H, b.H, 2 * b.Z, false, p->W2xT_16, p->W2x_16, b.H);
    tr_add(a, blocks, P + b.L.off[W1], nullptr, b.H, b.D, b.D, b.H, false, nullptr, p->W1_16, b.D);
  } else {
    tr_add(a, blocks, P + b.L.off[W4], p->W4T, b.D, b.H, b.H, b.D);
    tr_add(a, blocks, P + b.L.off[W3], p->W3T_set[p->w3par], b.H, b.Z, b.Z, b.H);
    tr_add(a, blocks, P + b.L.off[W2X], p->W2xT, 2 * b.Z, b.H, b.H, 2 * b.Z);
  }
  MDT_CUDA(launch_pdl(transpose_kernel, dim3(blocks), dim3(256), 0, s, a));
  TC_NOTE(p, "weights_T");
  return MDT_OK;
}

inline int tc05_params_changed(Tc05Plan* p, cudaStream_t s) {
  if (p->defer_shadows && !p->b.bf16) {
    MDT_CUDA(tc_event_record(p->ev_wt_fork, s));
    MDT_CUDA(tc_stream_wait(p->side_wt, p->ev_wt_fork));
    MDT_TRY(tc05_weight_shadows(p, p->side_wt));
    MDT_CUDA(tc_event_record(p->ev_wt_join, p->side_wt));
    p->shadows_pending = true;
    return MDT_OK;
  }
  return tc05_weight_shadows(p, s);
}
// Before the first kernel of a step that reads a transposed weight shadow.
inline int tc05_join_shadows(Tc05Plan* p, cudaStream_t s) {
  if (!p->shadows_pending) return MDT_OK;
  MDT_CUDA(tc_stream_wait(s, p->ev_wt_join));
  p->shadows_pending = false;
  return MDT_OK;
}

inline int tc05_sms(const Tc05Plan* p) {
  const int n = p->b.sm_count / (p->sm_share > 0 ? p->sm_share : 1);
  return n < 8 ? 8 : n;
}

// Where the fused training forward left the loss partials of a batch of B samples (BCE per fc4 tile, KLD per
// reparameterisation tile); the Adam kernel adds them.
inline void tc05_loss_partials(const Tc05Plan* p, int B, const float** a, int* na, const float** b, int* nb) {
  const int D = p->b.D;
  *a = p->bce_partials;
  // one slot per (fc4 tile, K-split): same split count as tc05_forward's launch of fc4
  int ns = 1;
  TcKnobScope knob_scope(&p->knobs);
  if (p->knobs.splitk) {
    Tc05Epi e{};
    e.ws = p->ws; e.ws_counters = p->ws_counters; e.bf16 = p->b.bf16 ? 1 : 0;
    ns = tc05_pick_splits(B, D, p->b.H, p->bn_fc4, tc05_sms(p), e, Tc05Plan::kWsFloats);
  }
  *na = ((D + p->bn_fc4 - 1) / p->bn_fc4) * ((B + kTcBM - 1) / kTcBM) * ns;
  *b = p->kl_partials;
  *nb = p->simt_latent ? latent_grid(B, latent_pick_sc(B, tc05_sms(p), false, p->knobs.latent_sc)) : (B + p->bn_skinny - 1) / p->bn_skinny;
}

inline void tc05_destroy(Tc05Plan* p) {
  if (!p) return;
  if (p->side) cudaStreamDestroy(p->side);
  if (p->side_pf) cudaStreamDestroy(p->side_pf);
  if (p->side_wt) cudaStreamDestroy(p->side_wt);
  if (p->ev_wt_fork) cudaEventDestroy(p->ev_wt_fork);
  if (p->ev_wt_join) cudaEventDestroy(p->ev_wt_join);
  if (p->ev_pf_fork) cudaEventDestroy(p->ev_pf_fork);
  if (p->ev_pf_join) cudaEventDestroy(p->ev_pf_join);
  for (cudaStream_t st : {p->side2, p->side3, p->side4}) if (st) cudaStreamDestroy(st);
  for (cudaEvent_t ev : {p->ev_join2, p->ev_join3, p->ev_join4}) if (ev) cudaEventDestroy(ev);
  for (auto& ev : p->ev_fork) if (ev) cudaEventDestroy(ev);
  if (p->ev_join) cudaEventDestroy(p->ev_join);
  if (p->slab) cudaFree(p->slab);
  delete p;
}

inline int tc05_create(const Tc05Buffers* bufs, cudaStream_t s, Tc05Plan** out) {
  const Tc05Buffers& b = *bufs;
  MDT_CHECK(b.Z <= 32, "tcgen05 path supports latent <= 32 (got %d)", b.Z);
  (void)tc05_dbg_buffer();          // allocate the optional phase-timing buffer outside any stream capture
  Tc05Plan* p = new Tc05Plan();
  p->b = b;
  p->ldB = (int)round_up(b.Bmax, b.bf16 ? 8 : 4);      // 16-byte rows of the transposed activations
  p->ldz16 = (int)round_up(b.Z, 8);
  p->bn_h = b.H >= 2048 ? 64 : 32;
  p->bn_d = b.Bmax > 512 ? 64 : 32;
  p->bn_fc4 = b.Bmax > 512 ? 64 : 32;   // 32: 25 tiles x 4 splits still fit as co-resident clusters (16: 49 x 3 do not)
  p->bn_dw = 64;
  p->bn_dwa = 32;
  // The CUDA-core latent kernels win while the step is latency-bound (measured: 90 vs 92 us at H=1024 B=128, 58 vs 70 us
  // at H=400 B=128); every CTA re-reads the 2Z x H weight rows, so at larger H x B the tensor-core GEMMs take over
  // (147 vs 143 us at H=400 B=1024, 274 vs 255 us at H=1024 B=1024, 248 vs 219 us at H=4096 B=128).
  // MDT_TC_LATENT=1 / 0 forces either.
  p->simt_latent = (long long)b.H * b.Bmax <= 256 * 1024;
  if (const char* v = getenv("MDT_TC_LATENT")) p->simt_latent = v[0] == '0';
  if (const char* v = getenv("MDT_BN_DWA")) p->bn_dwa = atoi(v);
  if (const char* v = getenv("MDT_BN_FC4")) p->bn_fc4 = atoi(v);          // tuning knobs (16 / 32 / 64)
  if (const char* v = getenv("MDT_BN_H")) p->bn_h = atoi(v);
  p->bn_skinny = b.Bmax > 1024 ? 64 : (b.Bmax > 256 ? 32 : 16);
  if (const char* v = getenv("MDT_BN_SKINNY")) { const int c = atoi(v); if (c == 16 || c == 32 || c == 64) p->bn_skinny = c; }
  p->preload = getenv("MDT_NO_PRELOAD") == nullptr ? 1 : 0;
  p->knobs = TcKnobs::from_env();
  const size_t ldB = p->ldB, D = b.D, H = b.H, Z = b.Z;
  struct Item { float** ptr; size_t n; };
  Item items[] = {{&p->xT, D * ldB}, {&p->h1T, H * ldB}, {&p->zT, Z * ldB}, {&p->h3T, H * ldB}, {&p->da4T, D * ldB},
                  {&p->da3T, H * ldB}, {&p->dmlT, 2 * Z * ldB}, {&p->da1T, H * ldB}, {&p->W4T, H * D}, {&p->W3T, Z * H},
                  {&p->W2xT, H * 2 * Z}, {&p->x_stage_set[1], (size_t)b.Bmax * D}, {&p->xT_set[1], D * ldB}, {&p->W3T_set[1], Z * H},
                  {&p->kl_partials, 1024}, {&p->bce_partials, 65536}, {&p->ws, Tc05Plan::kWsFloats}, {&p->ws_dw, 2 * Tc05Plan::kWsFloats}, {reinterpret_cast<float**>(&p->ws_counters), 8192}};
  // bf16 operand mirrors: sizes in bf16 elements
  struct Item16 { __nv_bfloat16** ptr; size_t n; };
  const size_t Bm = b.Bmax, lz = p->ldz16;
  Item16 items16[] = {{&p->x16, Bm * D}, {&p->h1_16, Bm * H}, {&p->z16, Bm * lz}, {&p->h3_16, Bm * H}, {&p->da4_16, Bm * D},
                      {&p->da3_16, Bm * H}, {&p->dml16, Bm * 2 * Z}, {&p->W1_16, H * D}, {&p->W2x_16, 2 * Z * H}, {&p->W3_16, H * lz},
                      {&p->W4_16, D * H}, {&p->W4T_16, H * D}, {&p->W3T_16, Z * H}, {&p->W2xT_16, H * 2 * Z}, {&p->x16_set[1], Bm * D}};
  size_t bytes = 256;
  for (auto& it : items) bytes += (it.n * sizeof(float) + 255) / 256 * 256;
  if (b.bf16)
    for (auto& it : items16) bytes += (it.n * sizeof(__nv_bfloat16) + 255) / 256 * 256;
  cudaError_t ce = cudaMalloc(&p->slab, bytes);
  if (ce != cudaSuccess) { delete p; return set_error(MDT_ERR_CUDA, "cudaMalloc(%zu) for the tcgen05 plan: %s", bytes, cudaGetErrorString(ce)); }
  cudaMemsetAsync(p->slab, 0, bytes, s);
  char* q = p->slab;
  p->ticket = reinterpret_cast<unsigned int*>(q);
  q += 256;
  for (auto& it : items) { *it.ptr = reinterpret_cast<float*>(q); q += (it.n * sizeof(float) + 255) / 256 * 256; }
  if (b.bf16)
    for (auto& it : items16) { *it.ptr = reinterpret_cast<__nv_bfloat16*>(q); q += (it.n * sizeof(__nv_bfloat16) + 255) / 256 * 256; }
  p->x_stage_set[0] = b.x_stage; p->xT_set[0] = p->xT; p->W3T_set[0] = p->W3T; p->x16_set[0] = p->x16;
  int r = MDT_OK;
  if (cudaStreamCreateWithFlags(&p->side_wt, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev_wt_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev_wt_join, cudaEventDisableTiming) != cudaSuccess ||
      cudaStreamCreateWithFlags(&p->side_pf, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev_pf_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev_pf_join, cudaEventDisableTiming) != cudaSuccess)
    r = set_error(MDT_ERR_CUDA, "prefetch stream");
  if (cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&p->side2, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&p->side3, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&p->side4, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev_join2, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev_join3, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev_join4, cudaEventDisableTiming) != cudaSuccess)
    r = set_error(MDT_ERR_CUDA, "side streams");
  for (auto& ev : p->ev_fork)
    if (r == MDT_OK && cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) r = set_error(MDT_ERR_CUDA, "event");
  if (r == MDT_OK && cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming) != cudaSuccess) r = set_error(MDT_ERR_CUDA, "event");
  const float* P = b.params;
  const bool bf = b.bf16;
  // weight operands: the fp32 parameters themselves (tf32 kinds), or their bf16 copies
  if (r == MDT_OK) r = tc05_encode(&p->mW1, bf ? (const void*)p->W1_16 : P + b.L.off[W1], b.H, b.D, b.D, p->bn_h, bf);
  if (r == MDT_OK) r = tc05_encode(&p->mW2x, bf ? (const void*)p->W2x_16 : P + b.L.off[W2X], 2 * b.Z, b.H, b.H, kTcBM, bf);
  if (r == MDT_OK) r = tc05_encode(&p->mW3, bf ? (const void*)p->W3_16 : P + b.L.off[W3], b.H, b.Z, bf ? p->ldz16 : b.Z, p->bn_h, bf);
  if (r == MDT_OK) r = tc05_encode(&p->mW4, bf ? (const void*)p->W4_16 : P + b.L.off[W4], b.D, b.H, b.H, p->bn_fc4, bf);
  if (r == MDT_OK) r = tc05_encode(&p->mW4T, bf ? (const void*)p->W4T_16 : p->W4T, b.H, b.D, b.D, p->bn_h, bf);
  if (r == MDT_OK) r = tc05_encode(&p->mW3T, bf ? (const void*)p->W3T_16 : p->W3T, b.Z, b.H, b.H, kTcBM, bf);
  if (r == MDT_OK) r = tc05_encode(&p->mW2xT, bf ? (const void*)p->W2xT_16 : p->W2xT, b.H, 2 * b.Z, 2 * b.Z, p->bn_h, bf);
  if (r == MDT_OK) r = tc05_weight_shadows(p, s);
  if (r != MDT_OK) { tc05_destroy(p); return r; }
  *out = p;
  return MDT_OK;
}

inline int tc05_act_maps(Tc05Plan* p, int B, Tc05ActMaps** out) {
  auto it = p->maps.find(B);
  if (it == p->maps.end()) {
    const Tc05Buffers& b = p->b;
    Tc05ActMaps m;
    const int ldB = p->ldB;
    const bool bf = b.bf16;
    auto act = [&](const float* f32, const __nv_bfloat16* b16) -> const void* { return bf ? (const void*)b16 : (const void*)f32; };
    MDT_TRY(tc05_encode(&m.x, act(b.x_stage, p->x16), B, b.D, b.D, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.h1, act(b.h1, p->h1_16), B, b.H, b.H, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.z, act(b.z, p->z16), B, b.Z, bf ? p->ldz16 : b.Z, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.h3, act(b.h3, p->h3_16), B, b.H, b.H, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.da4, act(b.da4, p->da4_16), B, b.D, b.D, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.da3, act(b.da3, p->da3_16), B, b.H, b.H, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.dml, act(b.dml, p->dml16), B, 2 * b.Z, 2 * b.Z, kTcBM, bf));
    // transposed activations: one buffer each, holding fp32 or bf16 as the mode says
    MDT_TRY(tc05_encode(&m.da4T, p->da4T, b.D, B, ldB, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.da3T, p->da3T, b.H, B, ldB, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.dmlT, p->dmlT, 2 * b.Z, B, ldB, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.da1T, p->da1T, b.H, B, ldB, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.h3T, p->h3T, b.H, B, ldB, p->bn_dw, bf));
    MDT_TRY(tc05_encode(&m.zT, p->zT, b.Z, B, ldB, 32, bf));
    MDT_TRY(tc05_encode(&m.h1T, p->h1T, b.H, B, ldB, p->bn_dw, bf));
    MDT_TRY(tc05_encode(&m.xT, p->xT, b.D, B, ldB, p->bn_dw, bf));
    MDT_TRY(tc05_encode(&m.h3Ta, p->h3T, b.H, B, ldB, p->bn_dwa, bf));
    MDT_TRY(tc05_encode(&m.h1Ta, p->h1T, b.H, B, ldB, p->bn_dwa, bf));
    MDT_TRY(tc05_encode(&m.xTa, p->xT, b.D, B, ldB, p->bn_dwa, bf));
    MDT_TRY(tc05_encode(&m.x_1, act(p->x_stage_set[1], p->x16_set[1]), B, b.D, b.D, kTcBM, bf));
    MDT_TRY(tc05_encode(&m.xT_1, p->xT_set[1], b.D, B, ldB, p->bn_dw, bf));
    MDT_TRY(tc05_encode(&m.xTa_1, p->xT_set[1], b.D, B, ldB, p->bn_dwa, bf));
    MDT_TRY(tc05_encode(&m.h1B, act(b.h1, p->h1_16), B, b.H, b.H, p->bn_skinny, bf));
    MDT_TRY(tc05_encode(&m.da3B, act(b.da3, p->da3_16), B, b.H, b.H, p->bn_skinny, bf));
    it = p->maps.emplace(B, m).first;
  }
  *out = &it->second;
  return MDT_OK;
}

// Split-K factor for a main-stream GEMM of the step (they run one after another, so they share the workspace).
inline int tc05_splits(Tc05Plan* p, Tc05Epi& e, int M, int N, int K, int bn) {
  if (!p->knobs.splitk) return 1;
  e.ws = p->ws;
  e.ws_counters = p->ws_counters;
  return tc05_pick_splits(M, N, K, bn, tc05_sms(p), e, Tc05Plan::kWsFloats);
}

// Same for the weight-gradient GEMMs, which run on the side stream concurrently with the main-stream GEMMs and
// therefore own a separate workspace and ticket range.
inline int tc05_splits_dw(Tc05Plan* p, Tc05Epi& e, int M, int N, int K, int bn, int slot = 0) {
  if (!p->knobs.splitk) return 1;
  // up to four weight-gradient GEMMs can be in flight at once (one side stream each): every slot owns 4 MB of workspace
  e.ws = p->ws_dw + (size_t)slot * (Tc05Plan::kWsFloats / 2);
  e.ws_counters = p->ws_counters + 2048 + slot * 1024;
  const int tiles = ((N + bn - 1) / bn) * ((M + kTcBM - 1) / kTcBM);
  if (tiles > 1024) return 1;
  // the side stream shares the GPU with the dX GEMM it overlaps: aim for about half the SMs
  return tc05_pick_splits(M, N, K, bn, tc05_sms(p) / 2, e, Tc05Plan::kWsFloats / 2);
}

// The step's prologue kernel into input set `set`: x -> stage copy (+ bf16 copy) and x^T; with `with_eps` also the
// step's eps from Philox (tensor-core latent path: the CUDA-core latent_fwd draws its own).  cursor_add = 1 fetches the
// NEXT step's batch of a mdt_train_steps source.
inline int tc05_prep(Tc05Plan* p, cudaStream_t s, const float* x, int B, int set, bool with_eps, int cursor_add, const char* label,
                     bool programmatic = true) {
  const Tc05Buffers& b = p->b;
  const int D = b.D, Z = b.Z;
  PrepArgs pa{};
  int blocks = 0;
  const float* src = (p->x_src != nullptr) ? p->x_src : x;
  float* const stage = p->x_stage_set[set];
  tr_add(pa.tr, blocks, src, p->xT_set[set], B, D, D, p->ldB, b.bf16, nullptr, b.bf16 ? p->x16_set[set] : nullptr, D);
  pa.tr_blocks = blocks;
  pa.x_copy = (src != stage) ? stage : nullptr;
  if (p->src_pool != nullptr || p->src_images != nullptr) {
    pa.pool = p->src_pool; pa.images = p->src_images; pa.indices = p->src_indices;
    pa.n_batches = p->src_n_batches; pa.n_images = p->src_n_images; pa.st = b.st;
    pa.x_copy = stage;
    pa.cursor_add = cursor_add;
  }
  if (with_eps) {
    pa.eps_out = b.eps; pa.n_eps = (long long)B * Z; pa.st = b.st;
    torch_randn_policy(pa.n_eps, b.sm_count, b.max_threads_sm, &pa.S, &pa.offset_inc);
    blocks += (int)((pa.n_eps + 255) / 256);
  }
  if (TcRecorder* r = tc_recorder()) {
    TcOp op; op.kind = TcOp::PREP; op.stream = s; op.pa = pa; op.blocks = blocks; op.programmatic = programmatic ? 1 : 0;
    r->ops.push_back(op);
    return MDT_OK;
  }
  if (programmatic) {
    MDT_CUDA(launch_pdl(prep_kernel, dim3(blocks), dim3(256), 0, s, pa));
  } else {
    prep_kernel<<<dim3(blocks), dim3(256), 0, s>>>(pa);
    MDT_CUDA(cudaGetLastError());
  }
  p->last_prep = pa;
  p->prep_launched = true;
  TC_NOTE(p, label);
  return MDT_OK;
}

// Forward.  fuse_loss: fc4's epilogue also produces the loss, d(logits) and its transpose (training step);
// otherwise fc4 stops at recon (mdt_forward / eval / decode) and the stand-alone loss kernel may follow.
inline int tc05_forward(Tc05Plan* p, cudaStream_t s, const float* x, int B, const float* eps_in, bool decode_only,
                        bool fuse_loss) {
  TcKnobScope knob_scope(&p->knobs);
  const Tc05Buffers& b = p->b;
  const bool sp = b.split3;
  const int D = b.D, H = b.H, Z = b.Z;
  const float* P = b.params;
  Tc05ActMaps* m = nullptr;
  MDT_TRY(tc05_act_maps(p, B, &m));
  const bool pipe = p->pipelined && p->simt_latent && !decode_only && fuse_loss;
  if (!pipe) p->xpar = 0;                        // everything but the pipelined steps of mdt_train_steps works on input set 0
  const int xs = p->xpar;
  const float* const xst = p->x_stage_set[xs];   // this step's x (BCE target, fc1's A operand through its tensor map)
  bool gen_eps_latent = false;                   // latent_fwd draws eps itself
  if (!decode_only) {
    MDT_CHECK(x == b.x_stage, "tcgen05 path reads x from the engine's stage buffer");
    bool prep_eps = false;
    if (eps_in == nullptr) {
      if (p->simt_latent) gen_eps_latent = true; // the CUDA-core latent_fwd draws eps itself, ahead of its dependency wait
      else { prep_eps = true; eps_in = b.eps; }  // tensor-core latent path: drawn by the prologue kernel
    }
    // pipelined steps have no prologue: set `xs` was filled while the previous step ran (or by mdt_train_steps' first call)
    if (!pipe) MDT_TRY(tc05_prep(p, s, x, B, xs, prep_eps, 0, "prep_xT_eps"));
    Tc05Epi e1{};
    e1.C = b.h1; e1.ldc = H; e1.CT = p->h1T; e1.ldct = p->ldB; e1.bias = P + b.L.off[B1]; e1.act = MDT_ACT_RELU;
    // W1 was last written by the previous step's update: its loads may go out before the dependency wait -- but only
    // when a prologue kernel sits between that update and this GEMM (a pipelined step follows the update directly)
    e1.pre_b = pipe ? 0 : p->preload;
    e1.bf16 = b.bf16; e1.Cb = b.bf16 ? p->h1_16 : nullptr; e1.ldcb = H;
    MDT_TRY(tc05_launch<EPI_BIAS_ACT>(p->bn_h, sp, xs ? m->x_1 : m->x, p->mW1, B, H, D, e1, s, tc05_splits(p, e1, B, H, D, p->bn_h)));
    TC_NOTE(p, "fc1_fwd");
    if (pipe && !p->no_prefetch) {
      // the next step's batch goes into the other input set on a side stream, behind this step's first GEMM (whose
      // start implies the previous step, the last reader of that set, is complete); joined where the step joins its
      // weight-gradient streams
      MDT_CUDA(tc_event_record(p->ev_pf_fork, s));
      MDT_CUDA(tc_stream_wait(p->side_pf, p->ev_pf_fork));
      // NOT a programmatic launch: its blocks would be resident from this GEMM's start, parked in the dependency wait
      // on the SMs the next kernel of the step needs for its own early start (measured: latent_fwd entered 3 us late)
      MDT_TRY(tc05_prep(p, p->side_pf, x, B, xs ^ 1, false, 1, "prefetch_xT", /*programmatic=*/false));
      MDT_CUDA(tc_event_record(p->ev_pf_join, p->side_pf));
      p->prefetch_pending = true;
    }
    if (p->simt_latent) {
      LatentFwdArgs la{};
      la.h1 = b.h1; la.H = H; la.W2x = P + b.L.off[W2X]; la.b2x = P + b.L.off[B2X]; la.W3 = P + b.L.off[W3]; la.b3 = P + b.L.off[B3];
      la.eps_in = gen_eps_latent ? nullptr : eps_in; la.ml = b.ml;
      la.st = b.st;
      torch_randn_policy((long long)B * Z, b.sm_count, b.max_threads_sm, &la.philox_S, &la.philox_inc); la.z = b.z; la.zT = p->zT; la.ldT = p->ldB; la.eps_out = b.eps; la.sd_out = b.sd;
      la.dml_kl = b.dml_kl; la.kl_partials = p->kl_partials; la.h3 = b.h3; la.h3T = p->h3T; la.B = B; la.Z = Z;
      la.bf16 = b.bf16; la.h3b = b.bf16 ? p->h3_16 : nullptr;
      la.dbg = tc05_dbg_buffer();
      if (la.dbg) la.dbg += 16 * (tc05_dbg_launch_index()++ % 32);
      if (TcRecorder* r = tc_recorder()) {
        TcOp op; op.kind = TcOp::LAT_FWD; op.stream = s; op.sc = latent_pick_sc(B, tc05_sms(p), false, p->knobs.latent_sc);
        la.dbg = nullptr; la.nchunks = (B + op.sc - 1) / op.sc; la.first_live_pass = 1; la.first_pass = 1;
        op.lf = la;
        r->ops.push_back(op);
      } else {
        MDT_CUDA(launch_latent_fwd(la, latent_pick_sc(B, tc05_sms(p), false, p->knobs.latent_sc), s));
      }
      TC_NOTE(p, "latent_fwd");
    } else {
    Tc05Epi e2{};
    e2.C = b.ml; e2.ldc = 2 * Z; e2.bias = P + b.L.off[B2X]; e2.eps_in = eps_in;
    e2.z = b.z; e2.zT = p->zT; e2.ldzT = p->ldB; e2.eps_out = b.eps; e2.sd_out = b.sd; e2.dml_kl = b.dml_kl; e2.Z = Z;
    e2.kl_partials = p->kl_partials;
    e2.pre_a = p->preload;                       // W2x
    e2.bf16 = b.bf16; e2.zb = b.bf16 ? p->z16 : nullptr; e2.ldzb = p->ldz16;
    MDT_TRY(tc05_launch<EPI_REPARAM>(p->bn_skinny, sp, p->mW2x, m->h1B, 2 * Z, B, H, e2, s,
                                      tc05_splits(p, e2, 2 * Z, B, H, p->bn_skinny)));
    TC_NOTE(p, "fc2x_reparam_fwd");
    }
  }
  if (decode_only && b.bf16) {                    // z came from the caller (fp32): make the bf16 operand
    MDT_CUDA(launch_pdl(rows_to_bf16_kernel, dim3((B * Z + 255) / 256), dim3(256), 0, s, (const float*)b.z, Z, p->z16, p->ldz16, B, Z));
    TC_NOTE(p, "z_bf16");
  }
  if (decode_only || !p->simt_latent) {
    Tc05Epi e3{};
    e3.C = b.h3; e3.ldc = H; e3.CT = p->h3T; e3.ldct = p->ldB; e3.bias = P + b.L.off[B3]; e3.act = MDT_ACT_RELU;
    e3.pre_b = p->preload;                       // W3
    e3.bf16 = b.bf16; e3.Cb = b.bf16 ? p->h3_16 : nullptr; e3.ldcb = H;
    MDT_TRY(tc05_launch<EPI_BIAS_ACT>(p->bn_h, sp, m->z, p->mW3, B, H, Z, e3, s));
    TC_NOTE(p, "fc3_fwd");
  }
  if (fuse_loss && !decode_only) {
    Tc05Epi e4{};
    e4.C = b.recon; e4.ldc = D; e4.bias = P + b.L.off[B4]; e4.x = xst; e4.ldx = D; e4.da4 = b.da4; e4.da4T = p->da4T;
    e4.ld_da4T = p->ldB; e4.bce_partials = p->bce_partials; e4.ticket = p->ticket; e4.loss_out = b.loss_dev;
    e4.kl_partials = p->kl_partials; e4.n_kl_partials = (B + p->bn_skinny - 1) / p->bn_skinny;
    e4.pre_b = p->preload;                       // W4
    e4.bf16 = b.bf16; e4.da4b = b.bf16 ? p->da4_16 : nullptr;
    MDT_TRY(tc05_launch<EPI_SIGMOID_BCE>(p->bn_fc4, sp, m->h3, p->mW4, B, D, H, e4, s, tc05_splits(p, e4, B, D, H, p->bn_fc4)));
    TC_NOTE(p, "fc4_bce_fwd");
  } else {
    Tc05Epi e4{};
    e4.C = b.recon; e4.ldc = D; e4.bias = P + b.L.off[B4]; e4.act = MDT_ACT_SIGMOID;
    e4.pre_b = p->preload; e4.bf16 = b.bf16;
    MDT_TRY(tc05_launch<EPI_BIAS_ACT>(p->bn_fc4, sp, m->h3, p->mW4, B, D, H, e4, s, tc05_splits(p, e4, B, D, H, p->bn_fc4)));
    TC_NOTE(p, "fc4_fwd");
  }
  return MDT_OK;
}

// da4 (row-major) was produced outside the fused forward (stand-alone loss kernel / compat backward): make da4^T.
inline int tc05_after_external_da4(Tc05Plan* p, cudaStream_t s, int B) {
  TrArgs a{};
  int blocks = 0;
  tr_add(a, blocks, p->b.da4, p->da4T, B, p->b.D, p->b.D, p->ldB, p->b.bf16, nullptr, p->b.bf16 ? p->da4_16 : nullptr, p->b.D);
  MDT_CUDA(launch_pdl(transpose_kernel, dim3(blocks), dim3(256), 0, s, a));
  TC_NOTE(p, "da4_T");
  return MDT_OK;
}

// dz -> reparameterisation backward -> fc21|fc22 input gradient as one CUDA-core kernel (latent.cuh)
inline int tc05_latent_bwd(Tc05Plan* p, cudaStream_t s, int B) {
  const Tc05Buffers& b = p->b;
  LatentBwdArgs la{};
  la.da3 = b.da3; la.H = b.H; la.W3T = p->W3T_set[p->w3par]; la.W2xT = p->W2xT; la.h1 = b.h1;
  la.eps = b.eps; la.sd = b.sd; la.dml_kl = b.dml_kl; la.dml = b.dml; la.dmlT = p->dmlT; la.ldT = p->ldB; la.da1T = p->da1T;
  la.B = B; la.Z = b.Z; la.bf16 = b.bf16;
  la.dbg = tc05_dbg_buffer();
  if (la.dbg) la.dbg += 16 * (tc05_dbg_launch_index()++ % 32);
  if (TcRecorder* r = tc_recorder()) {
    TcOp op; op.kind = TcOp::LAT_BWD; op.stream = s; op.sc = latent_pick_sc(B, tc05_sms(p), /*backward=*/true, p->knobs.latent_sc);
    la.dbg = nullptr; la.nchunks = (B + op.sc - 1) / op.sc; la.first_live_pass = 1; la.first_pass = 1;
    op.lb = la;
    r->ops.push_back(op);
  } else {
    MDT_CUDA(launch_latent_bwd(la, latent_pick_sc(B, tc05_sms(p), /*backward=*/true, p->knobs.latent_sc), s));
  }
  TC_NOTE(p, "latent_bwd");
  return MDT_OK;
}

// End of a pipelined step: the next batch (fetched on the side stream since this step's first GEMM) is in place; the
// step after this one works on the other input set.
inline int tc05_join_prefetch(Tc05Plan* p, cudaStream_t s) {
  if (!p->prefetch_pending) return MDT_OK;
  MDT_CUDA(tc_stream_wait(s, p->ev_pf_join));
  p->prefetch_pending = false;
  p->xpar ^= 1;
  return MDT_OK;
}

// Backward from da4 / da4^T and dml_kl: fills the flat gradient buffer (weights and biases).
inline int tc05_backward(Tc05Plan* p, cudaStream_t s, const float* x, int B) {
  TcKnobScope knob_scope(&p->knobs);
  const Tc05Buffers& b = p->b;
  const bool sp = b.split3;
  const int D = b.D, H = b.H, Z = b.Z, ldB = p->ldB;
  float* G = p->grads_override ? p->grads_override : b.grads;
  Tc05ActMaps* m = nullptr;
  MDT_TRY(tc05_act_maps(p, B, &m));
  (void)x;   // x^T was written by this step's forward prologue
  const bool fork = !(b.serial && *b.serial);
  // The four weight-gradient GEMMs are independent of each other: each trails on its own side stream (fc1's, the
  // last and largest, stays on the main stream behind the kernel that produces its operand).
  cudaStream_t sw = fork ? p->side : s;
  cudaStream_t sw3 = fork ? p->side2 : s, sw2x = fork ? p->side4 : s;
  int nf = 0;
  auto fork_to = [&](cudaStream_t t) -> int {    // everything enqueued on `s` so far is visible to later work on `t`
    if (!fork) return MDT_OK;
    MDT_CUDA(tc_event_record(p->ev_fork[nf], s));
    MDT_CUDA(tc_stream_wait(t, p->ev_fork[nf]));
    ++nf;
    return MDT_OK;
  };
  Tc05Epi e{};
  // critical path on `s`: fc4 dX -> dz/reparam -> fc2x dX
  MDT_TRY(fork_to(sw));                           // da4^T, h3^T ready
  e = Tc05Epi{};
  e.bf16 = b.bf16; e.C = G + b.L.off[W4]; e.ldc = H; e.a_src = p->da4T; e.lda_src = ldB; e.db = G + b.L.off[B4];
  MDT_TRY(tc05_launch<EPI_GRADW>(p->bn_dw, sp, m->da4T, m->h3T, D, H, B, e, sw, tc05_splits_dw(p, e, D, H, B, p->bn_dw)));
  TC_NOTE(p, "fc4_dW");
  if (b.bucket_ready) MDT_TRY(b.bucket_ready(b.note_ctx, 0, sw));   // fc4.weight + fc4.bias gradients: DDP's first bucket
  MDT_TRY(tc05_join_shadows(p, s));               // W4^T (and, further down, W3^T / W2x^T) refreshed by the previous step
  e = Tc05Epi{};
  e.C = b.da3; e.ldc = H; e.CT = p->da3T; e.ldct = ldB; e.mask = b.h3; e.ldmask = H;
  e.pre_b = p->preload;                          // W4^T: the shadow was last written by the previous step's update
  e.bf16 = b.bf16; e.Cb = b.bf16 ? p->da3_16 : nullptr; e.ldcb = H;
  MDT_TRY(tc05_launch<EPI_RELU_MASK>(p->bn_h, sp, m->da4, p->mW4T, B, H, D, e, s, tc05_splits(p, e, B, H, D, p->bn_h)));
  TC_NOTE(p, "fc4_dX");
  MDT_TRY(fork_to(sw3));                          // da3^T ready (z^T since the forward)
  e = Tc05Epi{};
  e.bf16 = b.bf16; e.C = G + b.L.off[W3]; e.ldc = Z; e.a_src = p->da3T; e.lda_src = ldB; e.db = G + b.L.off[B3];
  MDT_TRY(tc05_launch<EPI_GRADW>(32, sp, m->da3T, m->zT, H, Z, B, e, sw3, tc05_splits_dw(p, e, H, Z, B, 32, 1)));
  TC_NOTE(p, "fc3_dW");
  if (p->simt_latent) {
    MDT_TRY(tc05_latent_bwd(p, s, B));
  } else {
    e = Tc05Epi{};
    e.Z = Z; e.dml_kl = b.dml_kl; e.eps = b.eps; e.sd = b.sd; e.dml = b.dml; e.dmlT = p->dmlT; e.ld_dmlT = ldB;
    e.pre_a = p->preload;                        // W3^T
    e.bf16 = b.bf16; e.dmlb = b.bf16 ? p->dml16 : nullptr;
    MDT_TRY(tc05_launch<EPI_REPARAM_BWD>(p->bn_skinny, sp, p->mW3T, m->da3B, Z, B, H, e, s,
                                          tc05_splits(p, e, Z, B, H, p->bn_skinny)));
    TC_NOTE(p, "fc3_dX_reparam_bwd");
  }
  MDT_TRY(fork_to(sw2x));                         // dml^T ready (latent_bwd: da1^T too)
  e = Tc05Epi{};
  e.bf16 = b.bf16; e.C = G + b.L.off[W2X]; e.ldc = H; e.a_src = p->dmlT; e.lda_src = ldB; e.db = G + b.L.off[B2X];
  MDT_TRY(tc05_launch<EPI_GRADW>(p->bn_dw, sp, m->dmlT, m->h1T, 2 * Z, H, B, e, sw2x, tc05_splits_dw(p, e, 2 * Z, H, B, p->bn_dw, 2)));
  TC_NOTE(p, "fc2x_dW");
  if (!p->simt_latent) {
    e = Tc05Epi{};
    e.CT = p->da1T; e.ldct = ldB; e.mask = b.h1; e.ldmask = H;
    e.pre_b = p->preload;                        // W2x^T
    e.bf16 = b.bf16;
    MDT_TRY(tc05_launch<EPI_RELU_MASK>(p->bn_h, sp, m->dml, p->mW2xT, B, H, 2 * Z, e, s));
    TC_NOTE(p, "fc2x_dX");
  }
  e = Tc05Epi{};
  e.bf16 = b.bf16; e.C = G + b.L.off[W1]; e.ldc = D; e.a_src = p->da1T; e.lda_src = ldB; e.db = G + b.L.off[B1];
  MDT_TRY(tc05_launch<EPI_GRADW>(p->bn_dw, sp, m->da1T, p->xpar ? m->xT_1 : m->xT, H, D, B, e, s, tc05_splits_dw(p, e, H, D, B, p->bn_dw, 3)));
  TC_NOTE(p, "fc1_dW");
  if (fork) {                                     // join: the gradient buffer is complete once the side streams drain
    cudaStream_t st[3] = {sw, sw3, sw2x};
    cudaEvent_t ev[3] = {p->ev_join, p->ev_join2, p->ev_join4};
    for (int i = 0; i < 3; ++i) {
      MDT_CUDA(tc_event_record(ev[i], st[i]));
      MDT_CUDA(tc_stream_wait(s, ev[i]));
    }
  }
  MDT_TRY(tc05_join_prefetch(p, s));
  return MDT_OK;
}

// Backward with the Adam update fused into the weight-gradient GEMMs (single-rank trial groups: nothing to
// all-reduce).  Each dW_l GEMM updates W_l (and its transposed shadow) in its epilogue, so it may only start once the
// dX GEMM that still reads the old W_l^T has finished:
//   main  : fc4 dX --E1--> dz --E2--> fc2x dX --E3-->
//   side  :          E1: fc4 dW
//   side2 :                     E2: fc3 dW
//   side3 :                                            E3: fc1 dW
//   side4 :                                            E3: fc2x dW
// (one stream each: an update is bound by the SM<->L2 bandwidth of the SMs its tiles run on, so the four should not
// queue behind each other; narrow tiles -- bn_dwa -- spread each update over more SMs).  The CTA that takes the step's last ticket advances the optimizer state and sums the loss partials.
inline int tc05_backward_fused_adam(Tc05Plan* p, cudaStream_t s, int B) {
  TcKnobScope knob_scope(&p->knobs);
  const Tc05Buffers& b = p->b;
  const bool sp = b.split3;
  const int D = b.D, H = b.H, Z = b.Z, ldB = p->ldB;
  float* G = b.grads;
  Tc05ActMaps* m = nullptr;
  MDT_TRY(tc05_act_maps(p, B, &m));
  const bool fork = !(b.serial && *b.serial);
  cudaStream_t sa = fork ? p->side : s, sb = fork ? p->side2 : s, sc = fork ? p->side3 : s, sd = fork ? p->side4 : s;
  const int bna = p->bn_dwa;
  int nf = 0;
  auto after_main = [&](cudaStream_t t1, cudaStream_t t2) -> int {   // later work on t1 (and t2) sees `s` so far
    if (!fork) return MDT_OK;
    MDT_CUDA(tc_event_record(p->ev_fork[nf], s));
    MDT_CUDA(tc_stream_wait(t1, p->ev_fork[nf]));
    if (t2) MDT_CUDA(tc_stream_wait(t2, p->ev_fork[nf]));
    ++nf;
    return MDT_OK;
  };
  struct Dw { int seg_w, seg_b, M, N, bn, slot; const CUtensorMap *a, *bm; const float* a_src; float* pT; int ldpT; const char* label; int ns; };
  // MDT_W4_SPLIT=1: fc4's update runs apart from its weight-gradient GEMM (adam_tile_kernel, see there)
  // (measured: 60.5 vs 57.4 us/step -- the update's 325 small blocks crowd the SMs the other GEMMs need; kept as a knob)
  const bool w4_split = p->knobs.w4_split && !b.bf16;
  int bn4 = w4_split ? p->bn_dw : bna;
  if (p->knobs.bn_dw4 == p->bn_dw || p->knobs.bn_dw4 == bna) bn4 = p->knobs.bn_dw4;
  int bn1 = bna;
  if (p->knobs.bn_dw1 == p->bn_dw || p->knobs.bn_dw1 == bna) bn1 = p->knobs.bn_dw1;
  Dw dw[4] = {
      {W4, B4, D, H, bn4, 0, &m->da4T, bn4 == bna ? &m->h3Ta : &m->h3T, p->da4T, p->W4T, D, w4_split ? "fc4_dW" : "fc4_dW_adam", 1},
      // fc3's update writes the transpose into the W3^T set latent_bwd is NOT reading (CUDA-core latent path): see below
      {W3, B3, H, Z, 32, 1, &m->da3T, &m->zT, p->da3T, p->simt_latent ? p->W3T_set[p->w3par ^ 1] : p->W3T, H, "fc3_dW_adam", 1},
      {W2X, B2X, 2 * Z, H, bna, 2, &m->dmlT, &m->h1Ta, p->dmlT, p->W2xT, 2 * Z, "fc2x_dW_adam", 1},
      {W1, B1, H, D, bn1, 3, &m->da1T, bn1 == bna ? (p->xpar ? &m->xTa_1 : &m->xTa) : (p->xpar ? &m->xT_1 : &m->xT), p->da1T, nullptr, 0,
       "fc1_dW_adam", 1},
  };
  Tc05Epi ed[4];
  int total_ctas = 0;
  for (int i = 0; i < 4; ++i) {
    Tc05Epi& e = ed[i];
    e = Tc05Epi{};
    const int64_t ow = b.L.off[dw[i].seg_w], ob = b.L.off[dw[i].seg_b];
    e.C = G + ow; e.ldc = dw[i].N; e.a_src = dw[i].a_src; e.lda_src = ldB; e.db = G + ob;
    e.bf16 = b.bf16;
    const bool fused = !(i == 0 && w4_split);
    if (fused) {
      e.adam_p = b.params + ow; e.adam_m = b.exp_avg + ow; e.adam_v = b.exp_avg_sq + ow;
      e.adam_pT = dw[i].pT; e.adam_ldpT = dw[i].ldpT;
      if (b.bf16) {       // the next step's GEMMs read bf16 copies of the updated weight (and of its transpose)
        __nv_bfloat16* const p16[4] = {p->W4_16, p->W3_16, p->W2x_16, p->W1_16};
        __nv_bfloat16* const pT16[4] = {p->W4T_16, p->W3T_16, p->W2xT_16, nullptr};
        const int ld16[4] = {H, p->ldz16, H, D};
        e.adam_p16 = p16[i]; e.adam_ldp16 = ld16[i]; e.adam_pT16 = pT16[i];
        if (i == 0) e.adam_pT = nullptr;             // W4^T has no fp32 reader in this mode
      }
      e.adam_pb = b.params + ob; e.adam_mb = b.exp_avg + ob; e.adam_vb = b.exp_avg_sq + ob;
      e.adam_st = b.st;
    }
    dw[i].ns = tc05_splits_dw(p, e, dw[i].M, dw[i].N, B, dw[i].bn, dw[i].slot);
    if (fused) total_ctas += ((dw[i].N + dw[i].bn - 1) / dw[i].bn) * ((dw[i].M + kTcBM - 1) / kTcBM) * dw[i].ns;
  }
  AdamTileArgs at{};
  int at_blocks = 0;
  if (w4_split) {
    const int64_t ow = b.L.off[W4], ob = b.L.off[B4];
    at.p = b.params + ow; at.m = b.exp_avg + ow; at.v = b.exp_avg_sq + ow; at.g = G + ow;
    at.pb = b.params + ob; at.mb = b.exp_avg + ob; at.vb = b.exp_avg_sq + ob; at.gb = G + ob;
    at.pT = p->W4T; at.ldpT = D; at.R = D; at.C = H; at.tiles_c = (H + 31) / 32;
    at.st = b.st;
    at_blocks = ((D + 31) / 32) * at.tiles_c;
    total_ctas += at_blocks;
    at.total_tickets = total_ctas;
    tc05_loss_partials(p, B, &at.part_a, &at.n_a, &at.part_b, &at.n_b);
    at.loss_out = b.loss_dev;
  }
  for (int i = 0; i < 4; ++i) {
    if (i == 0 && w4_split) continue;
    ed[i].adam_total_ctas = total_ctas;
    tc05_loss_partials(p, B, &ed[i].loss_part_a, &ed[i].n_part_a, &ed[i].loss_part_b, &ed[i].n_part_b);
    ed[i].adam_loss_out = b.loss_dev;
  }
  if (w4_split) at.total_tickets = total_ctas;
  auto launch_dw = [&](int i, cudaStream_t st) -> int {
    MDT_TRY(tc05_launch<EPI_GRADW>(dw[i].bn, sp, *dw[i].a, *dw[i].bm, dw[i].M, dw[i].N, B, ed[i], st, dw[i].ns));
    if (tc_recorder() == nullptr) MDT_TRY(b.note(b.note_ctx, dw[i].label));
    if (i == 0 && w4_split) {
      MDT_CUDA(launch_pdl(adam_tile_kernel, dim3(at_blocks), dim3(256), 0, st, at));
      MDT_TRY(b.note(b.note_ctx, "fc4_adam"));
    }
    return MDT_OK;
  };
  MDT_TRY(tc05_join_shadows(p, s));
  Tc05Epi e{};
  e.C = b.da3; e.ldc = H; e.CT = p->da3T; e.ldct = ldB; e.mask = b.h3; e.ldmask = H;
  e.pre_b = p->preload;                          // W4^T: the shadow was last written by the previous step's update
  e.bf16 = b.bf16; e.Cb = b.bf16 ? p->da3_16 : nullptr; e.ldcb = H;
  MDT_TRY(tc05_launch<EPI_RELU_MASK>(p->bn_h, sp, m->da4, p->mW4T, B, H, D, e, s, tc05_splits(p, e, B, H, D, p->bn_h)));
  TC_NOTE(p, "fc4_dX");
  if (p->simt_latent) {
    // No event between fc4's input-gradient GEMM and latent_bwd: a full (event) dependent next to the programmatic
    // one costs latent_bwd its early launch.  All four updates start together once latent_bwd is done -- fc4's is
    // not on the critical path, the last of the other three is.
    const bool dw4_early = !p->knobs.dw4_late;
    // fc3's update could start with fc4's, right after fc4's input-gradient GEMM: its operands (da3^T, z^T) are ready,
    // W3 has no reader left in this step, and the updated W3^T goes into the set latent_bwd is not reading.
    // Measured: NOT a gain -- 60.0 vs 58.8 us/step (tf32x3), 48.3 vs 47.5 (bf16): its 4 CTAs and fc4's 91 hold SMs that
    // latent_bwd and then fc1's update want.  Off by default (MDT_DW3_EARLY=1 turns it on).
    const bool dw3_early = dw4_early && p->knobs.dw3_early;
    if (dw4_early) {
      MDT_TRY(after_main(sa, dw3_early ? sb : nullptr));   // E1: W4^T no longer needed, da3^T ready
      MDT_TRY(launch_dw(0, sa));
      if (dw3_early) MDT_TRY(launch_dw(1, sb));
    }
    MDT_TRY(tc05_latent_bwd(p, s, B));            // reads W3T_set[w3par] and W2x^T
    if (fork) {
      MDT_CUDA(tc_event_record(p->ev_fork[nf], s));
      for (cudaStream_t t : {sa, sb, sc, sd})
        if (!((t == sa && dw4_early) || (t == sb && dw3_early))) MDT_CUDA(tc_stream_wait(t, p->ev_fork[nf]));
      ++nf;
    }
    if (!dw3_early) MDT_TRY(launch_dw(1, sb));
    MDT_TRY(launch_dw(2, sd));
    // fc1's update is the longest of the three: it stays on the main stream, where it is launched programmatically
    // behind latent_bwd and has its set-up (barriers, TMEM, descriptor and parameter fetches: ~2.5 us for an
    // event-launched kernel) done before latent_bwd finishes
    MDT_TRY(launch_dw(3, p->knobs.dw1_side ? sc : s));
    if (!dw4_early) MDT_TRY(launch_dw(0, sa));
  } else {
    MDT_TRY(after_main(sa, nullptr));             // E1: W4^T no longer needed, da3^T ready
    MDT_TRY(launch_dw(0, sa));
    e = Tc05Epi{};
    e.Z = Z; e.dml_kl = b.dml_kl; e.eps = b.eps; e.sd = b.sd; e.dml = b.dml; e.dmlT = p->dmlT; e.ld_dmlT = ldB;
    e.pre_a = p->preload;                        // W3^T
    e.bf16 = b.bf16; e.dmlb = b.bf16 ? p->dml16 : nullptr;
    MDT_TRY(tc05_launch<EPI_REPARAM_BWD>(p->bn_skinny, sp, p->mW3T, m->da3B, Z, B, H, e, s,
                                          tc05_splits(p, e, Z, B, H, p->bn_skinny)));
    TC_NOTE(p, "fc3_dX_reparam_bwd");
    MDT_TRY(after_main(sb, nullptr));             // E2: W3^T no longer needed, dml^T ready
    MDT_TRY(launch_dw(1, sb));
    e = Tc05Epi{};
    e.CT = p->da1T; e.ldct = ldB; e.mask = b.h1; e.ldmask = H;
    e.pre_b = p->preload;                        // W2x^T
    e.bf16 = b.bf16;
    MDT_TRY(tc05_launch<EPI_RELU_MASK>(p->bn_h, sp, m->dml, p->mW2xT, B, H, 2 * Z, e, s));
    TC_NOTE(p, "fc2x_dX");
    MDT_TRY(after_main(sc, sd));                  // E3: W2x^T no longer needed, da1^T ready
    MDT_TRY(launch_dw(3, sc));
    MDT_TRY(launch_dw(2, sd));
  }
  if (fork) {
    cudaStream_t st[4] = {sa, sb, sc, sd};
    cudaEvent_t ev[4] = {p->ev_join, p->ev_join2, p->ev_join3, p->ev_join4};
    for (int i = 0; i < 4; ++i) {
      MDT_CUDA(tc_event_record(ev[i], st[i]));
      MDT_CUDA(tc_stream_wait(s, ev[i]));
    }
  }
  if (p->simt_latent) p->w3par ^= 1;              // the next reader of W3^T takes the set this step's update wrote
  MDT_TRY(tc05_join_prefetch(p, s));
  return MDT_OK;
}

// Stand-alone y = act(x W^T + b) on the tensor cores (mdt_linear): builds its two tensor maps per call.
inline int tc05_linear(const float* x, const float* w, const float* bias, int M, int N, int K, int act, bool split,
                       float* y, cudaStream_t s) {
  MDT_CHECK(K % 4 == 0 && N % 4 == 0, "tcgen05 mdt_linear needs K %% 4 == 0 and N %% 4 == 0 (16-byte rows), got K=%d N=%d", K, N);
  const TcKnobs knobs = TcKnobs::from_env();     // stand-alone call: no plan, the switches as they stand now
  TcKnobScope knob_scope(&knobs);
  CUtensorMap ma, mb;
  const int bn = (N > 32) ? 64 : 32;
  MDT_TRY(tc05_encode(&ma, x, M, K, K, kTcBM));
  MDT_TRY(tc05_encode(&mb, w, N, K, K, bn));
  Tc05Epi e{};
  e.C = y; e.ldc = N; e.bias = bias; e.act = act;
  return tc05_launch<EPI_BIAS_ACT>(bn, split, ma, mb, M, N, K, e, s);
}

}  // namespace mdt
