// The VAE's two latent-side layer groups as fused CUDA-core kernels.
//
//   latent_fwd :  (mu | logvar) = h1 W2x^T + b2x ; z = mu + eps * exp(logvar / 2) ; KL terms ;
//                 h3 = relu(z W3^T + b3)                                  (ref: vae-hpo.py:31-41, 47-50, 60-66)
//   latent_bwd :  dz = da3 W3 ; d(mu | logvar) through the reparameterisation + KL ;
//                 da1 = (d(mu | logvar) W2x) * (h1 > 0)                   (autograd of the same lines)
//
// These layers have 2Z = 40 / Z = 20 output features: as tensor-core GEMMs they are one 128-row tile that is 70-85 %
// padding, split over K, reduced through a workspace -- 8 us each of pure latency for 4 MFLOP.  Here one CTA owns a
// few samples: it keeps their activation rows in shared memory, every warp forms whole dot products over K with
// coalesced weight-row loads (fp32 FMA, fixed lane-strided order + xor-shuffle tree: deterministic), and the narrow
// layer that follows (fc3, resp. fc21|fc22's input gradient, K = Z / 2Z) is done by the same CTA from shared memory.
// One launch replaces two GEMMs and their split-K reductions; the work is bound by the L2 -> SM traffic of the
// weight rows (96 KB per CTA), about 1.5 us.
#pragma once
#include <cuda_bf16.h>

#include "common.cuh"
#include "elementwise.cuh"

namespace mdt {

constexpr int kLatThreads = 512;      // 16 warps: the kernels are chains of short dependent phases, so warps hide latency
constexpr int kLatWarps = kLatThreads / 32;
constexpr int kLatMaxOut = 64;        // 2Z <= 64

struct LatentFwdArgs {
  const float* h1; int H;             // [B, H] row-major
  const float* W2x; const float* b2x; // [2Z, H], [2Z]
  const float* W3; const float* b3;   // [H, Z], [H]
  const float* eps_in;                // [B, Z], or NULL: eps is drawn here from Philox exactly as torch.randn_like would on
                                      // CUDA (ref:vae-hpo.py:35), from the device-resident (seed, offset); the kernel's last
                                      // block advances the offset
  StepState* st; long long philox_S; unsigned long long philox_inc;
  float* ml;                          // [B, 2Z]  mu | logvar
  float* z; void* zT; int ldT;        // [B, Z], [Z, ldT] (z^T / h3^T: bf16 instead of fp32 when `bf16` is set)
  float* eps_out; float* sd_out;      // [B, Z] (eps_out may equal eps_in)
  float* dml_kl;                      // [B, 2Z]  dKL/d(mu | logvar)
  float* kl_partials;                 // [gridDim.x]
  float* h3; void* h3T;               // [B, H], [H, ldT]
  int bf16; __nv_bfloat16* h3b;       // MDT_GEMM_BF16: row-major bf16 mirror of h3 (operand of fc4), or NULL
  int B, Z, nchunks;
  long long* dbg;                     // optional phase stamps (MDT_TC05_DBG=1), same slot layout as the GEMM kernel's
  int first_live_pass;                // always 1 (a run-time value, so the compiler cannot specialise the two passes)
  int first_pass;                     // 1: live pass only; 0: a dry run first (a measurement aid; the launchers pass 1)
};

struct LatentBwdArgs {
  const float* da3; int H;            // [B, H] row-major
  const float* W3T;                   // [Z, H]   (transposed shadow of fc3.weight)
  const float* W2xT;                  // [H, 2Z]  (transposed shadow of fc21|fc22.weight)
  const float* h1;                    // [B, H]   ReLU mask of fc1
  const float* eps; const float* sd; const float* dml_kl;
  float* dml; void* dmlT; int ldT;    // [B, 2Z], [2Z, ldT] (dml^T / da1^T: bf16 instead of fp32 when `bf16` is set)
  void* da1T;                         // [H, ldT]
  int bf16;
  int B, Z, nchunks;
  long long* dbg;
  int first_live_pass, first_pass;
};

__device__ __forceinline__ void lat_st_operand(void* base, size_t idx, float v, bool b16) {
  if (b16) reinterpret_cast<__nv_bfloat16*>(base)[idx] = __float2bfloat16_rn(v);
  else reinterpret_cast<float*>(base)[idx] = v;
}

// SC consecutive elements of a transposed operand row (idx is a multiple of SC, rows are 16-byte aligned): one vector
// store when all SC samples exist, scalar stores for a ragged last chunk.
template <int SC>
__device__ __forceinline__ void lat_st_run(void* base, size_t idx, const float* v, int ns, bool b16) {
  if (ns == SC && SC == 4) {
    if (b16) {
      const __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]), hi = __floats2bfloat162_rn(v[2], v[3]);
      uint2 u;
      u.x = *reinterpret_cast<const uint32_t*>(&lo);
      u.y = *reinterpret_cast<const uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(base) + idx) = u;
    } else {
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(base) + idx) = make_float4(v[0], v[1], v[SC > 2 ? 2 : 0], v[SC > 3 ? 3 : 0]);
    }
  } else if (ns == SC && SC == 2) {
    if (b16) {
      const __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]);
      *reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(base) + idx) = lo;
    } else {
      *reinterpret_cast<float2*>(reinterpret_cast<float*>(base) + idx) = make_float2(v[0], v[1]);
    }
  } else {
#pragma unroll
    for (int s = 0; s < SC; ++s)
      if (s < ns) lat_st_operand(base, idx + s, v[s], b16);
  }
}

__device__ __forceinline__ void lat_stamp(long long* dbg, int i) {
  if (dbg != nullptr && blockIdx.x == 0 && threadIdx.x == 0) dbg[i] = clock64();
}
__device__ __forceinline__ void lat_stamp_go(long long* dbg) {
  if (dbg != nullptr && threadIdx.x == 0) {
    unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
    atomicMin(reinterpret_cast<unsigned long long*>(dbg + 13), g);
  }
}
__device__ __forceinline__ void lat_stamp_end(long long* dbg) {
  if (dbg != nullptr && threadIdx.x == 0) {
    unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
    atomicMax(reinterpret_cast<unsigned long long*>(dbg + 15), g);
  }
}

// ---- weight fragments ------------------------------------------------------------------------------------------
// Everything a thread needs from the weight matrices is fetched into registers in one batch of independent loads, and
// for the first chunk of samples that batch is issued BEFORE griddepcontrol.wait: the weights were last written in
// the previous step (every kernel of this step that has started implies the previous step is complete), so their
// L2 latency overlaps the tail of the predecessor kernel instead of sitting on the critical path.
// The loads are `asm volatile`: nvcc is free to sink an ordinary ld.global.nc below the (asm) wait down to its first
// use -- it did -- while volatile asm statements keep their order relative to griddepcontrol.wait.
__device__ __forceinline__ float4 lat_ldg4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float lat_ldg(const float* p) {
  float v;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

constexpr int kLatU = 4;              // k-iterations (of 32 lanes x float4) fetched together by lat_dots

template <int OPW>
struct LatDotW { float4 w[kLatU][OPW]; float bias[OPW]; };

// Rows o = warp, warp + 16, ... (< nout) of a row-major [nout, H] matrix; lane l holds k = 4 (k4b + l + 32 u).
template <int OPW>
__device__ __forceinline__ void lat_load_dot_w(LatDotW<OPW>& f, const float* __restrict__ w, int H, int nout, int k4b,
                                               const float* __restrict__ bias = nullptr) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int H4 = H >> 2;
  if (k4b == 0) {
#pragma unroll
    for (int i = 0; i < OPW; ++i) f.bias[i] = (bias != nullptr && warp + kLatWarps * i < nout) ? lat_ldg(bias + warp + kLatWarps * i) : 0.f;
  }
#pragma unroll
  for (int u = 0; u < kLatU; ++u) {
    const int k4 = k4b + lane + 32 * u;
#pragma unroll
    for (int i = 0; i < OPW; ++i) {
      const int o = warp + kLatWarps * i;
      f.w[u][i] = (o < nout && k4 < H4) ? lat_ldg4(reinterpret_cast<const float4*>(w + (size_t)o * H) + k4)
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// Dot products of those rows with the SC activation rows in shared memory; `first` = fragment already loaded for
// k4b = 0.  Results (+ bias) land in out[s * ldo + o].  Fixed lane-strided order + xor-shuffle tree: deterministic.
template <int SC, int OPW>
__device__ __forceinline__ void lat_dots(LatDotW<OPW>& first, const float* __restrict__ w, int H, int nout, const float* sA,
                                         float* out, int ldo) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int H4 = H >> 2;
  float acc[OPW][SC];
#pragma unroll
  for (int i = 0; i < OPW; ++i)
#pragma unroll
    for (int s = 0; s < SC; ++s) acc[i][s] = 0.f;
  for (int k4b = 0; k4b < H4; k4b += 32 * kLatU) {
    if (k4b > 0) lat_load_dot_w<OPW>(first, w, H, nout, k4b);
#pragma unroll
    for (int u = 0; u < kLatU; ++u) {
      // no branch on k4 < H4 (the weights of a lane past the row end are zero, its activation index is clamped): a
      // divergent region here makes the compiler wrap every shuffle below in WARPSYNC.COLLECTIVE
      const int k4 = min(k4b + lane + 32 * u, H4 - 1);
#pragma unroll
      for (int s = 0; s < SC; ++s) {
        const float4 a = *reinterpret_cast<const float4*>(sA + (size_t)s * H + 4 * k4);
#pragma unroll
        for (int i = 0; i < OPW; ++i) {
          acc[i][s] = fmaf(first.w[u][i].x, a.x, acc[i][s]);
          acc[i][s] = fmaf(first.w[u][i].y, a.y, acc[i][s]);
          acc[i][s] = fmaf(first.w[u][i].z, a.z, acc[i][s]);
          acc[i][s] = fmaf(first.w[u][i].w, a.w, acc[i][s]);
        }
      }
    }
  }
  // all OPW * SC butterflies advance together: the shuffles of one level are independent and pipeline
#pragma unroll
  for (int d = 16; d > 0; d >>= 1)
#pragma unroll
    for (int i = 0; i < OPW; ++i)
#pragma unroll
      for (int s = 0; s < SC; ++s) acc[i][s] += __shfl_xor_sync(0xffffffffu, acc[i][s], d);
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < OPW; ++i) {
      const int o = warp + kLatWarps * i;
      if (o < nout) {
#pragma unroll
        for (int s = 0; s < SC; ++s) out[s * ldo + o] = acc[i][s] + first.bias[i];
      }
    }
  }
}

// One row of a narrow row-major matrix [H, n] (n = 4 * nq floats, 16-byte aligned rows): the thread's own hidden unit.
template <int NQ>
struct LatRowW { float4 w[NQ]; };

template <int NQ>
__device__ __forceinline__ void lat_load_row_w(LatRowW<NQ>& f, const float* __restrict__ w, int n, int row, int H) {
  const int nq = n >> 2;
#pragma unroll
  for (int q = 0; q < NQ; ++q)
    f.w[q] = (row < H && q < nq) ? lat_ldg4(reinterpret_cast<const float4*>(w + (size_t)row * n) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
}

// acc[s] += sum_k row[k] * v[s * ldv + k], k ascending
template <int SC, int NQ>
__device__ __forceinline__ void lat_row_dot(const LatRowW<NQ>& f, int n, const float* v, int ldv, float* acc) {
  const int nq = n >> 2;
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    if (q < nq) {
#pragma unroll
      for (int s = 0; s < SC; ++s) {
        const float4 vv = *reinterpret_cast<const float4*>(v + s * ldv + 4 * q);   // 16-byte aligned broadcast read
        acc[s] = fmaf(f.w[q].x, vv.x, acc[s]);
        acc[s] = fmaf(f.w[q].y, vv.y, acc[s]);
        acc[s] = fmaf(f.w[q].z, vv.z, acc[s]);
        acc[s] = fmaf(f.w[q].w, vv.w, acc[s]);
      }
    }
  }
}

// Rows m0 .. m0 + ns of a row-major [B, H] matrix -> shared memory (float4, coalesced); missing rows are zero.
template <int SC>
__device__ __forceinline__ void lat_load_rows(const float* __restrict__ src, int H, int m0, int ns, float* sA) {
  const int H4 = H >> 2;
  for (int i = threadIdx.x; i < SC * H4; i += kLatThreads) {
    const int s = i / H4, c = i - s * H4;
    reinterpret_cast<float4*>(sA)[i] =
        (s < ns) ? __ldg(reinterpret_cast<const float4*>(src + (size_t)(m0 + s) * H) + c) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// ZQ: compile-time bound on Z / 4 (5 for Z <= 20, 8 for Z <= 32); the real Z is a run-time argument.
// bx / gx: this CTA's index and the CTA count of ITS trial (blockIdx.x / gridDim.x, or their per-trial parts in the
// grouped launch of several trials)
template <int SC, int ZQ>
__device__ __forceinline__ void latent_fwd_body(const LatentFwdArgs& a, const int bx, const int gx) {
  extern __shared__ __align__(16) float lat_smem[];
  float* sA = lat_smem;                               // [SC][H]
  __shared__ __align__(16) float sML[SC * kLatMaxOut];
  __shared__ __align__(16) float sZ[SC * 32];
  __shared__ float red[32];
  const int H = a.H, Z = a.Z;
  constexpr int OPW = (8 * ZQ + kLatWarps - 1) / kLatWarps;   // 2Z outputs over the warps
  lat_stamp(a.dbg, 0);
  LatDotW<OPW> fw;
  LatRowW<ZQ> f3[1];                                  // W3 row of hidden unit t
  lat_load_dot_w<OPW>(fw, a.W2x, H, 2 * Z, 0, a.b2x);
  lat_load_row_w<ZQ>(f3[0], a.W3, Z, threadIdx.x, H);
  // also safe before the wait: eps (written by the step's prologue, two kernels back) and the fc3 bias
  float b3r[1], e_pre = 0.f;
#pragma unroll
  for (int r = 0; r < 1; ++r) b3r[r] = (threadIdx.x + r * kLatThreads < H) ? lat_ldg(a.b3 + threadIdx.x + r * kLatThreads) : 0.f;
  // (seed, offset) were last written by the previous step's latent_fwd or by the host: safe before the wait, and the
  // Philox rounds of the first chunk's eps overlap the predecessor kernel
  const bool gen_eps = a.eps_in == nullptr;
  unsigned long long ph_seed = 0, ph_off = 0;
  if (gen_eps) { ph_seed = a.st->philox_seed; ph_off = a.st->philox_offset; }
  {
    const int m0 = bx * SC, ns = min(SC, a.B - m0);
    if ((int)threadIdx.x < ns * Z)
      e_pre = gen_eps ? torch_cuda_randn_elem(ph_seed, ph_off, (long long)m0 * Z + threadIdx.x, a.philox_S)
                      : lat_ldg(a.eps_in + (size_t)m0 * Z + threadIdx.x);
  }
  // (The pass loop below can start with a dry run of the same code before the dependency wait -- first_pass = 0 --
  // to pull the instructions into the caches early.  Measured: no gain, the kernel is not fetch-bound; left as a knob.)
  float kl = 0.f;
  bool fw_ok = true;
#pragma unroll 1
  for (int pass = a.first_pass; pass < 2; ++pass) {
  const bool live = pass >= a.first_live_pass;        // uniform, but not a compile-time constant
  if (live) {
    pdl_wait();
    pdl_trigger();
    lat_stamp(a.dbg, 1);
    lat_stamp_go(a.dbg);
  }
  bool first = true;
  const int chunk_end = live ? a.nchunks : min(a.nchunks, bx + 1);
  for (int chunk = bx; chunk < chunk_end; chunk += gx) {
    const int m0 = chunk * SC, ns = live ? min(SC, a.B - m0) : 0;   // dry run: every store is predicated off by ns = 0
    if (live) lat_load_rows<SC>(a.h1, H, m0, ns, sA);
    if (!fw_ok) lat_load_dot_w<OPW>(fw, a.W2x, H, 2 * Z, 0, a.b2x);
    __syncthreads();
    lat_stamp(a.dbg, 2);
    lat_dots<SC, OPW>(fw, a.W2x, H, 2 * Z, sA, sML, kLatMaxOut);
    fw_ok = (H >> 2) <= 32 * kLatU;                    // longer rows: the fragment now holds the last k-block
    __syncthreads();
    lat_stamp(a.dbg, 3);
    for (int idx = threadIdx.x; idx < SC * Z; idx += kLatThreads) {
      const int s = idx / Z, j = idx - s * Z;
      const bool ok = s < ns;
      const size_t m = ok ? (size_t)(m0 + s) : 0;
      const float mu = sML[s * kLatMaxOut + j], lv = sML[s * kLatMaxOut + Z + j];
      const float e = (first || !ok) ? e_pre        // SC * Z <= 128 < blockDim.x: one element per thread
                      : (gen_eps ? torch_cuda_randn_elem(ph_seed, ph_off, (long long)m * Z + j, a.philox_S) : __ldg(a.eps_in + m * Z + j));
      const float sd = __expf(0.5f * lv);
      const float ev = __expf(lv);
      const float zv = mu + e * sd;
      if (ok) {
        a.ml[m * 2 * Z + j] = mu;
        a.ml[m * 2 * Z + Z + j] = lv;
        a.z[m * Z + j] = zv;
        lat_st_operand(a.zT, (size_t)j * a.ldT + m, zv, a.bf16 != 0);
        if (a.eps_out != a.eps_in) a.eps_out[m * Z + j] = e;
        a.sd_out[m * Z + j] = sd;
        a.dml_kl[m * 2 * Z + j] = mu;
        a.dml_kl[m * 2 * Z + Z + j] = 0.5f * (ev - 1.0f);
        kl += -0.5f * (1.0f + lv - mu * mu - ev);
      }
      sZ[s * 32 + j] = zv;
    }
    __syncthreads();
    lat_stamp(a.dbg, 4);
    // fc3: thread = hidden unit j; its weight row W3[j, :Z] sits in registers for the first 512 units
    for (int jb = 0; jb < H; jb += kLatThreads) {
#pragma unroll
      for (int r = 0; r < 1; ++r) {
        const int j = jb + r * kLatThreads + threadIdx.x;
        if (jb > 0) lat_load_row_w<ZQ>(f3[r], a.W3, Z, j, H);
        if (j < H) {
          float acc[SC];
          const float bj = (jb == 0) ? b3r[r] : __ldg(a.b3 + j);
#pragma unroll
          for (int s = 0; s < SC; ++s) acc[s] = bj;
          lat_row_dot<SC, ZQ>(f3[r], Z, sZ, 32, acc);
#pragma unroll
          for (int s = 0; s < SC; ++s) {
            acc[s] = fmaxf(acc[s], 0.f);
            if (s < ns) {
              a.h3[(size_t)(m0 + s) * H + j] = acc[s];
              if (a.h3b != nullptr) a.h3b[(size_t)(m0 + s) * H + j] = __float2bfloat16_rn(acc[s]);
            }
          }
          lat_st_run<SC>(a.h3T, (size_t)j * a.ldT + m0, acc, ns, a.bf16 != 0);
        }
      }
    }
    if (H > kLatThreads) lat_load_row_w<ZQ>(f3[0], a.W3, Z, threadIdx.x, H);   // the row registers were reused: restore
    first = false;
    __syncthreads();                                   // sA / sML / sZ are rewritten by the next chunk
  }
  }
  lat_stamp(a.dbg, 5);
  kl = block_sum(kl, red);
  if (threadIdx.x == 0) a.kl_partials[bx] = kl;
  if (gen_eps && threadIdx.x == 0) {                  // every block has read the offset by the time the last ticket is taken
    __threadfence();
    if (atomicAdd(&a.st->ticket_reparam, 1u) == (unsigned)gx - 1u) {
      a.st->philox_offset = ph_off + a.philox_inc;
      a.st->ticket_reparam = 0;
      __threadfence();
    }
  }
  lat_stamp(a.dbg, 6);
  lat_stamp_end(a.dbg);
}

template <int SC, int ZQ>
__global__ void __launch_bounds__(kLatThreads) latent_fwd_kernel(LatentFwdArgs a) {
  latent_fwd_body<SC, ZQ>(a, (int)blockIdx.x, (int)gridDim.x);
}
// Grouped launch: blockIdx.x = trial * blocks_per_trial + block; arguments of trial t in table[t].
template <int SC, int ZQ>
__global__ void __launch_bounds__(kLatThreads) latent_fwd_group_kernel(const LatentFwdArgs* __restrict__ table, int blocks_per_trial) {
  const int trial = (int)blockIdx.x / blocks_per_trial;
  const LatentFwdArgs a = table[trial];
  latent_fwd_body<SC, ZQ>(a, (int)blockIdx.x - trial * blocks_per_trial, blocks_per_trial);
}

template <int SC, int ZQ>
__device__ __forceinline__ void latent_bwd_body(const LatentBwdArgs& a, const int bx, const int gx) {
  extern __shared__ __align__(16) float lat_smem[];
  float* sA = lat_smem;                               // [SC][H]
  __shared__ __align__(16) float sDZ[SC * 32];
  __shared__ __align__(16) float sDML[SC * kLatMaxOut];
  const int H = a.H, Z = a.Z;
  constexpr int OPW = (4 * ZQ + kLatWarps - 1) / kLatWarps;   // Z outputs over the warps
  lat_stamp(a.dbg, 0);
  LatDotW<OPW> fw;
  LatRowW<2 * ZQ> f2[1];                              // W2x^T row of hidden unit t
  lat_load_dot_w<OPW>(fw, a.W3T, H, Z, 0);
  lat_load_row_w<2 * ZQ>(f2[0], a.W2xT, 2 * Z, threadIdx.x, H);
  // per-sample inputs of the first chunk (all written by forward kernels of this step, long complete)
  float kmu_pre = 0.f, klv_pre = 0.f, es_pre = 0.f, mk_pre[1][SC];
  {
    const int m0 = bx * SC, ns = min(SC, a.B - m0);
    if ((int)threadIdx.x < ns * Z) {
      const int s = threadIdx.x / Z, j = threadIdx.x - s * Z;
      const size_t m = (size_t)(m0 + s);
      kmu_pre = lat_ldg(a.dml_kl + m * 2 * Z + j);
      klv_pre = lat_ldg(a.dml_kl + m * 2 * Z + Z + j);
      es_pre = lat_ldg(a.eps + m * Z + j) * lat_ldg(a.sd + m * Z + j);
    }
#pragma unroll
    for (int r = 0; r < 1; ++r)
#pragma unroll
      for (int s = 0; s < SC; ++s) {
        const int h = threadIdx.x + r * kLatThreads;
        mk_pre[r][s] = (h < H && s < ns) ? lat_ldg(a.h1 + (size_t)(m0 + s) * H + h) : 0.f;
      }
  }
  bool fw_ok = true;
#pragma unroll 1
  for (int pass = a.first_pass; pass < 2; ++pass) {     // pass 0: optional dry run (see latent_fwd)
  const bool live = pass >= a.first_live_pass;
  if (live) {
    pdl_wait();
    pdl_trigger();
    lat_stamp(a.dbg, 1);
    lat_stamp_go(a.dbg);
  }
  bool first = true;
  const int chunk_end = live ? a.nchunks : min(a.nchunks, bx + 1);
  for (int chunk = bx; chunk < chunk_end; chunk += gx) {
    const int m0 = chunk * SC, ns = live ? min(SC, a.B - m0) : 0;
    if (live) lat_load_rows<SC>(a.da3, H, m0, ns, sA);
    if (!fw_ok) lat_load_dot_w<OPW>(fw, a.W3T, H, Z, 0);
    __syncthreads();
    lat_stamp(a.dbg, 2);
    lat_dots<SC, OPW>(fw, a.W3T, H, Z, sA, sDZ, 32);
    fw_ok = (H >> 2) <= 32 * kLatU;
    __syncthreads();
    lat_stamp(a.dbg, 3);
    for (int idx = threadIdx.x; idx < SC * Z; idx += kLatThreads) {
      const int s = idx / Z, j = idx - s * Z;
      const bool ok = s < ns;
      const size_t m = ok ? (size_t)(m0 + s) : 0;
      const float g = sDZ[s * 32 + j];
      const bool pre = first || !ok;
      const float kmu = pre ? kmu_pre : __ldg(a.dml_kl + m * 2 * Z + j);
      const float klv = pre ? klv_pre : __ldg(a.dml_kl + m * 2 * Z + Z + j);
      const float es = pre ? es_pre : __ldg(a.eps + m * Z + j) * __ldg(a.sd + m * Z + j);
      const float dmu = g + kmu;
      const float dlv = g * es * 0.5f + klv;
      if (ok) {
        a.dml[m * 2 * Z + j] = dmu;
        a.dml[m * 2 * Z + Z + j] = dlv;
        lat_st_operand(a.dmlT, (size_t)j * a.ldT + m, dmu, a.bf16 != 0);
        lat_st_operand(a.dmlT, (size_t)(Z + j) * a.ldT + m, dlv, a.bf16 != 0);
      }
      sDML[s * kLatMaxOut + j] = dmu;
      sDML[s * kLatMaxOut + Z + j] = dlv;
    }
    __syncthreads();
    lat_stamp(a.dbg, 4);
    // da1[s, h] = sum_o dml[s, o] W2x[o, h], masked by fc1's ReLU; thread = hidden unit h, its W2x^T row in registers
    for (int hb = 0; hb < H; hb += kLatThreads) {
#pragma unroll
      for (int r = 0; r < 1; ++r) {
        const int h = hb + r * kLatThreads + threadIdx.x;
        if (hb > 0) lat_load_row_w<2 * ZQ>(f2[r], a.W2xT, 2 * Z, h, H);
        if (h < H) {
          float mk[SC];
#pragma unroll
          for (int s = 0; s < SC; ++s)
            mk[s] = (first && hb == 0) ? mk_pre[r][s] : ((s < ns) ? __ldg(a.h1 + (size_t)(m0 + s) * H + h) : 0.f);
          float acc[SC];
#pragma unroll
          for (int s = 0; s < SC; ++s) acc[s] = 0.f;
          lat_row_dot<SC, 2 * ZQ>(f2[r], 2 * Z, sDML, kLatMaxOut, acc);
#pragma unroll
          for (int s = 0; s < SC; ++s) acc[s] = mk[s] > 0.f ? acc[s] : 0.f;
          lat_st_run<SC>(a.da1T, (size_t)h * a.ldT + m0, acc, ns, a.bf16 != 0);
        }
      }
    }
    if (H > kLatThreads) lat_load_row_w<2 * ZQ>(f2[0], a.W2xT, 2 * Z, threadIdx.x, H);
    first = false;
    __syncthreads();
  }
  }
  lat_stamp(a.dbg, 6);
  lat_stamp_end(a.dbg);
}

template <int SC, int ZQ>
__global__ void __launch_bounds__(kLatThreads) latent_bwd_kernel(LatentBwdArgs a) {
  latent_bwd_body<SC, ZQ>(a, (int)blockIdx.x, (int)gridDim.x);
}
template <int SC, int ZQ>
__global__ void __launch_bounds__(kLatThreads) latent_bwd_group_kernel(const LatentBwdArgs* __restrict__ table, int blocks_per_trial) {
  const int trial = (int)blockIdx.x / blocks_per_trial;
  const LatentBwdArgs a = table[trial];
  latent_bwd_body<SC, ZQ>(a, (int)blockIdx.x - trial * blocks_per_trial, blocks_per_trial);
}

// Samples per CTA.  A CTA takes a whole SM (512 threads x 128 registers), and the kernels only pay off when all their
// CTAs are resident before the predecessor finishes (so the weight preload is off the critical path): forward shares the
// GPU with fc1's ~78 CTAs, backward with fc4's input-gradient GEMM and then with the 91 CTAs of fc4's weight-gradient
// GEMM on the side stream -- measured at B = 128: forward 4.2 / 3.3 / 4.6 us for 1 / 2 / 4 samples per CTA, backward
// 12.3 / 10.0 / 4.2 us (with 1 or 2 the last CTAs queue behind the weight-gradient GEMM).
// `forced`: MDT_LATENT_SC as read when the plan was created (0 = choose)
inline int latent_pick_sc(int B, int sm_count, bool backward = false, int forced = 0) {
  if (forced == 1 || forced == 2 || forced == 4) return forced;
  const int room = backward ? sm_count / 3 : sm_count / 2;
  return B <= room ? 1 : (B <= 2 * room ? 2 : 4);
}
inline int latent_grid(int B, int sc) { const int n = (B + sc - 1) / sc; return n < 1024 ? n : 1024; }

// Launch with `smem` bytes of dynamic shared memory, raising the kernel's limit first when it is above the 48 KB default.
template <class K, class A>
inline cudaError_t latent_launch(K kernel, size_t& limit, dim3 grid, size_t smem, cudaStream_t s, const A& a) {
  if (smem > 48 * 1024 && smem > limit) {
    const cudaError_t ce = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (ce != cudaSuccess) return ce;
    limit = smem;
  }
  return launch_pdl(kernel, grid, dim3(kLatThreads), smem, s, a);
}

template <int ZQ>
inline cudaError_t launch_latent_fwd_z(const LatentFwdArgs& a, int sc, dim3 grid, size_t smem, cudaStream_t s) {
  static size_t lim[3] = {0, 0, 0};
  if (sc == 1) return latent_launch(latent_fwd_kernel<1, ZQ>, lim[0], grid, smem, s, a);
  if (sc == 2) return latent_launch(latent_fwd_kernel<2, ZQ>, lim[1], grid, smem, s, a);
  return latent_launch(latent_fwd_kernel<4, ZQ>, lim[2], grid, smem, s, a);
}
inline cudaError_t launch_latent_fwd(LatentFwdArgs a, int sc, cudaStream_t s) {
  a.nchunks = (a.B + sc - 1) / sc;
  a.first_live_pass = 1;
  a.first_pass = 1;      // (0 would add a dry run of the pass before the dependency wait: measured no gain, see latent_fwd_kernel)
  const dim3 grid(latent_grid(a.B, sc));
  const size_t smem = (size_t)sc * a.H * sizeof(float);
  return a.Z <= 20 ? launch_latent_fwd_z<5>(a, sc, grid, smem, s) : launch_latent_fwd_z<8>(a, sc, grid, smem, s);
}

template <int ZQ>
inline cudaError_t launch_latent_bwd_z(const LatentBwdArgs& a, int sc, dim3 grid, size_t smem, cudaStream_t s) {
  static size_t lim[3] = {0, 0, 0};
  if (sc == 1) return latent_launch(latent_bwd_kernel<1, ZQ>, lim[0], grid, smem, s, a);
  if (sc == 2) return latent_launch(latent_bwd_kernel<2, ZQ>, lim[1], grid, smem, s, a);
  return latent_launch(latent_bwd_kernel<4, ZQ>, lim[2], grid, smem, s, a);
}
inline cudaError_t launch_latent_bwd(LatentBwdArgs a, int sc, cudaStream_t s) {
  a.nchunks = (a.B + sc - 1) / sc;
  a.first_live_pass = 1;
  a.first_pass = 1;      // (0 would add a dry run of the pass before the dependency wait: measured no gain, see latent_fwd_kernel)
  const dim3 grid(latent_grid(a.B, sc));
  const size_t smem = (size_t)sc * a.H * sizeof(float);
  return a.Z <= 20 ? launch_latent_bwd_z<5>(a, sc, grid, smem, s) : launch_latent_bwd_z<8>(a, sc, grid, smem, s);
}

}  // namespace mdt
