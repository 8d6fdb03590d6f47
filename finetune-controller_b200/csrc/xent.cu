// Fused softmax-cross-entropy forward + backward over bf16 logits, in place.
//
// Two kernels, chosen by row length (launch_xent below; algorithmic bytes 2*rows*ld*2 + 12*rows either way):
//
//   softmax_xent_staged_kernel   rows of at most 32768 logits.  The row is brought into shared memory by one bulk copy
//       (TMA), three rows deep, passes through registers ONCE (one MUFU exponential per logit) and leaves by one bulk
//       store: DRAM traffic is exactly one read + one write.
//   softmax_xent_kernel          longer rows.  One CTA per row, two sweeps: online (max, sum) from HBM, then the
//       gradient with the row re-read from L2.  The number of CTAs per SM is capped so that the rows in flight fit
//       the 126 MB L2 (uncapped, V = 128256 put 303 MB in flight and the second sweep went back to HBM).
//
// Targets outside [0, vocab) that are not ignore_index are treated exactly like ignored ones — no loss, zero
// gradient, not counted in n_valid — and counted in scratch word [2] so the host can report them.
//
// Scratch (the caller's workspace): [0] number of valid targets, [1] completion ticket, [2] number of
// out-of-range targets, [16 .. 16+rows) per-row losses.  The last CTA to finish folds the row losses in a fixed
// order, so the loss is bit-reproducible from run to run.
#include <stdlib.h>
#include <mutex>
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

constexpr int kUnroll = 4;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

__global__ void __launch_bounds__(256)
count_valid_kernel(const int64_t* __restrict__ targets, int64_t rows, int64_t ignore_index, int vocab,
                   unsigned int* __restrict__ ws) {
  grid_dep_sync();
  // single CTA: rows is at most a few 10^5, this is ~1 us and keeps the count exact
  unsigned int c = 0, bad = 0;
  for (int64_t i = threadIdx.x; i < rows; i += blockDim.x) {
    const int64_t t = targets[i];
    const bool in_range = t >= 0 && t < vocab;
    c += (t != ignore_index && in_range);
    bad += (t != ignore_index && !in_range);
  }
  __shared__ unsigned int red[8], redb[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    c += __shfl_xor_sync(0xffffffffu, c, o);
    bad += __shfl_xor_sync(0xffffffffu, bad, o);
  }
  if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = c; redb[threadIdx.x >> 5] = bad; }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = 0, b = 0;
    for (int w = 0; w < 8; ++w) { t += red[w]; b += redb[w]; }
    ws[0] = t;
    ws[1] = 0;  // ticket
    ws[2] = b;  // targets that are neither ignore_index nor inside the vocabulary
  }
}

struct MS { float m, s; };
__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ MS ms_merge(MS a, MS b) {
  float M = fmaxf(a.m, b.m);
  MS r;
  r.m = M;
  // guard (-inf) - (-inf): a fully masked partial contributes 0
  float ea = (a.m == -INFINITY) ? 0.f : exp2f(a.m - M);
  float eb = (b.m == -INFINITY) ? 0.f : exp2f(b.m - M);
  r.s = a.s * ea + b.s * eb;
  return r;
}

template <int kThreads>
__global__ void __launch_bounds__(kThreads)
softmax_xent_kernel(__nv_bfloat16* __restrict__ logits, int64_t rows, int vocab, int64_t ld,
                    const int64_t* __restrict__ targets, int64_t ignore_index, float grad_scale,
                    float* __restrict__ loss_out, unsigned int* __restrict__ ws) {
  grid_dep_sync();
  const int64_t row = blockIdx.x;
  __nv_bfloat16* z = logits + row * ld;
  const int nvec = (int)(ld >> 3);  // the whole padded row is rewritten
  const int64_t tgt = targets[row];
  const bool valid = (tgt != ignore_index) && tgt >= 0 && tgt < vocab;
  float* row_loss = reinterpret_cast<float*>(ws + 16);
  __shared__ MS red[kThreads / 32];
  __shared__ float s_lse2;
  __shared__ bool s_last;

  if (valid) {
    // ---- sweep 1: online max / sum in the log2 domain, kUnroll independent 16-byte loads in
    // flight per thread (the running max is revisited once per 8*kUnroll elements) ----
    // Packed math throughout: the maximum on bf16 pairs (HMNMX2), then per pair two unpack ops, one FFMA2, two MUFU and
    // one FADD2 — half the issue slots of the scalar form, which matters inside a step, where the SM clock sits at
    // the power cap and this kernel's issue rate, not HBM, sets its time.
    MS acc{-INFINITY, 0.f};
    const f32x2_t l2e2 = pk2(kLog2e, kLog2e);
    for (int i0 = threadIdx.x; i0 < nvec; i0 += kThreads * kUnroll) {
      uint4 q[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {   // all loads of the iteration in flight before the first use
        const int i = i0 + u * kThreads;
        q[u] = make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);   // (-inf, -inf) pairs in bf16
        if (i < nvec) q[u] = ld_cg16(z + ((int64_t)i << 3));
      }
      uint32_t mx2 = 0xff80ff80u;
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int c0 = (i0 + u * kThreads) << 3;
        if (c0 + 8 > vocab && c0 < (nvec << 3)) {  // the vector straddling (or beyond) the vocabulary: padding -> -inf
          uint32_t* w = reinterpret_cast<uint32_t*>(&q[u]);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (c0 + 2 * j >= vocab) w[j] = (w[j] & 0xffff0000u) | 0xff80u;
            if (c0 + 2 * j + 1 >= vocab) w[j] = (w[j] & 0x0000ffffu) | 0xff800000u;
          }
        }
        mx2 = bf16x2_max(bf16x2_max(mx2, q[u].x), bf16x2_max(q[u].y, bf16x2_max(q[u].z, q[u].w)));
      }
      const float vmax = fmaxf(bf16lo(mx2), bf16hi(mx2)) * kLog2e;
      if (vmax > acc.m) {  // rare after the first few iterations
        acc.s *= ex2_approx(acc.m - vmax);  // ex2(-inf) = 0 on the first hit
        acc.m = vmax;
      }
      if (acc.m != -INFINITY) {
        const f32x2_t nm2 = pk2(-acc.m, -acc.m);
        f32x2_t s2a = pk2(0.f, 0.f), s2b = pk2(0.f, 0.f);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const uint32_t* w = reinterpret_cast<const uint32_t*>(&q[u]);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float lo, hi;
            upk2(fma2(bf16x2_to_f32x2(w[j]), l2e2, nm2), lo, hi);
            const f32x2_t e = pk2(ex2_approx(lo), ex2_approx(hi));   // padding: exp2(-inf) = 0
            if (j & 1) s2b = add2(s2b, e); else s2a = add2(s2a, e);
          }
        }
        float s_lo, s_hi;
        upk2(add2(s2a, s2b), s_lo, s_hi);
        acc.s += s_lo + s_hi;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      MS other;
      other.m = __shfl_xor_sync(0xffffffffu, acc.m, o);
      other.s = __shfl_xor_sync(0xffffffffu, acc.s, o);
      acc = ms_merge(acc, other);
    }
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      MS t = red[0];
      for (int w = 1; w < kThreads / 32; ++w) t = ms_merge(t, red[w]);
      float lse2 = t.m + log2f(t.s);
      s_lse2 = lse2;
      float zt = __bfloat162float(z[tgt]) * kLog2e;
      row_loss[row] = (lse2 - zt) * kLn2;
    }
    __syncthreads();
  } else if (threadIdx.x == 0) {
    row_loss[row] = 0.f;
  }

  // ---- sweep 2: write the gradient in place (row re-read from L2) ----
  const unsigned int n_valid = ws[0];
  const float scale = valid ? grad_scale / (float)(n_valid > 0 ? n_valid : 1) : 0.f;
  const float lse2 = valid ? s_lse2 : 0.f;
  const int tcol = valid ? (int)tgt : -1;
  if (valid) {
    const f32x2_t l2e2s = pk2(kLog2e, kLog2e), nl2 = pk2(-lse2, -lse2), sc2 = pk2(scale, scale);
    for (int i0 = threadIdx.x; i0 < nvec; i0 += kThreads * kUnroll) {
      uint4 q[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int i = i0 + u * kThreads;
        if (i < nvec) q[u] = ld_cg16(z + ((int64_t)i << 3));
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int i = i0 + u * kThreads;
        if (i < nvec) {
          const int c0 = i << 3;
          const uint32_t* w = reinterpret_cast<const uint32_t*>(&q[u]);
          uint4 o;
          uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
          if ((tcol >= c0 && tcol < c0 + 8) || c0 + 8 > vocab) {   // the vector holding the target, or padding: scalar
            float f[8];
            unpack8(q[u], f);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              f[j] = ex2_approx(f[j] * kLog2e - lse2) * scale;
              if (c0 + j == tcol) f[j] -= scale;
              if (c0 + j >= vocab) f[j] = 0.f;
            }
            o = pack8(f);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float lo, hi;
              upk2(fma2(bf16x2_to_f32x2(w[j]), l2e2s, nl2), lo, hi);
              ow[j] = f32x2_to_bf16x2(mul2(pk2(ex2_approx(lo), ex2_approx(hi)), sc2));
            }
          }
          st_stream16(z + ((int64_t)i << 3), o);
        }
      }
    }
  } else {
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    for (int i = threadIdx.x; i < nvec; i += kThreads) st_stream16(z + ((int64_t)i << 3), zero);
  }

  // ---- last CTA folds the per-row losses ----
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = (atomicAdd(ws + 1, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    __threadfence();
    float t = 0.f;
    for (int64_t i = threadIdx.x; i < rows; i += kThreads) t += __ldcg(row_loss + i);
    t = warp_sum(t);
    __shared__ float fr[kThreads / 32];
    if ((threadIdx.x & 31) == 0) fr[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
      float sum = 0.f;
      for (int w = 0; w < kThreads / 32; ++w) sum += fr[w];
      loss_out[0] = sum / (float)(n_valid > 0 ? n_valid : 1);
    }
  }
}

// ---------------------------------------------------------------------------------------
// Staged kernel (rows of at most 32768 logits): the row stays on the chip between its read and its write.
// ---------------------------------------------------------------------------------------
constexpr int kCThreads = 1024;
constexpr int kCU = 4;                           // 16-byte vectors (8 logits each) per thread
constexpr int kCChunkVecs = kCThreads * kCU;     // 4096 vectors = 32768 logits = 64 KB per row
constexpr int kCStages = 3;
constexpr int kCStageBytes = kCChunkVecs * 16;
constexpr int kCSmemBytes = kCStages * kCStageBytes;   // 192 KB dynamic; barriers are static
constexpr long long kCWaitLimit = 4000000000ll;  // clock64 ticks (~2 s) before a barrier wait traps

__device__ __forceinline__ uint32_t cx_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cx_mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = 0;
  while (true) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > kCWaitLimit) __trap();   // never hang the GPU
  }
}

// Persistent: one CTA per SM walks rows r = blockIdx.x, += gridDim.x.  A row is brought into shared memory by ONE bulk
// copy (TMA), three rows deep, so the HBM reads of the next two rows are in flight while the current one is processed.
// Each of the 1024 threads takes its 32 logits into registers once: CTA maximum first (packed-bf16 HMNMX2, shuffles,
// one barrier), then e = exp2(z*log2e - max) — ONE MUFU op per logit; a re-reading kernel needs two, and at 16/clk/SM
// the special-function unit is what the two-sweep kernel runs into — then the sum, the scaling by grad_scale /
// (n_valid * sum) in registers, bf16 back into the same shared-memory bytes and one bulk store to HBM.  DRAM traffic is
// exactly one read and one write of the row, whatever L2 does.
__global__ void __launch_bounds__(kCThreads, 1)
softmax_xent_staged_kernel(__nv_bfloat16* __restrict__ logits, int64_t rows, int vocab, int64_t ld,
                           const int64_t* __restrict__ targets, int64_t ignore_index, float grad_scale,
                           float* __restrict__ loss_out, unsigned int* __restrict__ ws) {
  grid_dep_sync();
  extern __shared__ __align__(128) uint8_t cx_smem[];
  __shared__ __align__(8) unsigned long long full_bar[kCStages];
  __shared__ float red_m[kCThreads / 32], red_s[kCThreads / 32];
  __shared__ bool s_last;
  __shared__ float fr[kCThreads / 32];

  const int tid = threadIdx.x;
  const int64_t G = gridDim.x;
  const int nvec = (int)(ld >> 3);
  const uint32_t row_bytes = (uint32_t)nvec * 16u;
  const uint32_t smem0 = cx_smem_u32(cx_smem);
  float* row_loss = reinterpret_cast<float*>(ws + 16);
  const unsigned int n_valid = ws[0];
  const float scale = grad_scale / (float)(n_valid > 0 ? n_valid : 1);

  if (tid == 0) {
    for (int s = 0; s < kCStages; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(cx_smem_u32(&full_bar[s])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  auto is_valid = [&](int64_t t) { return t != ignore_index && t >= 0 && t < vocab; };
  // producer state (thread 0 only): the next row to look at, how many valid rows have been requested
  // (the target of `load_row` is fetched one call ahead: its global-memory latency never sits in the row loop)
  int64_t load_row = blockIdx.x;
  int64_t load_tgt = (tid == 0 && load_row < rows) ? targets[load_row] : ignore_index;
  uint32_t loads_issued = 0;
  auto issue_next_load = [&]() {
    while (load_row < rows && !is_valid(load_tgt)) {
      load_row += G;
      if (load_row < rows) load_tgt = targets[load_row];
    }
    if (load_row < rows) {
      const uint32_t st = loads_issued % kCStages;
      const uint32_t bar = cx_smem_u32(&full_bar[st]);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(row_bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem0 + st * kCStageBytes), "l"(logits + load_row * ld), "r"(row_bytes), "r"(bar) : "memory");
      ++loads_issued;
      load_row += G;
      if (load_row < rows) load_tgt = targets[load_row];
    }
  };
  if (tid == 0) { issue_next_load(); issue_next_load(); }

  // row-invariant: how many leading logits of each of this thread's vectors are real (inside the row AND inside the
  // vocabulary); 8 everywhere for all but the last few warps
  int nkeep[kCU];
  bool partial = false;
#pragma unroll
  for (int u = 0; u < kCU; ++u) {
    const int v = u * kCThreads + tid;
    nkeep[u] = v < nvec ? min(8, max(0, vocab - (v << 3))) : 0;
    partial = partial || nkeep[u] < 8;
  }
  uint32_t k = 0;   // valid rows processed so far (stage = k % 3)
  int64_t tgt_next = blockIdx.x < rows ? targets[blockIdx.x] : ignore_index;
  for (int64_t r = blockIdx.x; r < rows; r += G) {
    const int64_t tgt = tgt_next;
    if (r + G < rows) tgt_next = targets[r + G];   // used one row later
    if (!is_valid(tgt)) {   // ignored (or out-of-range) target: zero gradient, no loss; nothing was loaded
      const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
      __nv_bfloat16* z = logits + r * ld;
      for (int v = tid; v < nvec; v += kCThreads) st_stream16(z + ((int64_t)v << 3), zero);
      if (tid == 0) row_loss[r] = 0.f;
      continue;
    }
    const uint32_t st = k % kCStages;
    const uint32_t sbase = smem0 + st * kCStageBytes;
    cx_mbar_wait(cx_smem_u32(&full_bar[st]), (k / kCStages) & 1u);
    const int tcol = (int)tgt;
    float zt = 0.f;
    if (tid == 0)   // read before anyone overwrites the stage
      zt = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(cx_smem + st * kCStageBytes)[tcol]) * kLog2e;

    // ---- the thread's 32 logits: packed-bf16 maximum first (one HMNMX2 per two logits).  No per-vector guards: a
    // vector index past the row still lies inside the 64 KB stage (stale bytes, masked below), and which of a thread's
    // vectors reach past the vocabulary is the same for every row (`partial`, `nkeep`) ----
    uint4 q[kCU];
#pragma unroll
    for (int u = 0; u < kCU; ++u)
      asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q[u].x), "=r"(q[u].y), "=r"(q[u].z), "=r"(q[u].w)
                   : "r"(sbase + (uint32_t)(u * kCThreads + tid) * 16u) : "memory");
    if (partial) {   // a few warps at most: padding and out-of-row logits -> -inf
#pragma unroll
      for (int u = 0; u < kCU; ++u) {
        uint32_t* w = reinterpret_cast<uint32_t*>(&q[u]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (2 * j >= nkeep[u]) w[j] = (w[j] & 0xffff0000u) | 0xff80u;
          if (2 * j + 1 >= nkeep[u]) w[j] = (w[j] & 0x0000ffffu) | 0xff800000u;
        }
      }
    }
    uint32_t mx2 = bf16x2_max(bf16x2_max(q[0].x, q[0].y), bf16x2_max(q[0].z, q[0].w));
#pragma unroll
    for (int u = 1; u < kCU; ++u)
      mx2 = bf16x2_max(bf16x2_max(mx2, q[u].x), bf16x2_max(q[u].y, bf16x2_max(q[u].z, q[u].w)));
    // ---- row maximum BEFORE any exponential (max-merges cost FMNMX + SHFL; merging (max, sum) pairs instead would
    // cost two MUFU ops per thread and round, as many again as the exponentials themselves) ----
    float m = fmaxf(bf16lo(mx2), bf16hi(mx2));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((tid & 31) == 0) red_m[tid >> 5] = m;
    __syncthreads();
    m = red_m[tid & 31];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    m *= kLog2e;                                      // row maximum, log2 domain (finite: the target is a real column)
    // ---- one exponential per logit: FFMA2 + 2 MUFU + FADD2 per pair ----
    f32x2_t e2[kCU][4];
    const f32x2_t l2e2 = pk2(kLog2e, kLog2e), nm2 = pk2(-m, -m);
    f32x2_t s2a = pk2(0.f, 0.f), s2b = pk2(0.f, 0.f);
#pragma unroll
    for (int u = 0; u < kCU; ++u) {
      const uint32_t* w = reinterpret_cast<const uint32_t*>(&q[u]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float lo, hi;
        upk2(fma2(bf16x2_to_f32x2(w[j]), l2e2, nm2), lo, hi);
        e2[u][j] = pk2(ex2_approx(lo), ex2_approx(hi));   // padding: exp2(-inf) = 0
        if (j & 1) s2b = add2(s2b, e2[u][j]); else s2a = add2(s2a, e2[u][j]);
      }
    }
    float s_lo, s_hi;
    upk2(add2(s2a, s2b), s_lo, s_hi);
    float ssum = s_lo + s_hi;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, o);
    if ((tid & 31) == 0) red_s[tid >> 5] = ssum;
    __syncthreads();
    ssum = red_s[tid & 31];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, o);
    // ---- gradient: e * scale / sum; bf16 back into the same bytes (the one-hot is applied by thread 0 below) ----
    const float fac = __fdividef(scale, ssum);
    const f32x2_t fac2 = pk2(fac, fac);
#pragma unroll
    for (int u = 0; u < kCU; ++u) {   // unguarded: bytes past the row are never copied out
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) w[j] = f32x2_to_bf16x2(mul2(e2[u][j], fac2));   // padding columns carry e = 0
      asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(sbase + (uint32_t)(u * kCThreads + tid) * 16u), "r"(w[0]),
                   "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the bulk store
    __syncthreads();
    if (tid == 0) {
      // the one-hot: softmax(target) * scale - scale in one rounding, over the plain product stored for that column
      const float gt = ex2_approx(zt - m) * fac - scale;
      reinterpret_cast<__nv_bfloat16*>(cx_smem + st * kCStageBytes)[tcol] = __float2bfloat16_rn(gt);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                   ::"l"(logits + r * ld), "r"(sbase), "r"(row_bytes) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      // the stage the next load goes into was stored from one row ago: that store must have read it.  (Issuing the load
      // earlier, in the middle of the row, measured 7 % slower: a 64 KB bulk store needs more than a max + exp phase to
      // drain its stage, and the wait then sits on the row's critical path.)
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      issue_next_load();
      row_loss[r] = (m + log2f(ssum) - zt) * kLn2;
    }
    ++k;
  }
  if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // every store of this CTA has landed

  // ---- the last CTA to finish folds the per-row losses (fixed order: bit-reproducible) ----
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    s_last = (atomicAdd(ws + 1, 1u) == (unsigned int)G - 1u);
  }
  __syncthreads();
  if (s_last) {
    __threadfence();
    float t = 0.f;
    for (int64_t i = tid; i < rows; i += kCThreads) t += __ldcg(row_loss + i);
    t = warp_sum(t);
    if ((tid & 31) == 0) fr[tid >> 5] = t;
    __syncthreads();
    if (tid == 0) {
      float sum = 0.f;
      for (int w = 0; w < kCThreads / 32; ++w) sum += fr[w];
      loss_out[0] = sum / (float)(n_valid > 0 ? n_valid : 1);
    }
  }
}

__global__ void shift_targets_kernel(const int64_t* __restrict__ labels, int64_t* __restrict__ targets,
                                     int64_t total, int seq, int64_t ignore_index) {
  grid_dep_sync();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int t = (int)(i % seq);
  targets[i] = (t == seq - 1) ? ignore_index : labels[i + 1];
}

}  // namespace

namespace {
// Both entry points launch through here.  ws: the scratch words described at the top of the file.
int launch_xent(void* logits, int64_t rows, int vocab, int64_t ld, const int64_t* targets, int64_t ignore_index,
                float grad_scale, float* loss_out, unsigned int* ws, cudaStream_t st) {
  launch_k(count_valid_kernel, 1, 256, 0, st, targets, rows, ignore_index, vocab, ws);
  B200FT_LAUNCH_CHECK();
  const int64_t nvec = ld >> 3;
  static const bool legacy = getenv("B200FT_XENT_LEGACY") != nullptr;   // A/B measurements: two-sweep kernel only
  // Measured (tools/hbm_probe.py, B200, GB/s of algorithmic bytes): V = 32000: staged 5474, two-sweep 4882.  Rows
  // longer than one CTA's registers hold were also tried as clusters of 2-8 CTAs exchanging (max, sum) through
  // distributed shared memory: V = 50257 3600, V = 128256 4350-4500 — behind the two-sweep kernel (5089, 4740-4860
  // with its rows in flight capped to what L2 holds), because the per-row chain (max, exp at the MUFU rate, exchange,
  // store drain) stays while the bytes per CTA shrink; they are not built.
  if (!legacy && nvec <= kCChunkVecs) {
    static thread_local int attr_dev = -1;
    int dev = -1;
    B200FT_CUDA(cudaGetDevice(&dev));
    if (dev != attr_dev) {
      B200FT_CUDA(cudaFuncSetAttribute(softmax_xent_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kCSmemBytes));
      attr_dev = dev;
    }
    const unsigned grid = (unsigned)(rows < B200FT_NUM_SMS ? rows : B200FT_NUM_SMS);   // persistent, one CTA per SM
    launch_k(softmax_xent_staged_kernel, grid, kCThreads, kCSmemBytes, st, (__nv_bfloat16*)logits, rows, vocab, ld, targets,
                                                                     ignore_index, grad_scale, loss_out, ws);
    B200FT_LAUNCH_CHECK();
    return 0;
  }
  // Two-sweep kernel.  Its second sweep must find the row in L2 (126 MB): rows in flight = CTAs per SM x 148.  Rows up to
  // ~105 KB run 8 CTAs of 256 threads per SM; longer rows run 512-thread CTAs whose number per SM is capped, by a dynamic
  // shared-memory request they do not use, so that the rows in flight stay below ~120 MB.
  // Measured at V = 128256 (256 KB rows), GB/s: 512 threads x 1 / 2 / 3 CTAs per SM 3500-3630 / 4480-4630 / 4250-4860 (114 MB
  // in flight: fits L2 on some boxes and runs, not on others); 1024 threads x 2 4200-4320; the uncapped 303 MB in
  // flight (third sweep from HBM) 4270-4370.  V = 50257: 256 threads x 4 per SM 5150, x 8 5250.  Rule: rows in flight
  // <= 90 MB.
  const int64_t row_bytes = ld * 2;
  int64_t cap = (90ll << 20) / (row_bytes * B200FT_NUM_SMS);
  static const int force_cap = getenv("B200FT_XENT_CAP") ? atoi(getenv("B200FT_XENT_CAP")) : 0;   // experiments
  if (force_cap > 0) cap = force_cap;
  if (row_bytes * 8 * B200FT_NUM_SMS <= (126ll << 20) || force_cap >= 8) {
    static thread_local int attr_dev3 = -1;
    int dev3 = -1;
    B200FT_CUDA(cudaGetDevice(&dev3));
    if (dev3 != attr_dev3) {
      B200FT_CUDA(cudaFuncSetAttribute(softmax_xent_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
      attr_dev3 = dev3;
    }
    const int ctas256 = 8;   // measured at V = 50257 (119 MB in flight): 8 per SM 5250 GB/s, 6 4700, 4 5150
    launch_k(softmax_xent_kernel<256>, (unsigned)rows, 256, 220 * 1024 / ctas256 - 2048, st,
        (__nv_bfloat16*)logits, rows, vocab, ld, targets, ignore_index, grad_scale, loss_out, ws);
  } else {
    if (cap < 1) cap = 1;
    if (cap > 3) cap = 3;
    const int pad = (int)(220 * 1024 / cap) - 2048;
    static const int force_threads = getenv("B200FT_XENT_THREADS") ? atoi(getenv("B200FT_XENT_THREADS")) : 0;   // experiments
    const int threads = force_threads ? force_threads : 512;   // 1024-thread CTAs measured 7 % slower at the same bytes in flight
    static thread_local int attr_dev2 = -1;
    int dev = -1;
    B200FT_CUDA(cudaGetDevice(&dev));
    if (dev != attr_dev2) {
      B200FT_CUDA(cudaFuncSetAttribute(softmax_xent_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
      B200FT_CUDA(cudaFuncSetAttribute(softmax_xent_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
      attr_dev2 = dev;
    }
    if (threads == 1024)
      launch_k(softmax_xent_kernel<1024>, (unsigned)rows, 1024, pad, st, (__nv_bfloat16*)logits, rows, vocab, ld, targets,
                                                                  ignore_index, grad_scale, loss_out, ws);
    else
      launch_k(softmax_xent_kernel<512>, (unsigned)rows, 512, pad, st, (__nv_bfloat16*)logits, rows, vocab, ld, targets,
                                                                ignore_index, grad_scale, loss_out, ws);
  }
  B200FT_LAUNCH_CHECK();
  return 0;
}
}  // namespace

extern "C" size_t b200ft_softmax_xent_workspace_bytes(int64_t rows) {
  return (size_t)(rows + 16) * sizeof(float);
}

extern "C" int b200ft_softmax_xent_fwd_bwd(void* logits, int64_t rows, int vocab, int64_t ld,
                                           const int64_t* targets, int64_t ignore_index,
                                           float grad_scale, float* loss_out, void* workspace,
                                           size_t workspace_bytes, void* stream) {
  B200FT_CHECK_ARG(logits && targets && loss_out && workspace, "softmax_xent: null buffer");
  B200FT_CHECK_ARG(rows > 0 && rows < (1ll << 31) && vocab > 0 && ld >= vocab && ld % 8 == 0,
                   "softmax_xent: rows=%lld vocab=%d ld=%lld (ld must be a multiple of 8 and >= vocab)",
                   (long long)rows, vocab, (long long)ld);
  B200FT_CHECK_ARG((uintptr_t)logits % 16 == 0, "softmax_xent: logits must be 16-byte aligned");
  if (workspace_bytes < b200ft_softmax_xent_workspace_bytes(rows)) {
    b200ft_set_error("softmax_xent: workspace %zu < %zu", workspace_bytes,
                     b200ft_softmax_xent_workspace_bytes(rows));
    return B200FT_ERR_WORKSPACE;
  }
  return launch_xent(logits, rows, vocab, ld, targets, ignore_index, grad_scale, loss_out, (unsigned int*)workspace,
                     (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant.  Some callers of this library are not in a position to hand over a
// workspace: a host that has given every byte of device memory to its own pool and may not
// call cudaMalloc again, or a plug-in running inside another framework's allocator where a
// fresh allocation would fragment or stall it.  For them the scratch of the cross-entropy
// kernel (valid-target count, completion ticket, one loss per row) lives in this file-scope
// device array instead.  It costs (kNoAllocMaxRows + 16) * 4 bytes on every GPU that loads the
// library, until the process exits.
//
// The array's name is only a host-side handle; the kernels need its device address, which
// cudaGetSymbolAddress() yields for the device that is current at the time.  That look-up is
// done once per device, on the first call made with that device current.
// ---------------------------------------------------------------------------------------
namespace {
constexpr int64_t kNoAllocMaxRows = 65536;
static __device__ __align__(256) unsigned int g_xent_static_scratch[kNoAllocMaxRows + 16];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
unsigned int* g_scratch_addr[kMaxDevices] = {nullptr};

// Device address of the static scratch on the CURRENT device (resolved once per device).
int static_scratch_for_current_device(unsigned int** out) {
  int dev = -1;
  B200FT_CUDA(cudaGetDevice(&dev));
  B200FT_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "softmax_xent_noalloc: device ordinal %d unsupported", dev);
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (g_scratch_addr[dev] == nullptr) {
    unsigned int* ptr = nullptr;
    B200FT_CUDA(cudaGetSymbolAddress((void**)&ptr, g_xent_static_scratch));
    g_scratch_addr[dev] = ptr;
  }
  *out = g_scratch_addr[dev];
  return 0;
}
}  // namespace

extern "C" int64_t b200ft_softmax_xent_noalloc_max_rows(void) { return kNoAllocMaxRows; }

extern "C" int b200ft_softmax_xent_fwd_bwd_noalloc(void* logits, int64_t rows, int vocab, int64_t ld,
                                                   const int64_t* targets, int64_t ignore_index,
                                                   float grad_scale, float* loss_out, void* stream) {
  B200FT_CHECK_ARG(logits && targets && loss_out, "softmax_xent_noalloc: null buffer");
  B200FT_CHECK_ARG(rows > 0 && vocab > 0 && ld >= vocab && ld % 8 == 0,
                   "softmax_xent_noalloc: rows=%lld vocab=%d ld=%lld (ld must be a multiple of 8 and >= vocab)",
                   (long long)rows, vocab, (long long)ld);
  if (rows > kNoAllocMaxRows) {
    b200ft_set_error("softmax_xent_noalloc: rows=%lld exceeds the built-in scratch (max %lld rows); split the "
                     "batch or use b200ft_softmax_xent_fwd_bwd with a caller workspace",
                     (long long)rows, (long long)kNoAllocMaxRows);
    return B200FT_ERR_WORKSPACE;
  }
  B200FT_CHECK_ARG((uintptr_t)logits % 16 == 0, "softmax_xent_noalloc: logits must be 16-byte aligned");
  unsigned int* ws = nullptr;
  int rc = static_scratch_for_current_device(&ws);
  if (rc) return rc;
  // count_valid_kernel rewrites the counts and zeroes the ticket; every row's loss slot is
  // rewritten by the CTA that owns the row, so nothing of the previous call's contents is read.
  return launch_xent(logits, rows, vocab, ld, targets, ignore_index, grad_scale, loss_out, ws, (cudaStream_t)stream);
}

extern "C" int b200ft_shift_targets(const int64_t* labels, int64_t* targets, int64_t batch, int seq,
                                    int64_t ignore_index, void* stream) {
  B200FT_CHECK_ARG(labels && targets && batch > 0 && seq > 0, "shift_targets: bad arguments");
  int64_t total = batch * seq;
  launch_k(shift_targets_kernel, (unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream,
      labels, targets, total, seq, ignore_index);
  B200FT_LAUNCH_CHECK();
  return 0;
}
