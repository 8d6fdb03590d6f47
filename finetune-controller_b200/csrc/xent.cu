// Fused softmax-cross-entropy forward + backward over bf16 logits, in place.
//
// Main kernel (softmax_xent_cluster_kernel): a row never leaves the chip between its read and its write.
// A thread-block CLUSTER of 1, 2, 4 or 8 CTAs owns a row; each CTA's share (at most 32768 logits = 64 KB) is
// brought into shared memory by one bulk-copy (TMA) per row, three rows deep, so HBM reads for the next two
// rows are in flight while the current one is processed.  Each of the 512 threads takes its 64 logits into
// registers ONCE: e = exp2(z*log2e - m_thread) against the thread's own maximum (one MUFU op per element; a
// re-reading kernel needs two, and MUFU at 16/clk/SM is what bounded the previous version at 62 % busy),
// (max, sum) pairs are merged across the warp, the CTA and — through distributed shared memory, one cluster
// barrier per row — the cluster; the thread then scales its registers by 2^(m_thread - M) * grad_scale /
// (n_valid * sum), subtracts the one-hot, writes bf16 back into the same shared-memory bytes, and one bulk
// store per CTA sends them to HBM.  DRAM traffic is exactly one read + one write of the row whatever the
// vocabulary (the previous kernel re-read the row from L2, which at V = 128256 with 8 CTAs x 148 SMs x 256 KB
// in flight no longer held it: ~3 sweeps instead of 2).  Algorithmic bytes: 2*rows*ld*2 + 12*rows.
//
// Fallback (softmax_xent_kernel, rows longer than 8 x 32768 logits, or B200FT_XENT_LEGACY=1): one CTA per row,
// two sweeps with the second served by L2.
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

constexpr int kThreads = 256;
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

// the same merge on the MUFU approximation (inputs are maxima and sums of exp2: ~2^-22 relative is ample)
__device__ __forceinline__ MS ms_merge_fast(MS a, MS b) {
  const float M = fmaxf(a.m, b.m);
  MS r;
  r.m = M;
  const float ea = (a.m == -INFINITY) ? 0.f : ex2_approx(a.m - M);
  const float eb = (b.m == -INFINITY) ? 0.f : ex2_approx(b.m - M);
  r.s = a.s * ea + b.s * eb;
  return r;
}

__global__ void __launch_bounds__(kThreads)
softmax_xent_kernel(__nv_bfloat16* __restrict__ logits, int64_t rows, int vocab, int64_t ld,
                    const int64_t* __restrict__ targets, int64_t ignore_index, float grad_scale,
                    float* __restrict__ loss_out, unsigned int* __restrict__ ws) {
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
    MS acc{-INFINITY, 0.f};
    for (int i0 = threadIdx.x; i0 < nvec; i0 += kThreads * kUnroll) {
      uint4 q[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int i = i0 + u * kThreads;
        if (i < nvec) q[u] = ld_cg16(z + ((int64_t)i << 3));
      }
      float f[kUnroll][8];
      float vmax = -INFINITY;
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int i = i0 + u * kThreads;
        const int c0 = i << 3;
        if (i < nvec) {
          unpack8(q[u], f[u]);
          if (c0 + 8 <= vocab) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { f[u][j] *= kLog2e; vmax = fmaxf(vmax, f[u][j]); }
          } else {  // the vector straddling (or beyond) the end of the vocabulary
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              f[u][j] = (c0 + j < vocab) ? f[u][j] * kLog2e : -INFINITY;
              vmax = fmaxf(vmax, f[u][j]);
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) f[u][j] = -INFINITY;
        }
      }
      if (vmax > acc.m) {  // rare after the first few iterations
        acc.s *= ex2_approx(acc.m - vmax);  // ex2(-inf) = 0 on the first hit
        acc.m = vmax;
      }
      if (acc.m != -INFINITY) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc.s += ex2_approx(f[u][j] - acc.m);
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
          float f[8];
          unpack8(q[u], f);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = ex2_approx(f[j] * kLog2e - lse2) * scale;
          if (tcol >= c0 && tcol < c0 + 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) if (c0 + j == tcol) f[j] -= scale;
          }
          if (c0 + 8 > vocab) {
#pragma unroll
            for (int j = 0; j < 8; ++j) if (c0 + j >= vocab) f[j] = 0.f;
          }
          st_stream16(z + ((int64_t)i << 3), pack8(f));
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
// Cluster kernel: rows staged in shared memory by bulk copies, one pass through registers.
// ---------------------------------------------------------------------------------------
constexpr int kCThreads = 1024;
constexpr int kCU = 4;                           // 16-byte vectors (8 logits each) per thread
constexpr int kCChunkVecs = kCThreads * kCU;     // 4096 vectors = 32768 logits = 64 KB per CTA and row
constexpr int kCStages = 3;
constexpr int kCStageBytes = kCChunkVecs * 16;
constexpr int kCSmemBytes = kCStages * kCStageBytes;   // 192 KB dynamic; barriers and exchange slots are static
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
// wait for a phase completed by another CTA's st.async: acquire at cluster scope
__device__ __forceinline__ void cx_mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = 0;
  while (true) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > kCWaitLimit) __trap();
  }
}
__device__ __forceinline__ uint32_t cx_cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cx_cluster_size() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cx_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float cx_ld_remote_f32(uint32_t local_addr, uint32_t rank) {
  uint32_t ra;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_addr), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
  return v;
}

__global__ void __launch_bounds__(kCThreads, 1)
softmax_xent_cluster_kernel(__nv_bfloat16* __restrict__ logits, int64_t rows, int vocab, int64_t ld,
                            const int64_t* __restrict__ targets, int64_t ignore_index, float grad_scale,
                            float* __restrict__ loss_out, unsigned int* __restrict__ ws, int chunk_vecs) {
  extern __shared__ __align__(128) uint8_t cx_smem[];
  __shared__ __align__(8) unsigned long long full_bar[kCStages];
  __shared__ __align__(16) float xch[2][8][4];   // [row parity][sending CTA]: (max, sum, target logit, -)
  __shared__ __align__(8) unsigned long long xch_bar[2];
  __shared__ float red_m[kCThreads / 32], red_s[kCThreads / 32];
  __shared__ bool s_last;
  __shared__ float fr[kCThreads / 32];

  const int tid = threadIdx.x;
  const uint32_t crank = cx_cluster_rank(), csize = cx_cluster_size();
  const int64_t cid = blockIdx.x / csize, G = gridDim.x / csize;
  const int nvec = (int)(ld >> 3);
  const int v0 = (int)crank * chunk_vecs;                                   // first vector of this CTA's share
  const int my_vecs = max(0, min(chunk_vecs, nvec - v0));
  const uint32_t my_bytes = (uint32_t)my_vecs * 16u;
  const int col_lo = v0 << 3, col_hi = (v0 + my_vecs) << 3;
  const uint32_t smem0 = cx_smem_u32(cx_smem);
  float* row_loss = reinterpret_cast<float*>(ws + 16);
  const unsigned int n_valid = ws[0];
  const float scale = grad_scale / (float)(n_valid > 0 ? n_valid : 1);

  if (tid == 0) {
    for (int s = 0; s < kCStages; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(cx_smem_u32(&full_bar[s])) : "memory");
    for (int s = 0; s < 2; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(cx_smem_u32(&xch_bar[s])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (csize > 1) cx_cluster_sync();   // once: nobody sends into a neighbour's exchange barrier before it exists

  auto is_valid = [&](int64_t t) { return t != ignore_index && t >= 0 && t < vocab; };
  // producer state (thread 0 only): the next row to look at, how many valid rows have been requested
  int64_t load_row = cid;
  uint32_t loads_issued = 0;
  auto issue_next_load = [&]() {
    while (load_row < rows && !is_valid(targets[load_row])) load_row += G;
    if (load_row < rows) {
      if (my_bytes) {
        const uint32_t st = loads_issued % kCStages;
        const uint32_t bar = cx_smem_u32(&full_bar[st]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(my_bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem0 + st * kCStageBytes), "l"(logits + load_row * ld + col_lo), "r"(my_bytes), "r"(bar)
                     : "memory");
      }
      ++loads_issued;
      load_row += G;
    }
  };
  if (tid == 0) { issue_next_load(); issue_next_load(); }

  uint32_t k = 0;   // valid rows processed so far (stage = k % 3, exchange slot = k & 1)
  for (int64_t r = cid; r < rows; r += G) {
    const int64_t tgt = targets[r];
    if (!is_valid(tgt)) {   // ignored (or out-of-range) target: zero gradient, no loss; nothing was loaded
      const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
      __nv_bfloat16* z = logits + r * ld + col_lo;
      for (int v = tid; v < my_vecs; v += kCThreads) st_stream16(z + ((int64_t)v << 3), zero);
      if (crank == 0 && tid == 0) row_loss[r] = 0.f;
      continue;
    }
    const uint32_t st = k % kCStages;
    const uint32_t sbase = smem0 + st * kCStageBytes;
    if (my_bytes) cx_mbar_wait(cx_smem_u32(&full_bar[st]), (k / kCStages) & 1u);
    const int tcol = (int)tgt;
    float zt = -INFINITY;
    if (tid == 0 && tcol >= col_lo && tcol < col_hi)   // read before anyone overwrites the stage
      zt = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(cx_smem + st * kCStageBytes)[tcol - col_lo]) * kLog2e;

    // ---- the thread's 32 logits: packed-bf16 maximum first (one HMNMX2 per two logits) ----
    uint4 q[kCU];
    uint32_t mx2 = 0xff80ff80u;   // (-inf, -inf) in bf16
#pragma unroll
    for (int u = 0; u < kCU; ++u) {
      const int v = u * kCThreads + tid;
      q[u] = make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);
      if (v < my_vecs) {
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q[u].x), "=r"(q[u].y), "=r"(q[u].z), "=r"(q[u].w)
                     : "r"(sbase + (uint32_t)v * 16u) : "memory");
        const int c0 = (v0 + v) << 3;
        if (c0 + 8 > vocab) {   // the vector straddling (or beyond) the end of the vocabulary: padding -> -inf
          uint32_t* w = reinterpret_cast<uint32_t*>(&q[u]);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (c0 + 2 * j >= vocab) w[j] = (w[j] & 0xffff0000u) | 0xff80u;
            if (c0 + 2 * j + 1 >= vocab) w[j] = (w[j] & 0x0000ffffu) | 0xff800000u;
          }
        }
        mx2 = bf16x2_max(bf16x2_max(mx2, q[u].x), bf16x2_max(q[u].y, bf16x2_max(q[u].z, q[u].w)));
      }
    }
    // ---- CTA maximum BEFORE any exponential (max-merges cost FMNMX + SHFL; merging (max, sum) pairs instead would
    // cost two MUFU ops per thread and round, as many again as the exponentials themselves) ----
    float m = fmaxf(bf16lo(mx2), bf16hi(mx2));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((tid & 31) == 0) red_m[tid >> 5] = m;
    __syncthreads();
    m = red_m[tid & 31];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    m *= kLog2e;                                      // CTA maximum, log2 domain
    // ---- one exponential per logit: FFMA2 + 2 MUFU + FADD2 per pair ----
    f32x2_t e2[kCU][4];
    const float nm = (m == -INFINITY) ? 0.f : -m;    // all-padding share: exp2(-inf + 0) = 0
    const f32x2_t l2e2 = pk2(kLog2e, kLog2e), nm2 = pk2(nm, nm);
    f32x2_t s2a = pk2(0.f, 0.f), s2b = pk2(0.f, 0.f);
#pragma unroll
    for (int u = 0; u < kCU; ++u) {
      const uint32_t* w = reinterpret_cast<const uint32_t*>(&q[u]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float lo, hi;
        upk2(fma2(bf16x2_to_f32x2(w[j]), l2e2, nm2), lo, hi);
        e2[u][j] = pk2(ex2_approx(lo), ex2_approx(hi));
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
    MS acc{m, ssum};                                  // this CTA's share of the row
    MS tot = acc;
    float zt_all = zt;
    if (csize > 1) {
      // Exchange without a cluster-wide barrier: thread 0 PUSHES this CTA's triple into slot [k & 1][crank] of every
      // CTA of the cluster (itself included) with st.async, which signals the receiver's mbarrier by byte count; every
      // thread then waits on its own CTA's barrier and reads local shared memory.  Slot k & 1 is reused two rows
      // later: a sender gets there only after it has received this CTA's message of row k+1, which this CTA sends
      // after all of its threads have read the slots of row k.
      const uint32_t xb = cx_smem_u32(&xch_bar[k & 1]);
      if (tid == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(xb), "r"(csize * 16u) : "memory");
        const uint32_t slot = cx_smem_u32(&xch[k & 1][crank][0]);
        for (uint32_t c = 0; c < csize; ++c) {
          uint32_t rslot, rbar;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rslot) : "r"(slot), "r"(c));
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rbar) : "r"(xb), "r"(c));
          asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];"
                       ::"r"(rslot), "f"(acc.m), "f"(acc.s), "f"(zt), "f"(0.f), "r"(rbar) : "memory");
        }
      }
      cx_mbar_wait_cluster(xb, (k >> 1) & 1u);
      const uint32_t c = (uint32_t)(tid & 7);            // lane c takes CTA c's triple
      MS o{-INFINITY, 0.f};
      float ztc = -INFINITY;
      if (c < csize) { o.m = xch[k & 1][c][0]; o.s = xch[k & 1][c][1]; ztc = xch[k & 1][c][2]; }
#pragma unroll
      for (int sh = 4; sh > 0; sh >>= 1) {
        MS other;
        other.m = __shfl_xor_sync(0xffffffffu, o.m, sh);
        other.s = __shfl_xor_sync(0xffffffffu, o.s, sh);
        ztc = fmaxf(ztc, __shfl_xor_sync(0xffffffffu, ztc, sh));
        o = ms_merge_fast(o, other);
      }
      tot = o; zt_all = ztc;
    } else {
      // zt_all = zt: only thread 0 holds the target logit, and only thread 0 uses it (the row loss)
    }
    // ---- gradient: e * 2^(m_thread - M) * scale / sum, minus the one-hot; bf16 back into the same bytes ----
    const float fac = (m == -INFINITY) ? 0.f : ex2_approx(m - tot.m) * __fdividef(scale, tot.s);
    const f32x2_t fac2 = pk2(fac, fac);
    const int tvec = tcol >> 3, tj = tcol & 7;
#pragma unroll
    for (int u = 0; u < kCU; ++u) {
      const int v = u * kCThreads + tid;
      if (v < my_vecs) {
        uint32_t w[4];
        if (v0 + v == tvec) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float lo, hi;
            upk2(mul2(e2[u][j], fac2), lo, hi);
            if (2 * j == tj) lo -= scale;
            if (2 * j + 1 == tj) hi -= scale;
            w[j] = pack_bf16x2(lo, hi);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) w[j] = f32x2_to_bf16x2(mul2(e2[u][j], fac2));   // padding columns carry e = 0
        }
        asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(sbase + (uint32_t)v * 16u), "r"(w[0]), "r"(w[1]),
                     "r"(w[2]), "r"(w[3]) : "memory");
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the bulk store
    __syncthreads();
    if (tid == 0) {
      if (my_bytes) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     ::"l"(logits + r * ld + col_lo), "r"(sbase), "r"(my_bytes) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        // the stage the next load goes into was stored from one row ago: that store must have read it
        asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      }
      issue_next_load();
      if (crank == 0) row_loss[r] = (tot.m + log2f(tot.s) - zt_all) * kLn2;
    }
    ++k;
  }
  if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // every store of this CTA has landed
  // (every message addressed to this CTA has been waited for above, so it may leave without a closing cluster barrier)

  // ---- the last cluster to finish folds the per-row losses (fixed order: bit-reproducible) ----
  if (crank != 0) return;
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
  count_valid_kernel<<<1, 256, 0, st>>>(targets, rows, ignore_index, vocab, ws);
  B200FT_LAUNCH_CHECK();
  const int64_t nvec = ld >> 3;
  int csize = 1;
  while (csize < 8 && (nvec + csize - 1) / csize > kCChunkVecs) csize <<= 1;
  static const bool legacy = getenv("B200FT_XENT_LEGACY") != nullptr;
  if (!legacy && (nvec + csize - 1) / csize <= kCChunkVecs) {
    static thread_local int attr_dev = -1;
    static thread_local int max_clusters[4] = {0, 0, 0, 0};   // co-resident clusters for sizes 1, 2, 4, 8
    int dev = -1;
    B200FT_CUDA(cudaGetDevice(&dev));
    if (dev != attr_dev) {
      B200FT_CUDA(cudaFuncSetAttribute(softmax_xent_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kCSmemBytes));
      for (int i = 0; i < 4; ++i) max_clusters[i] = 0;
      attr_dev = dev;
    }
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)csize; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.blockDim = dim3(kCThreads); cfg.dynamicSmemBytes = kCSmemBytes; cfg.stream = st;
    cfg.attrs = attr; cfg.numAttrs = 1;
    const int slot = csize == 1 ? 0 : csize == 2 ? 1 : csize == 4 ? 2 : 3;
    if (max_clusters[slot] == 0) {
      // persistent grid = what the chip holds at once (GPCs whose SM count is not a multiple of the cluster
      // size strand a few SMs; a grid larger than this would run a second, mostly empty wave)
      cfg.gridDim = dim3((unsigned)(B200FT_NUM_SMS / csize * csize));
      int n = 0;
      B200FT_CUDA(cudaOccupancyMaxActiveClusters(&n, softmax_xent_cluster_kernel, &cfg));
      B200FT_CHECK_ARG(n > 0, "softmax_xent: no cluster of %d CTAs fits on this device", csize);
      max_clusters[slot] = n;
    }
    const int64_t clusters = rows < max_clusters[slot] ? rows : max_clusters[slot];
    cfg.gridDim = dim3((unsigned)(clusters * csize));
    const int chunk_vecs = (int)((nvec + csize - 1) / csize);
    B200FT_CUDA(cudaLaunchKernelEx(&cfg, softmax_xent_cluster_kernel, (__nv_bfloat16*)logits, rows, vocab, ld, targets,
                                   ignore_index, grad_scale, loss_out, ws, chunk_vecs));
    return 0;
  }
  softmax_xent_kernel<<<(unsigned)rows, kThreads, 0, st>>>((__nv_bfloat16*)logits, rows, vocab, ld, targets,
                                                           ignore_index, grad_scale, loss_out, ws);
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
  shift_targets_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      labels, targets, total, seq, ignore_index);
  B200FT_LAUNCH_CHECK();
  return 0;
}
