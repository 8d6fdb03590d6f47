// Causal flash-attention BACKWARD on tcgen05 for d_head = 64, the counterpart of attention.cu: reads q, k, v
// from the fused [tokens, q|k|v] projection buffer, o / do from [tokens, heads*64] buffers and the forward's
// per-row log-sum-exp, and writes dq | dk | dv into a buffer with the projection's own column layout.
//
// Three launches:
//   delta     D[row] = sum_d do[row,d] * o[row,d]                              (one warp per token and head)
//   MODE 0    one CTA per (kv tile j, batch, kv head): K_j, V_j resident; streams (Q_i, dO_i) over the q heads
//             of the group and the q tiles i >= j; dV_j += P^T dO_i and dK_j += dS^T Q_i accumulate in TMEM
//   MODE 1    one CTA per (q tile i, batch, head): Q_i, dO_i resident; streams (K_j, V_j), j <= i;
//             dQ_i += dS K_j accumulates in TMEM
// Both modes recompute S = Q K^T and dP = dO V^T per tile pair on the tensor core (no atomics, no fp32 dQ
// buffer; 7 instead of 5 tile GEMMs).  Per pair: warps 4-7 (one thread per query row) read S and dP from
// TMEM, form P = exp2(S*scale*log2e - lse2) and dS = P * (dP - D) * scale, and write them as bf16 into
// 128B-swizzled shared tiles [q rows][kv columns] — the byte layout TMA would produce — which the second
// group of UMMAs consumes MN-major (P^T, dS^T: reduction over q) or K-major (dS: reduction over kv).
// Q, dO, K, V tiles are used in both roles too: K-major for S / dP, MN-major as the B operand of the
// gradient GEMMs.  256 threads: warp 0 TMA, warp 1 MMA issue, warp 2 TMEM allocation, warps 4-7 softmax.
#include <cuda.h>
#include <mutex>
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

constexpr int kDH = 64;
constexpr int kB = 128;                       // tile edge in queries and in keys
constexpr int kThreads = 256;
constexpr int kTile = kB * kDH * 2;           // 16 KB: one Q / dO / K / V tile
constexpr int kPTile = kB * kB * 2;           // 32 KB: P or dS, two 64-column chunks of 16 KB
constexpr int kOffF = 0;                      // resident pair
constexpr int kOffS = 2 * kTile;              // stream ring: 2 stages x (tile0, tile1)
constexpr int kOffP = kOffS + 4 * kTile;
constexpr int kOffDS = kOffP + kPTile;
constexpr int kOffBar = kOffDS + kPTile;      // 160 KB
constexpr int kSmemBytes = kOffBar + 256;
constexpr uint32_t kColS = 0, kColDP = 128, kColAcc0 = 256, kColAcc1 = 320;
constexpr long long kWaitLimit = 4000000000ll;

struct BwdParams {
  int batch, seq, n_head, n_kv_head;
  int q_off, k_off, v_off;
  int n_tiles;
  float scale, scale_log2;
  const float* lse2;     // [batch*n_head, seq]
  const float* delta;    // [batch*n_head, seq]
  __nv_bfloat16* dqkv; int64_t ld_d;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > kWaitLimit) __trap();  // deadlock guard
  }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// smem matrix descriptor (SW128): start>>4 | LBO>>4 << 16 || SBO>>4 | version 1 | swizzle 128B
__device__ __forceinline__ uint64_t sdesc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  const uint32_t lo = ((addr & 0x3FFFFu) >> 4) | (((lbo >> 4) & 0x3FFFu) << 16);
  const uint32_t hi = ((sbo >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
  return ((uint64_t)hi << 32) | lo;
}
// kind::f16 instruction descriptor: fp32 accumulate, bf16 x bf16, M = 128
__device__ __forceinline__ uint32_t idesc(int n, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
         ((uint32_t)(n >> 3) << 17) | (8u << 24);
}

// ------------------------------------------------------------------------------------------------
// D[bh, t] = sum_d do[t, h*64 + d] * o[t, h*64 + d]; one warp per (token, head)
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ o, int64_t ld_o, const __nv_bfloat16* __restrict__ dout,
                  int64_t ld_do, float* __restrict__ delta, int batch, int seq, int n_head) {
  const int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int64_t total = (int64_t)batch * seq * n_head;
  if (w >= total) return;
  const int lane = threadIdx.x & 31;
  const int h = (int)(w % n_head);
  const int64_t tok = w / n_head;           // b * seq + t
  const int b = (int)(tok / seq), t = (int)(tok % seq);
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(o + tok * ld_o + h * kDH + 2 * lane));
  const float2 g = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(dout + tok * ld_do + h * kDH + 2 * lane));
  const float s = warp_sum(a.x * g.x + a.y * g.y);
  if (lane == 0) delta[((int64_t)b * n_head + h) * seq + t] = s;
}

// ------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                const BwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  if (raw & 1023u) __trap();  // dynamic shared memory starts 1024-aligned (no static smem in this kernel)
  const uint32_t sF0 = raw + kOffF, sF1 = sF0 + kTile;
  auto sS0 = [&](int st) { return raw + kOffS + (uint32_t)st * 2u * kTile; };
  auto sS1 = [&](int st) { return raw + kOffS + (uint32_t)st * 2u * kTile + kTile; };
  const uint32_t sP = raw + kOffP, sDS = raw + kOffDS;
  const uint32_t bars = raw + kOffBar;
  const uint32_t f_full = bars;
  auto st_full = [&](int s) { return bars + 8u * (1 + s); };
  auto st_empty = [&](int s) { return bars + 8u * (3 + s); };
  const uint32_t sdp_full = bars + 8u * 5, sdp_free = bars + 8u * 6;
  const uint32_t pds_full = bars + 8u * 7, pds_free = bars + 8u * 8, acc_full = bars + 8u * 9;
  const uint32_t tmem_slot = bars + 8u * 10;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = p.n_head / p.n_kv_head;
  // work item of this CTA, heaviest first.  MODE 0: (kv tile jt ascending, batch, kv head);
  // MODE 1: (q tile it descending, batch, head)
  const int per_tile = p.batch * (MODE == 0 ? p.n_kv_head : p.n_head);
  const int tile_ord = (int)(blockIdx.x / per_tile);
  const int rest = (int)(blockIdx.x % per_tile);
  const int fixed_tile = MODE == 0 ? tile_ord : p.n_tiles - 1 - tile_ord;
  const int b = rest / (MODE == 0 ? p.n_kv_head : p.n_head);
  const int hx = rest % (MODE == 0 ? p.n_kv_head : p.n_head);   // kv head (MODE 0) or q head (MODE 1)
  const int hk = MODE == 0 ? hx : hx / G;
  const int per_head = MODE == 0 ? p.n_tiles - fixed_tile : fixed_tile + 1;  // streamed tiles per head
  const int n_iter = MODE == 0 ? G * per_head : per_head;

  if (warp == 1 && lane == 0) {
    mbar_init(f_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(st_full(s), 1); mbar_init(st_empty(s), 1); }
    mbar_init(sdp_full, 1); mbar_init(sdp_free, 4); mbar_init(pds_full, 4); mbar_init(pds_free, 1);
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_ptr;

  if (warp == 0) {
    // ================================ TMA producer ================================
    const bool issuer = elect_one();
    if (issuer) {
      mbar_expect_tx(f_full, 2 * kTile);
      if (MODE == 0) {
        tma_load_3d(sF0, &tm_qkv, p.k_off + hk * kDH, fixed_tile * kB, b, f_full);
        tma_load_3d(sF1, &tm_qkv, p.v_off + hk * kDH, fixed_tile * kB, b, f_full);
      } else {
        tma_load_3d(sF0, &tm_qkv, p.q_off + hx * kDH, fixed_tile * kB, b, f_full);
        tma_load_3d(sF1, &tm_do, hx * kDH, fixed_tile * kB, b, f_full);
      }
    }
    for (int t = 0; t < n_iter; ++t) {
      const int st = t & 1;
      const uint32_t ph = (uint32_t)(t >> 1) & 1u;
      mbar_wait(st_empty(st), ph ^ 1u);
      if (issuer) {
        mbar_expect_tx(st_full(st), 2 * kTile);
        if (MODE == 0) {
          const int hq = hk * G + t / per_head, it = fixed_tile + t % per_head;
          tma_load_3d(sS0(st), &tm_qkv, p.q_off + hq * kDH, it * kB, b, st_full(st));
          tma_load_3d(sS1(st), &tm_do, hq * kDH, it * kB, b, st_full(st));
        } else {
          tma_load_3d(sS0(st), &tm_qkv, p.k_off + hk * kDH, t * kB, b, st_full(st));
          tma_load_3d(sS1(st), &tm_qkv, p.v_off + hk * kDH, t * kB, b, st_full(st));
        }
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    const bool issuer = elect_one();
    const uint32_t id_s = idesc(kB, 0, 0);     // S, dP: 128 x 128 x 64, both operands K-major
    const uint32_t id_g0 = idesc(kDH, 1, 1);   // dV, dK: 128 (keys) x 64 x 128 (queries): A = P^T / dS^T, B = dO / Q, MN-major
    const uint32_t id_g1 = idesc(kDH, 0, 1);   // dQ: 128 (queries) x 64 x 128 (keys): A = dS K-major, B = K MN-major
    auto issue_sdp = [&](int st) {
      const uint32_t aq = MODE == 0 ? sS0(st) : sF0, bk = MODE == 0 ? sF0 : sS0(st);
      const uint32_t ado = MODE == 0 ? sS1(st) : sF1, bv = MODE == 0 ? sF1 : sS1(st);
      if (issuer) {
#pragma unroll
        for (int k = 0; k < kDH / 16; ++k)
          umma(tmem + kColS, sdesc(aq + 32 * k, 16, 1024), sdesc(bk + 32 * k, 16, 1024), id_s, k > 0);
#pragma unroll
        for (int k = 0; k < kDH / 16; ++k)
          umma(tmem + kColDP, sdesc(ado + 32 * k, 16, 1024), sdesc(bv + 32 * k, 16, 1024), id_s, k > 0);
        umma_commit(sdp_full);
      }
      __syncwarp();
    };
    auto issue_grads = [&](int st, bool first) {
      if (issuer) {
        if (MODE == 0) {
#pragma unroll
          for (int k = 0; k < kB / 16; ++k)   // reduction over 16 query rows per step
            umma(tmem + kColAcc0, sdesc(sP + 2048 * k, 16384, 1024), sdesc(sS1(st) + 2048 * k, 8192, 1024), id_g0,
                 !(first && k == 0));
#pragma unroll
          for (int k = 0; k < kB / 16; ++k)
            umma(tmem + kColAcc1, sdesc(sDS + 2048 * k, 16384, 1024), sdesc(sS0(st) + 2048 * k, 8192, 1024), id_g0,
                 !(first && k == 0));
        } else {
#pragma unroll
          for (int k = 0; k < kB / 16; ++k)   // reduction over 16 keys per step: dS chunk k/4, 32 B per step inside
            umma(tmem + kColAcc0, sdesc(sDS + 16384 * (k >> 2) + 32 * (k & 3), 16, 1024),
                 sdesc(sS0(st) + 2048 * k, 8192, 1024), id_g1, !(first && k == 0));
        }
        umma_commit(pds_free);      // P / dS tiles may be overwritten
        umma_commit(st_empty(st));  // and the streamed pair's slot refilled
      }
      __syncwarp();
    };
    mbar_wait(f_full, 0);
    mbar_wait(st_full(0), 0);
    tc_fence_after();
    issue_sdp(0);
    for (int t = 0; t < n_iter; ++t) {
      const int st = t & 1;
      const uint32_t par = (uint32_t)(t & 1);
      // the next pair's scores first (the softmax warps will be waiting for them), then this pair's gradients
      mbar_wait(sdp_free, par);           // S(t), dP(t) have been read out of TMEM
      if (t + 1 < n_iter) {
        mbar_wait(st_full(st ^ 1), (uint32_t)((t + 1) >> 1) & 1u);
        tc_fence_after();
        issue_sdp(st ^ 1);
      }
      mbar_wait(pds_full, par);           // P(t), dS(t) are in shared memory
      tc_fence_after();
      issue_grads(st, t == 0);
    }
    if (issuer) umma_commit(acc_full);
    __syncwarp();
  } else if (warp >= 4) {
    // ================================ softmax / dS, one thread per query row ================================
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t t_s = tmem + ((uint32_t)(quarter * 32) << 16) + kColS;
    const uint32_t t_dp = tmem + ((uint32_t)(quarter * 32) << 16) + kColDP;
    const uint32_t swz = (uint32_t)(row & 7);
    const uint32_t row_off = (uint32_t)row * 128u;
    float lse = INFINITY, dl = 0.f;
    if (MODE == 1) {
      const int qrow = fixed_tile * kB + row;
      if (qrow < p.seq) {
        const int64_t idx = ((int64_t)b * p.n_head + hx) * p.seq + qrow;
        lse = p.lse2[idx]; dl = p.delta[idx];
      }
    }
    for (int t = 0; t < n_iter; ++t) {
      const uint32_t par = (uint32_t)(t & 1);
      int qt, kt;
      if (MODE == 0) {
        const int hq = hk * G + t / per_head;
        qt = fixed_tile + t % per_head; kt = fixed_tile;
        const int qrow = qt * kB + row;
        lse = INFINITY; dl = 0.f;          // rows past the end of the sequence: P = dS = 0
        if (qrow < p.seq) {
          const int64_t idx = ((int64_t)b * p.n_head + hq) * p.seq + qrow;
          lse = p.lse2[idx]; dl = p.delta[idx];
        }
      } else {
        qt = fixed_tile; kt = t;
      }
      const bool diag = qt == kt;
      const int qrow = qt * kB + row;
      mbar_wait(sdp_full, par);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < kB / 32; ++c) {
        uint32_t rs[32], rd[32];
        tmem_ld32(t_s + 32 * c, rs);
        tmem_ld32(t_dp + 32 * c, rd);
        tmem_wait_ld();
        uint32_t wp[16], wd[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = ex2(fmaf(__uint_as_float(rs[i]), p.scale_log2, -lse));
          float p1 = ex2(fmaf(__uint_as_float(rs[i + 1]), p.scale_log2, -lse));
          if (diag) {
            if (kt * kB + 32 * c + i > qrow) p0 = 0.f;
            if (kt * kB + 32 * c + i + 1 > qrow) p1 = 0.f;
          }
          const float d0 = p0 * (__uint_as_float(rd[i]) - dl) * p.scale;
          const float d1 = p1 * (__uint_as_float(rd[i + 1]) - dl) * p.scale;
          wp[i >> 1] = pack_bf16x2(p0, p1);
          wd[i >> 1] = pack_bf16x2(d0, d1);
        }
        if (c == 0 && t > 0) mbar_wait(pds_free, par ^ 1u);  // the previous pair's gradient UMMAs are done with the tiles
        // columns 32c .. 32c+31 of this row: four 16-byte units in chunk c/2, units 4*(c&1) .. +3, XOR-swizzled
        const uint32_t chunk = (uint32_t)(c >> 1) * 16384u + row_off;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t unit = ((uint32_t)(4 * (c & 1) + u)) ^ swz;
          if (MODE == 0) st_shared_v4(sP + chunk + unit * 16u, wp[4 * u], wp[4 * u + 1], wp[4 * u + 2], wp[4 * u + 3]);
          st_shared_v4(sDS + chunk + unit * 16u, wd[4 * u], wd[4 * u + 1], wd[4 * u + 2], wd[4 * u + 3]);
        }
      }
      tc_fence_before();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> visible to the UMMAs
      __syncwarp();
      if (lane == 0) { mbar_arrive(sdp_free); mbar_arrive(pds_full); }
    }
    // ---- accumulators -> bf16 -> the gradient buffer ----
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int orow = fixed_tile * kB + row;   // key row (MODE 0) or query row (MODE 1)
    const int n_out = MODE == 0 ? 2 : 1;
    for (int a = 0; a < n_out; ++a) {
      const uint32_t t_acc = tmem + ((uint32_t)(quarter * 32) << 16) + (a == 0 ? kColAcc0 : kColAcc1);
      // MODE 0: accumulator 0 = dV, 1 = dK; MODE 1: dQ
      const int col = MODE == 0 ? (a == 0 ? p.v_off : p.k_off) + hk * kDH : p.q_off + hx * kDH;
#pragma unroll
      for (int c = 0; c < kDH / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(t_acc + 32 * c, r);
        tmem_wait_ld();
        if (orow < p.seq) {
          __nv_bfloat16* dst = p.dqkv + ((int64_t)b * p.seq + orow) * p.ld_d + col + 32 * c;
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(r[i]);
#pragma unroll
          for (int v = 0; v < 4; ++v) *reinterpret_cast<uint4*>(dst + 8 * v) = pack8(f + 8 * v);
        }
      }
    }
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)f;
  });
  return fn;
}

// [batch][seq][columns] view of a [batch*seq, ld] bf16 buffer, boxes of 64 columns x 128 rows, 128B swizzle;
// rows past the end of a sequence are zero-filled
int make_map(EncodeTiledFn enc, CUtensorMap* tm, const void* base, int64_t ld, int seq, int batch) {
  cuuint64_t dims[3] = {(cuuint64_t)ld, (cuuint64_t)seq, (cuuint64_t)batch};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)ld * 2 * (cuuint64_t)seq};
  cuuint32_t box[3] = {kDH, kB, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { b200ft_set_error("attn_bwd: cuTensorMapEncodeTiled failed (%d)", (int)r); return B200FT_ERR_ARG; }
  return 0;
}

}  // namespace

extern "C" int b200ft_attn_bwd(const void* qkv, int64_t ld, int q_off, int k_off, int v_off, const void* out,
                               int64_t ld_out, const void* dout, int64_t ld_dout, const float* lse2, float* delta,
                               void* dqkv, int64_t ld_dqkv, int batch, int seq, int n_head, int n_kv_head,
                               int d_head, float scale, void* stream) {
  B200FT_CHECK_ARG(qkv && out && dout && lse2 && delta && dqkv, "attn_bwd: null buffer");
  if (d_head != kDH) {
    b200ft_set_error("attn_bwd: d_head=%d is not built (only 64)", d_head);
    return B200FT_ERR_UNSUPPORTED;
  }
  B200FT_CHECK_ARG(batch > 0 && seq > 0 && n_head > 0 && n_kv_head > 0 && n_head % n_kv_head == 0,
                   "attn_bwd: batch=%d seq=%d heads=%d kv_heads=%d", batch, seq, n_head, n_kv_head);
  B200FT_CHECK_ARG(ld % 8 == 0 && ld_out % 8 == 0 && ld_dout % 8 == 0 && ld_dqkv % 8 == 0 && q_off % 8 == 0 &&
                       k_off % 8 == 0 && v_off % 8 == 0 &&
                       ((uintptr_t)qkv | (uintptr_t)out | (uintptr_t)dout | (uintptr_t)dqkv) % 16 == 0 &&
                       ld_out >= (int64_t)n_head * kDH && ld_dout >= (int64_t)n_head * kDH,
                   "attn_bwd: alignment (ld=%lld ld_out=%lld ld_dout=%lld ld_dqkv=%lld offsets %d/%d/%d)",
                   (long long)ld, (long long)ld_out, (long long)ld_dout, (long long)ld_dqkv, q_off, k_off, v_off);
  EncodeTiledFn enc = encode_fn();
  if (!enc) { b200ft_set_error("attn_bwd: cuTensorMapEncodeTiled unavailable"); return B200FT_ERR_UNSUPPORTED; }
  CUtensorMap tm_qkv, tm_do;
  int rc = make_map(enc, &tm_qkv, qkv, ld, seq, batch);
  if (rc) return rc;
  rc = make_map(enc, &tm_do, dout, ld_dout, seq, batch);
  if (rc) return rc;
  BwdParams p;
  p.batch = batch; p.seq = seq; p.n_head = n_head; p.n_kv_head = n_kv_head;
  p.q_off = q_off; p.k_off = k_off; p.v_off = v_off;
  p.n_tiles = (seq + kB - 1) / kB;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
  p.lse2 = lse2; p.delta = delta;
  p.dqkv = (__nv_bfloat16*)dqkv; p.ld_d = ld_dqkv;
  static thread_local int attr_dev = -1;
  int dev = -1;
  B200FT_CUDA(cudaGetDevice(&dev));
  if (dev != attr_dev) {
    B200FT_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    B200FT_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_dev = dev;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows = (int64_t)batch * seq * n_head;
  attn_delta_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>((const __nv_bfloat16*)out, ld_out,
                                                                (const __nv_bfloat16*)dout, ld_dout, delta, batch,
                                                                seq, n_head);
  B200FT_LAUNCH_CHECK();
  attn_bwd_kernel<0><<<(unsigned)(p.n_tiles * batch * n_kv_head), kThreads, kSmemBytes, st>>>(tm_qkv, tm_do, p);
  B200FT_LAUNCH_CHECK();
  attn_bwd_kernel<1><<<(unsigned)(p.n_tiles * batch * n_head), kThreads, kSmemBytes, st>>>(tm_qkv, tm_do, p);
  B200FT_LAUNCH_CHECK();
  return 0;
}
