// Causal flash-attention BACKWARD on tcgen05 for d_head = 64, the counterpart of attention.cu: reads q, k, v
// from the fused [tokens, q|k|v] projection buffer, o / do from [tokens, heads*64] buffers and the forward's
// per-row log-sum-exp, and writes dq | dk | dv into a buffer with the projection's own column layout.
//
// Launches: delta, then either MODE 2 — the default: MODE 0's work item, which also forms the dQ partial of every
// tile pair (dS K_j) and reduce-adds it, scaled and rounded to bf16, into the pre-zeroed q columns of the gradient
// buffer by TMA (5 tile GEMMs per pair, S / dP / exponentials computed once) — or MODE 0 + MODE 1 (deterministic):
//   delta     D[row] = sum_d do[row,d] * o[row,d]                              (one warp per token)
//   MODE 0    one CTA per (kv tile j, batch, kv head): K_j, V_j resident; streams (Q_i, dO_i) over the q heads
//             of the group and the q tiles i >= j; dV_j += P^T dO_i and dK_j += dS^T Q_i accumulate in TMEM
//   MODE 1    one CTA per (q tile i, batch, head): Q_i, dO_i resident; streams (K_j, V_j), j <= i;
//             dQ_i += dS K_j accumulates in TMEM
// Both modes recompute S = Q K^T and dP = dO V^T per tile pair on the tensor core (no atomics, no fp32 dQ
// buffer; 7 instead of 5 tile GEMMs).  Per pair: warps 4-7 (one thread per query row) read S and dP from
// TMEM, form P = exp2(S*scale*log2e - lse2) and dS' = P * (dP - D) (the softmax scale multiplies the dK / dQ
// accumulators once, at the end), and write them as bf16 into
// 128B-swizzled shared tiles [q rows][kv columns] — the byte layout TMA would produce — which the second
// group of UMMAs consumes MN-major (P^T, dS^T: reduction over q) or K-major (dS: reduction over kv).
// Q, dO, K, V tiles are used in both roles too: K-major for S / dP, MN-major as the B operand of the
// gradient GEMMs.  640 threads: warp 0 TMA, warp 1 MMA issue, warp 2 TMEM allocation, warps 4-19 softmax / dS
// (four per TMEM lane quarter, each owning one 32-key slice of the tile).
#include "attn_common.cuh"

namespace {
using namespace b200ft_attn;

constexpr int kDH = 64;
constexpr int kB = 128;                       // tile edge in queries and in keys
constexpr int kThreads = 640;                 // warps 0-2 TMA / MMA / TMEM, 4-19 softmax
constexpr int kTile = kB * kDH * 2;           // 16 KB: one Q / dO / K / V tile
constexpr int kPTile = kB * kB * 2;           // 32 KB: P or dS, two 64-column chunks of 16 KB
constexpr int kOffF = 0;                      // resident pair
constexpr int kRing = 3;                      // stream ring: pair g feeds its gradient UMMAs while g+1 feeds dP and g+2 feeds S
constexpr int kOffS = 2 * kTile;              // ring stages of (tile0, tile1)
constexpr int kOffP = kOffS + kRing * 2 * kTile;
constexpr int kOffDS = kOffP + kPTile;
constexpr int kOffDQ = kOffDS + kPTile;       // MODE 2: eight 2 KB staging tiles (32 rows x 32 bf16) for the dQ reduce,
constexpr int kOffK1 = kOffDQ + 8 * 2048;     //         then the second buffer of the resident K tile (16 KB)
constexpr int kOffBar = kOffDQ + 8 * 4096;    // 224 KB
constexpr int kSmemBytes = kOffBar + 256;
// S is double-buffered (+128) in the split modes; the fused mode keeps one S buffer and puts its two dQ-partial
// buffers (64 columns each) where the second one would be
constexpr uint32_t kColS = 0, kColDQ = 128, kColDP = 256, kColAcc0 = 384, kColAcc1 = 448;

struct BwdParams {
  int batch, seq, n_head, n_kv_head;
  int q_off, k_off, v_off;
  int n_tiles;
  float scale, scale_log2;
  const float* lse2;     // [batch*n_head, seq]
  const float* delta;    // [batch*n_head, seq]
  __nv_bfloat16* dqkv; int64_t ld_d;
};

// ------------------------------------------------------------------------------------------------
// D[bh, t] = sum_d do[t, h*DH + d] * o[t, h*DH + d].  One warp per token: lanes walk the row in 16-byte units
// (8 or 16 units = one head), partial dot products meet across the lanes of a head by shuffle.
template <int DH>
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __nv_bfloat16* __restrict__ o, int64_t ld_o, const __nv_bfloat16* __restrict__ dout,
                  int64_t ld_do, float* __restrict__ delta, int batch, int seq, int n_head) {
  constexpr int kUPH = DH / 8;                                        // 16-byte units per head: 8 or 16 lanes
  grid_dep_sync();
  const int64_t tok = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);   // b * seq + t
  if (tok >= (int64_t)batch * seq) return;
  const int lane = threadIdx.x & 31;
  const int b = (int)(tok / seq), t = (int)(tok % seq);
  const int units = n_head * kUPH;
  for (int u0 = 0; u0 < units; u0 += 32) {       // 32 units = 4 heads per pass; all lanes stay in the shuffles
    const int u = u0 + lane;
    float acc = 0.f;
    if (u < units) {
      float a[8], g[8];
      unpack8(ld_stream16(o + tok * ld_o + 8 * u), a);
      unpack8(ld_stream16(dout + tok * ld_do + 8 * u), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc = fmaf(a[j], g[j], acc);
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if (kUPH == 16) acc += __shfl_xor_sync(0xffffffffu, acc, 8);
    if (u < units && (lane & (kUPH - 1)) == 0) delta[((int64_t)b * n_head + u / kUPH) * seq + t] = acc;
  }
}

// Work item of a CTA.  MODE 0: (kv tile ascending, batch, kv head); MODE 1: (q tile descending, batch, head):
// heaviest first.  Every warp role walks the same item sequence w = blockIdx.x, += gridDim.x.
#ifdef B200FT_ATTN_TIMING   // per-phase clock readings of CTA 0 (tools/attn_timing.py); never defined in the shipped build
__device__ long long g_dlog[2][4096][2];
__device__ int g_dcnt[2];
#define DLOG(role, tag) do { if (MODE == 2 && blockIdx.x == 0) { int i__ = g_dcnt[role]++; if (i__ < 4096) { g_dlog[role][i__][0] = (tag); g_dlog[role][i__][1] = clock64(); } } } while (0)
#else
#define DLOG(role, tag) do { } while (0)
#endif

struct Item {
  int fixed_tile, b, hx, hk, per_head, n_iter;
};
template <int MODE>
__device__ __forceinline__ Item decode_item(const BwdParams& p, int w, int G) {
  Item it;
  const int heads = MODE != 1 ? p.n_kv_head : p.n_head;
  const int per_tile = p.batch * heads;
  const int tile_ord = w / per_tile, rest = w % per_tile;
  it.fixed_tile = MODE != 1 ? tile_ord : p.n_tiles - 1 - tile_ord;
  it.b = rest / heads;
  it.hx = rest % heads;                       // kv head (MODE 0) or q head (MODE 1)
  it.hk = MODE != 1 ? it.hx : it.hx / G;
  it.per_head = MODE != 1 ? p.n_tiles - it.fixed_tile : it.fixed_tile + 1;   // streamed tiles per head
  it.n_iter = MODE != 1 ? G * it.per_head : it.per_head;
  return it;
}

// Persistent: one CTA per SM walks work items.  Barrier parities run on two counters shared by all roles:
// `g` = tile pairs processed so far by this CTA (stream ring stage g % 3; S buffer g & 1; dP and P/dS hand-offs),
// `n` = items processed so far (resident tiles, accumulators).
//
// Software pipeline (what hides the tensor core behind the softmax warps and vice versa): S is double-buffered
// in TMEM, so S(g+1) is already there when the softmax warps finish pair g and they start exponentiating at
// once (phase A: P = exp2(S*c - lse), the MUFU-bound part, needs S only); meanwhile the tensor core produces
// dP(g+1) and the gradient UMMAs of pair g; phase B (dS' = P * (dP - D)) picks dP(g+1) up when it is done.
// MMA warp, at "softmax of pair g complete":  dP(g+1), gradients(g), S(g+2).
template <int MODE>
__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                const __grid_constant__ CUtensorMap tm_dq, const BwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  if (raw & 1023u) __trap();  // dynamic shared memory starts 1024-aligned (no static smem in this kernel)
  const uint32_t sF0 = raw + kOffF, sF1 = sF0 + kTile;
  auto sS0 = [&](uint32_t st) { return raw + kOffS + st * 2u * kTile; };
  auto sS1 = [&](uint32_t st) { return raw + kOffS + st * 2u * kTile + kTile; };
  const uint32_t sP = raw + kOffP, sDS = raw + kOffDS;
  const uint32_t bars = raw + kOffBar;
  const uint32_t f_full = bars, f_empty = bars + 8u;
  auto st_full = [&](uint32_t s) { return bars + 8u * (2 + s); };
  auto st_empty = [&](uint32_t s) { return bars + 8u * (5 + s); };
  auto s_full = [&](uint32_t bf) { return bars + 8u * (8 + bf); };
  auto s_free = [&](uint32_t bf) { return bars + 8u * (10 + bf); };
  const uint32_t dp_full = bars + 8u * 12, pds_full = bars + 8u * 13, pds_free = bars + 8u * 14;
  const uint32_t acc_full = bars + 8u * 15, acc_free = bars + 8u * 16;
  auto dq_full = [&](uint32_t bf) { return bars + 8u * (17 + bf); };   // MODE 2: dQ partial of a pair is in TMEM
  auto dq_free = [&](uint32_t bf) { return bars + 8u * (19 + bf); };   //         ... and has been read out
  // MODE 2: the resident K tile is double-buffered (item n uses buffer n & 1) and V is released after the item's last
  // dP product, so that the next item's K and V are in shared memory when the current item ends.  At T = 1024 an item
  // has 4.5 pairs on average, and reloading the resident pair after the last UMMA of an item (the dQ partials read K
  // to the end) left the tensor core idle for ~4000 cycles at every boundary.
  auto k_full = [&](uint32_t b) { return bars + 8u * (21 + b); };
  auto k_empty = [&](uint32_t b) { return bars + 8u * (23 + b); };
  const uint32_t v_full = bars + 8u * 25, v_empty = bars + 8u * 26;
  const uint32_t sK1 = raw + kOffK1;
  const uint32_t tmem_slot = bars + 8u * 27;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = p.n_head / p.n_kv_head;
  const int num_items = p.n_tiles * p.batch * (MODE != 1 ? p.n_kv_head : p.n_head);

  if (warp == 1 && lane == 0) {
    mbar_init(f_full, 1); mbar_init(f_empty, 1);
    for (uint32_t s = 0; s < kRing; ++s) { mbar_init(st_full(s), 1); mbar_init(st_empty(s), 1); }
    for (uint32_t bf = 0; bf < 2; ++bf) { mbar_init(s_full(bf), 1); mbar_init(s_free(bf), 16); }
    mbar_init(dp_full, 1); mbar_init(pds_full, 16); mbar_init(pds_free, 1);
    mbar_init(acc_full, 1); mbar_init(acc_free, 16);
    for (uint32_t bf = 0; bf < 2; ++bf) { mbar_init(dq_full(bf), 1); mbar_init(dq_free(bf), 8); }
    for (uint32_t b = 0; b < 2; ++b) { mbar_init(k_full(b), 1); mbar_init(k_empty(b), 1); }
    mbar_init(v_full, 1); mbar_init(v_empty, 1);
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
  grid_dep_sync();  // the prologue above overlaps the previous kernel's tail; global memory only from here on

  // Items are sorted heaviest first and dealt out in a snake: round n hands item n*grid + b to CTA b when n is even and
  // n*grid + (grid-1-b) when n is odd, so a CTA that drew a heavy item in one round draws a light one in the next.
  auto item_at = [&](uint32_t round) -> int {
    const int i = (round & 1u) ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x;
    return (int)(round * gridDim.x) + i;
  };

  if (warp == 0) {
    // ================================ TMA producer ================================
    const bool issuer = elect_one();
    uint32_t g = 0, n = 0;
    for (int w = item_at(0); w < num_items; w = item_at(++n)) {
      const Item it = decode_item<MODE>(p, w, G);
      if (MODE == 2) {
        if (n == 0 && issuer) {
          mbar_expect_tx(k_full(0), kTile);
          tma_load_3d(sF0, &tm_qkv, p.k_off + it.hk * kDH, it.fixed_tile * kB, it.b, k_full(0));
        }
        if (n > 0) mbar_wait(v_empty, (n - 1) & 1u);   // the previous item's last dP product has read V
        if (issuer) {
          mbar_expect_tx(v_full, kTile);
          tma_load_3d(sF1, &tm_qkv, p.v_off + it.hk * kDH, it.fixed_tile * kB, it.b, v_full);
        }
        __syncwarp();
      }
      if (MODE != 2 && n > 0) mbar_wait(f_empty, (n - 1) & 1u);   // the previous item's last dP UMMAs have read the resident pair
      if (MODE != 2 && issuer) {
        mbar_expect_tx(f_full, 2 * kTile);
        if (MODE != 1) {
          tma_load_3d(sF0, &tm_qkv, p.k_off + it.hk * kDH, it.fixed_tile * kB, it.b, f_full);
          tma_load_3d(sF1, &tm_qkv, p.v_off + it.hk * kDH, it.fixed_tile * kB, it.b, f_full);
        } else {
          tma_load_3d(sF0, &tm_qkv, p.q_off + it.hx * kDH, it.fixed_tile * kB, it.b, f_full);
          tma_load_3d(sF1, &tm_do, it.hx * kDH, it.fixed_tile * kB, it.b, f_full);
        }
      }
      __syncwarp();
      for (int t = 0; t < it.n_iter; ++t, ++g) {
        const uint32_t st = g % kRing;
        mbar_wait(st_empty(st), ((g / kRing) & 1u) ^ 1u);
        if (issuer) {
          mbar_expect_tx(st_full(st), 2 * kTile);
          if (MODE != 1) {
            const int hq = it.hk * G + t / it.per_head, qt = it.fixed_tile + t % it.per_head;
            tma_load_3d(sS0(st), &tm_qkv, p.q_off + hq * kDH, qt * kB, it.b, st_full(st));
            tma_load_3d(sS1(st), &tm_do, hq * kDH, qt * kB, it.b, st_full(st));
          } else {
            tma_load_3d(sS0(st), &tm_qkv, p.k_off + it.hk * kDH, t * kB, it.b, st_full(st));
            tma_load_3d(sS1(st), &tm_qkv, p.v_off + it.hk * kDH, t * kB, it.b, st_full(st));
          }
        }
        __syncwarp();
      }
      if (MODE == 2) {
        // K of the NEXT item into the other buffer, requested once this item's streamed tiles are all on their way
        // (the producer runs up to three pairs ahead of the tensor core): the buffer's last user was item n - 1
        const int w_next = item_at(n + 1);
        if (w_next < num_items) {
          const Item nx = decode_item<MODE>(p, w_next, G);
          const uint32_t nb = (n + 1) & 1u;
          if (n >= 1) mbar_wait(k_empty(nb), ((n - 1) >> 1) & 1u);
          if (issuer) {
            mbar_expect_tx(k_full(nb), kTile);
            tma_load_3d(nb ? sK1 : sF0, &tm_qkv, p.k_off + nx.hk * kDH, nx.fixed_tile * kB, nx.b, k_full(nb));
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    const bool issuer = elect_one();
    const uint32_t id_s = idesc(kB, 0, 0);     // S, dP: 128 x 128 x 64, both operands K-major
    const uint32_t id_g0 = idesc(kDH, 1, 1);   // dV, dK: 128 (keys) x 64 x 128 (queries): A = P^T / dS^T, B = dO / Q, MN-major
    const uint32_t id_g1 = idesc(kDH, 0, 1);   // dQ: 128 (queries) x 64 x 128 (keys): A = dS K-major, B = K MN-major
    uint32_t sK = sF0;                         // MODE 2: the current item's buffer of the resident K tile
    auto issue_s = [&](uint32_t gg) {          // S(gg) = Q K^T into S buffer gg & 1 (MODE 2: the only buffer)
      const uint32_t st = gg % kRing, bf = MODE == 2 ? 0u : gg & 1u;
      mbar_wait(st_full(st), (gg / kRing) & 1u);
      if (MODE == 2) { if (gg >= 1) mbar_wait(s_free(0), (gg - 1) & 1u); }   // phase A of the previous pair has read S
      else if (gg >= 2) mbar_wait(s_free(bf), ((gg >> 1) - 1) & 1u);         // phase A of pair gg-2 has read this buffer
      tc_fence_after();
      const uint32_t aq = MODE != 1 ? sS0(st) : sF0, bk = MODE != 1 ? sK : sS0(st);
      if (issuer) {
#ifndef B200FT_EXP_NO_SDP
        // descriptors once per product, the k-th step's offset added to the address field (16-byte units; the tiles lie
        // below 256 KB, so the add never carries out of it): the single issuing thread is on every pair's critical path
        const uint64_t da = sdesc(aq, 16, 1024), db = sdesc(bk, 16, 1024);
#pragma unroll
        for (int k = 0; k < kDH / 16; ++k)
          umma(tmem + kColS + 128u * bf, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), id_s, k > 0);
#endif
        umma_commit(s_full(bf));
      }
      __syncwarp();
    };
    auto issue_dp = [&](uint32_t gg, bool last_of_item) {   // dP(gg) = dO V^T (the caller knows the dP buffer is free)
      const uint32_t st = gg % kRing;
      const uint32_t ado = MODE != 1 ? sS1(st) : sF1, bv = MODE != 1 ? sF1 : sS1(st);
      if (issuer) {
#ifndef B200FT_EXP_NO_SDP
        const uint64_t da = sdesc(ado, 16, 1024), db = sdesc(bv, 16, 1024);
#pragma unroll
        for (int k = 0; k < kDH / 16; ++k)
          umma(tmem + kColDP, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), id_s, k > 0);
#endif
        umma_commit(dp_full);
        if (last_of_item) umma_commit(MODE == 2 ? v_empty : f_empty);   // nothing of this item reads the resident pair (MODE 2: V) any more
      }
      __syncwarp();
    };
    auto issue_grads = [&](uint32_t gg, bool first) {
      const uint32_t st = gg % kRing;
      if (issuer) {
#ifndef B200FT_EXP_NO_GRADS   // bottleneck experiment: skip the gradient UMMAs (commits stay), wrong numbers
        if (MODE != 1) {
          const uint64_t dp_ = sdesc(sP, 16384, 1024), dds = sdesc(sDS, 16384, 1024);
          const uint64_t d1 = sdesc(sS1(st), 8192, 1024), d0 = sdesc(sS0(st), 8192, 1024);
#pragma unroll
          for (int k = 0; k < kB / 16; ++k)   // reduction over 16 query rows per step (2 KB = 128 units)
            umma(tmem + kColAcc0, dp_ + (uint64_t)(128 * k), d1 + (uint64_t)(128 * k), id_g0, !(first && k == 0));
#pragma unroll
          for (int k = 0; k < kB / 16; ++k)
            umma(tmem + kColAcc1, dds + (uint64_t)(128 * k), d0 + (uint64_t)(128 * k), id_g0, !(first && k == 0));
        } else {
#pragma unroll
          for (int k = 0; k < kB / 16; ++k)   // reduction over 16 keys per step: dS chunk k/4, 32 B per step inside
            umma(tmem + kColAcc0, sdesc(sDS + 16384 * (k >> 2) + 32 * (k & 3), 16, 1024),
                 sdesc(sS0(st) + 2048 * k, 8192, 1024), id_g1, !(first && k == 0));
        }
        if (MODE == 2) {
          // dQ partial of this pair: 128 (queries) x 64 x 128 (keys), A = dS K-major, B = the resident K tile MN-major
          const uint32_t dq_col = kColDQ + 64u * (gg & 1u);
          const uint64_t dds1 = sdesc(sDS, 16, 1024), dk1 = sdesc(sK, 8192, 1024);
#pragma unroll
          for (int k = 0; k < kB / 16; ++k)
            umma(tmem + dq_col, dds1 + (uint64_t)(1024 * (k >> 2) + 2 * (k & 3)), dk1 + (uint64_t)(128 * k), id_g1, k > 0);
        }
#endif
        umma_commit(pds_free);      // P / dS tiles may be overwritten
        umma_commit(st_empty(st));  // and the streamed pair's slot refilled
        if (MODE == 2) umma_commit(dq_full(gg & 1u));
      }
      __syncwarp();
    };
    uint32_t g = 0, n = 0;
    for (int w = item_at(0); w < num_items; w = item_at(++n)) {
      const Item it = decode_item<MODE>(p, w, G);
      const int N = it.n_iter;
      if (MODE == 2) { sK = (n & 1u) ? sK1 : sF0; mbar_wait(k_full(n & 1u), (n >> 1) & 1u); }
      else mbar_wait(f_full, n & 1u);
      // item prologue: S(first), dP(first), S(second).  The dP buffer is free: the previous item's last pair
      // was waited for below before its gradients were issued.
      issue_s(g);
      if (MODE == 2) { mbar_wait(v_full, n & 1u); tc_fence_after(); }
      issue_dp(g, N == 1);
      if (MODE != 2 && N > 1) issue_s(g + 1);
      for (int t = 0; t < N; ++t, ++g) {
        if (issuer) DLOG(1, 10);
        if (MODE == 2 && t + 1 < N) issue_s(g + 1);   // waits for phase A of pair g; S is single-buffered here
        if (issuer) DLOG(1, 11);
        mbar_wait(pds_full, g & 1u);        // softmax warps are done with pair g: P, dS in smem, dP read out
        tc_fence_after();
        if (issuer) DLOG(1, 12);
        if (t + 1 < N) issue_dp(g + 1, t + 2 == N);
        if (t == 0 && n > 0) { mbar_wait(acc_free, (n - 1) & 1u); tc_fence_after(); }  // previous item's accumulators read out
        if (issuer) DLOG(1, 13);
        if (MODE == 2 && g >= 2) { mbar_wait(dq_free(g & 1u), ((g >> 1) - 1) & 1u); tc_fence_after(); }
        if (issuer) DLOG(1, 14);
        issue_grads(g, t == 0);
        if (issuer) DLOG(1, 15);
        if (MODE != 2 && t + 2 < N) issue_s(g + 2);
      }
      if (issuer) {
        umma_commit(acc_full);
        if (MODE == 2) umma_commit(k_empty(n & 1u));   // the dQ partials read the resident K tile to the end of the item
      }
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ================================ softmax / dS ================================
    // Four warps per TMEM lane quarter (= per scheduler: they hide each other's tcgen05.ld / MUFU / st.shared
    // latencies): a thread owns one query row and one 32-key slice of the tile.
    const int quarter = warp & 3;
    const int part = (warp - 4) >> 2;            // 32-key slice 0..3
    const int row = quarter * 32 + lane;
    const uint32_t t_lane = tmem + ((uint32_t)(quarter * 32) << 16);
    const uint32_t t_dp = t_lane + kColDP + 32u * part;
    const uint32_t swz = (uint32_t)(row & 7);
    // slice -> 64-key smem chunk part/2, 16-byte units 4*(part&1) .. +3 of the row
    const uint32_t tile_off = (uint32_t)(part >> 1) * 16384u + (uint32_t)row * 128u;
    const uint32_t unit0 = 4u * (uint32_t)(part & 1);
    const f32x2_t sc2 = pk2(p.scale_log2, p.scale_log2);
    auto store_slice = [&](uint32_t base, const uint32_t* wv) {   // 32 bf16 of this row -> four swizzled 16-byte units
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t unit = (unit0 + (uint32_t)u) ^ swz;
#ifdef B200FT_EXP_NO_STORE
        if (wv[4 * u] == 0x12345678u)   // practically never: the stores are compiled but not executed
#endif
        st_shared_v4(base + tile_off + unit * 16u, wv[4 * u], wv[4 * u + 1], wv[4 * u + 2], wv[4 * u + 3]);
      }
    };
    // MODE 2: dQ partial of pair gg -> staging tile -> TMA reduce-add (bf16) into the q columns of the gradient buffer
    // [batch][seq][columns], which the host zeroed; rows past the end of the sequence are clipped by TMA.  Slices 0 and 1 do it (32 rows x
    // 32 columns per warp); the other two warps of the quarter run ahead into the next pair's exponentials.
    const uint32_t stage_dq = raw + kOffDQ + (uint32_t)((part & 1) * 4 + quarter) * 2048u;
    auto drain_dq = [&](uint32_t gg, int hq, int qt, int bb) {
      const uint32_t bq = gg & 1u;
      mbar_wait(dq_full(bq), (gg >> 1) & 1u);
      tc_fence_after();
#ifdef B200FT_EXP_NO_DRAIN
      __syncwarp();
      if (lane == 0) mbar_arrive(dq_free(bq));
      (void)hq; (void)qt; (void)bb;
      return;
#endif
      uint32_t r[32];
      tmem_ld32(t_lane + kColDQ + 64u * bq + 32u * part, r);
      tmem_wait_ld();
      tc_fence_before();
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the previous reduce has read the tile
      __syncwarp();
      if (lane == 0) mbar_arrive(dq_free(bq));
      // x softmax scale, to bf16: the partial is ADDED in bf16 straight into the q columns of the gradient buffer
      // (a 32 x 32 bf16 tile, 64-byte rows, unswizzled)
      uint32_t wq[16];
#pragma unroll
      for (int i = 0; i < 16; ++i)
        wq[i] = pack_bf16x2(__uint_as_float(r[2 * i]) * p.scale, __uint_as_float(r[2 * i + 1]) * p.scale);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        st_shared_v4(stage_dq + (uint32_t)lane * 64u + ((uint32_t)u << 4), wq[4 * u], wq[4 * u + 1], wq[4 * u + 2], wq[4 * u + 3]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        asm volatile(
            "cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
            ::"l"(reinterpret_cast<uint64_t>(&tm_dq)), "r"(stage_dq), "r"(p.q_off + hq * kDH + 32 * part),
              "r"(qt * kB + 32 * quarter), "r"(bb) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    };
    const bool drains = MODE == 2 && part < 2;
    uint32_t g = 0, n = 0;
    for (int w = item_at(0); w < num_items; w = item_at(++n)) {
      const Item it = decode_item<MODE>(p, w, G);
      // per-row log-sum-exp and D.  Rows past the end of the sequence get lse = +inf: P = dS = 0.  With the kv tile
      // fixed (MODE 0, 2) every pair has its own query rows: the values of pair t+1 are requested while pair t
      // is processed, so their global-memory latency is never on the critical path.
      // (head, q tile) of pair t = (it.hk * G + t / per_head, fixed_tile + t % per_head) when the kv tile is fixed: kept
      // as counters that advance with t — the divisions, done per pair by each of the 512 threads, cost about as many
      // instructions as the pair's arithmetic
      auto load_row_stats = [&](int hq, int qt_, float& lse_o, float& dl_o) {
        const int qr = qt_ * kB + row;
        lse_o = INFINITY; dl_o = 0.f;
        if (qr < p.seq) {
          const int64_t idx = ((int64_t)it.b * p.n_head + hq) * p.seq + qr;
          lse_o = __ldg(p.lse2 + idx); dl_o = __ldg(p.delta + idx);
        }
      };
      float lse, dl, lse_next = INFINITY, dl_next = 0.f;
      int hq_cur = MODE == 1 ? it.hx : it.hk * G, qt_cur = it.fixed_tile;   // pair t
      int hq_prev = hq_cur, qt_prev = qt_cur;                               // pair t - 1 (for the dQ drain)
      load_row_stats(hq_cur, qt_cur, lse, dl);
      for (int t = 0; t < it.n_iter; ++t, ++g) {
        int qt, kt;
        int hq_next = hq_cur, qt_next = qt_cur + 1;                         // pair t + 1
        if (MODE != 1) {
          if (qt_next == it.fixed_tile + it.per_head) { qt_next = it.fixed_tile; ++hq_next; }
          qt = qt_cur; kt = it.fixed_tile;
          if (t > 0) { lse = lse_next; dl = dl_next; }
          if (t + 1 < it.n_iter) load_row_stats(hq_next, qt_next, lse_next, dl_next);
        } else {
          qt = it.fixed_tile; kt = t;
        }
        const int qrow = qt * kB + row;
        const int kcol0 = kt * kB + 32 * part;
        const bool diag = qt == kt;
        const f32x2_t nl2 = pk2(-lse, -lse), dl2 = pk2(dl, dl);
        const uint32_t bf = MODE == 2 ? 0u : g & 1u;
        const uint32_t t_s = t_lane + kColS + 128u * bf + 32u * part;
        // ---- phase A: P = exp2(S*c - lse) for this thread's 32 keys (kept in registers for phase B) ----
        f32x2_t P2[16];
        const bool tl = (warp == 4 && lane == 0);
        if (tl) DLOG(0, 0);
        mbar_wait(s_full(bf), MODE == 2 ? g & 1u : (g >> 1) & 1u);
        tc_fence_after();
        if (tl) DLOG(0, 1);
        {
          uint32_t rs[32];
          tmem_ld32(t_s, rs);
          tmem_wait_ld();
          if (diag) {                   // warp-uniform; only the diagonal pair carries masking code
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float e0, e1;
              upk2(fma2(pk2(__uint_as_float(rs[i]), __uint_as_float(rs[i + 1])), sc2, nl2), e0, e1);
              const float p0 = kcol0 + i > qrow ? 0.f : ex2(e0);
              const float p1 = kcol0 + i + 1 > qrow ? 0.f : ex2(e1);
              P2[i >> 1] = pk2(p0, p1);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float e0, e1;
              upk2(fma2(pk2(__uint_as_float(rs[i]), __uint_as_float(rs[i + 1])), sc2, nl2), e0, e1);
              P2[i >> 1] = pk2(ex2(e0), ex2(e1));
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free(bf));          // the next S into this buffer may be issued
        if (tl) DLOG(0, 2);
        if (MODE != 1) {
          if (g > 0) mbar_wait(pds_free, (g - 1) & 1u);  // the previous pair's gradient UMMAs are done with the tiles
          if (tl) DLOG(0, 3);
          uint32_t wp[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) wp[i] = f32x2_to_bf16x2(P2[i]);
          store_slice(sP, wp);
        }
        // ---- phase B: dS' = P * (dP - D) ----
        if (tl) DLOG(0, 4);
        mbar_wait(dp_full, g & 1u);
        tc_fence_after();
        if (tl) DLOG(0, 5);
        {
          uint32_t rd[32];
          tmem_ld32(t_dp, rd);
          tmem_wait_ld();
          if (MODE == 1 && g > 0) mbar_wait(pds_free, (g - 1) & 1u);
          uint32_t wd[16];
#pragma unroll
          for (int i = 0; i < 16; ++i)
            wd[i] = f32x2_to_bf16x2(mul2(P2[i], sub2(pk2(__uint_as_float(rd[2 * i]), __uint_as_float(rd[2 * i + 1])), dl2)));
          store_slice(sDS, wd);
        }
        tc_fence_before();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> visible to the UMMAs
        __syncwarp();
        if (lane == 0) mbar_arrive(pds_full);
        if (tl) DLOG(0, 6);
        if (drains && t > 0)   // the previous pair's dQ partial: its UMMAs ran while this pair was exponentiated
          drain_dq(g - 1, hq_prev, qt_prev, it.b);
        if (tl) DLOG(0, 7);
        hq_prev = hq_cur; qt_prev = qt_cur;
        hq_cur = hq_next; qt_cur = qt_next;
      }
      if (drains) drain_dq(g - 1, hq_prev, qt_prev, it.b);
      // ---- accumulators -> (x softmax scale for dK) -> bf16 -> the gradient buffer ----
      mbar_wait(acc_full, n & 1u);
      tc_fence_after();
      const int orow = it.fixed_tile * kB + row;   // key row (MODE 0, 2) or query row (MODE 1)
      if (MODE != 1) {
        // slices 0, 1 write the two 32-column halves of dV (accumulator 0), slices 2, 3 those of dK (accumulator 1)
        const uint32_t t_acc = t_lane + ((part >> 1) == 0 ? kColAcc0 : kColAcc1) + 32u * (uint32_t)(part & 1);
        const int col = ((part >> 1) == 0 ? p.v_off : p.k_off) + it.hk * kDH + 32 * (part & 1);
        const float oscale = (part >> 1) == 0 ? 1.f : p.scale;
        uint32_t r[32];
        tmem_ld32(t_acc, r);
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_free);    // the next item's gradient UMMAs may overwrite the accumulators
        if (orow < p.seq) {
          __nv_bfloat16* dst = p.dqkv + ((int64_t)it.b * p.seq + orow) * p.ld_d + col;
          float fv[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) fv[i] = __uint_as_float(r[i]) * oscale;
#pragma unroll
          for (int v = 0; v < 4; ++v) *reinterpret_cast<uint4*>(dst + 8 * v) = pack8(fv + 8 * v);
        }
      } else {
        // dQ: the four slices write 16 columns each
        uint32_t r[16];
        tmem_ld16(t_lane + kColAcc0 + 16u * part, r);
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_free);
        if (orow < p.seq) {
          __nv_bfloat16* dst = p.dqkv + ((int64_t)it.b * p.seq + orow) * p.ld_d + p.q_off + it.hx * kDH + 16 * part;
          float fv[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) fv[i] = __uint_as_float(r[i]) * p.scale;
#pragma unroll
          for (int v = 0; v < 2; ++v) *reinterpret_cast<uint4*>(dst + 8 * v) = pack8(fv + 8 * v);
        }
      }
    }
    if (drains && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // reductions complete before exit
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
}

}  // namespace

// workspace: delta [batch*n_head*seq] fp32
static size_t delta_bytes(int batch, int seq, int n_head) {
  return (((size_t)batch * n_head * seq * sizeof(float)) + 255) & ~(size_t)255;
}

extern "C" size_t b200ft_attn_bwd_workspace_bytes(int batch, int seq, int n_head, int d_head) {
  (void)d_head;
  return delta_bytes(batch, seq, n_head);
}

extern "C" int b200ft_attn_bwd(const void* qkv, int64_t ld, int q_off, int k_off, int v_off, const void* out,
                               int64_t ld_out, const void* dout, int64_t ld_dout, const float* lse2, void* workspace,
                               size_t workspace_bytes, void* dqkv, int64_t ld_dqkv, int batch, int seq, int n_head,
                               int n_kv_head, int d_head, float scale, int flags, void* stream) {
  B200FT_CHECK_ARG(qkv && out && dout && lse2 && workspace && dqkv, "attn_bwd: null buffer");
  if (d_head != 64 && d_head != 128) {
    b200ft_set_error("attn_bwd: d_head=%d is not built (64 and 128 are)", d_head);
    return B200FT_ERR_UNSUPPORTED;
  }
  B200FT_CHECK_ARG(batch > 0 && seq > 0 && n_head > 0 && n_kv_head > 0 && n_head % n_kv_head == 0,
                   "attn_bwd: batch=%d seq=%d heads=%d kv_heads=%d", batch, seq, n_head, n_kv_head);
  B200FT_CHECK_ARG(ld % 8 == 0 && ld_out % 8 == 0 && ld_dout % 8 == 0 && ld_dqkv % 8 == 0 && q_off % 8 == 0 &&
                       k_off % 8 == 0 && v_off % 8 == 0 &&
                       ((uintptr_t)qkv | (uintptr_t)out | (uintptr_t)dout | (uintptr_t)dqkv) % 16 == 0 &&
                       (uintptr_t)workspace % 256 == 0 && ld_out >= (int64_t)n_head * d_head &&
                       ld_dout >= (int64_t)n_head * d_head,
                   "attn_bwd: alignment (ld=%lld ld_out=%lld ld_dout=%lld ld_dqkv=%lld offsets %d/%d/%d)",
                   (long long)ld, (long long)ld_out, (long long)ld_dout, (long long)ld_dqkv, q_off, k_off, v_off);
  // d_head 128: always the two-kernel form (dK|dV per kv tile, dQ per q tile; no atomics, bit-reproducible) of
  // attention_bwd_d128.cu — TMEM has no room for a dQ partial beside S, dP, dK and dV at that head size
  const bool split = (flags & B200FT_ATTN_BWD_DETERMINISTIC) != 0 || d_head == 128;
  const size_t need = delta_bytes(batch, seq, n_head);
  if (workspace_bytes < need) {
    b200ft_set_error("attn_bwd: workspace %zu < %zu", workspace_bytes, need);
    return B200FT_ERR_WORKSPACE;
  }
  float* delta = (float*)workspace;
  EncodeTiledFn enc = encode_fn();
  if (!enc) { b200ft_set_error("attn_bwd: cuTensorMapEncodeTiled unavailable"); return B200FT_ERR_UNSUPPORTED; }
  CUtensorMap tm_qkv, tm_do, tm_dq;
  int rc = make_map(enc, &tm_qkv, qkv, ld, seq, batch);
  if (rc) return rc;
  rc = make_map(enc, &tm_do, dout, ld_dout, seq, batch);
  if (rc) return rc;
  tm_dq = tm_do;  // unused by the split kernels
  if (!split) {
    // the gradient buffer itself as [batch][seq][columns] bf16, boxes of 32 columns x 32 rows, unswizzled: the dQ
    // partials are reduce-added into its q columns in bf16 (one rounding per visible kv tile: T/128 of them per element)
    cuuint64_t dims[3] = {(cuuint64_t)ld_dqkv, (cuuint64_t)seq, (cuuint64_t)batch};
    cuuint64_t strides[2] = {(cuuint64_t)ld_dqkv * 2, (cuuint64_t)ld_dqkv * 2 * (cuuint64_t)seq};
    cuuint32_t box[3] = {32, 32, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&tm_dq, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, dqkv, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { b200ft_set_error("attn_bwd: cuTensorMapEncodeTiled(dq) failed (%d)", (int)r); return B200FT_ERR_ARG; }
  }
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
    B200FT_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_dev = dev;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t rows = (int64_t)batch * seq;
  if (d_head == 128) {
    B200FT_CUDA(launch_k(attn_delta_kernel<128>, (unsigned)((rows + 7) / 8), 256, 0, st, (const __nv_bfloat16*)out, ld_out,
                         (const __nv_bfloat16*)dout, ld_dout, delta, batch, seq, n_head));
    Bwd128Args a;
    a.batch = batch; a.seq = seq; a.n_head = n_head; a.n_kv_head = n_kv_head;
    a.q_off = q_off; a.k_off = k_off; a.v_off = v_off; a.scale = scale;
    a.lse2 = lse2; a.delta = delta; a.dqkv = (__nv_bfloat16*)dqkv; a.ld_d = ld_dqkv;
    a.qkv = (const __nv_bfloat16*)qkv; a.ld_qkv = ld; a.dout = (const __nv_bfloat16*)dout; a.ld_dout = ld_dout;
    return launch_bwd_d128(tm_qkv, tm_do, a, st);
  }
  B200FT_CUDA(launch_k(attn_delta_kernel<64>, (unsigned)((rows + 7) / 8), 256, 0, st, (const __nv_bfloat16*)out, ld_out,
                       (const __nv_bfloat16*)dout, ld_dout, delta, batch, seq, n_head));
  // persistent: one CTA per SM walks the work items (heaviest first, round-robin)
  const int items0 = p.n_tiles * batch * n_kv_head, items1 = p.n_tiles * batch * n_head;
  const unsigned grid0 = (unsigned)(items0 < B200FT_NUM_SMS ? items0 : B200FT_NUM_SMS);
  if (split) {
    B200FT_CUDA(launch_k(attn_bwd_kernel<0>, grid0, kThreads, kSmemBytes, st, tm_qkv, tm_do, tm_dq, p));
    B200FT_CUDA(launch_k(attn_bwd_kernel<1>, (unsigned)(items1 < B200FT_NUM_SMS ? items1 : B200FT_NUM_SMS), kThreads,
                         kSmemBytes, st, tm_qkv, tm_do, tm_dq, p));
    return 0;
  }
  // zero the q columns (a strided DMA fill, no kernel); dk | dv are written whole by the kernel
  B200FT_CUDA(cudaMemset2DAsync((__nv_bfloat16*)dqkv + q_off, (size_t)ld_dqkv * 2, 0, (size_t)n_head * kDH * 2, (size_t)rows, st));
  B200FT_CUDA(launch_k(attn_bwd_kernel<2>, grid0, kThreads, kSmemBytes, st, tm_qkv, tm_do, tm_dq, p));
  return 0;
}

#ifdef B200FT_ATTN_TIMING
extern "C" int b200ft_debug_attn_bwd64_timing(long long* log_out, int* counts_out, int reset) {
  if (reset) {
    int z[2] = {0, 0};
    B200FT_CUDA(cudaMemcpyToSymbol(g_dcnt, z, sizeof(z)));
    return 0;
  }
  B200FT_CUDA(cudaDeviceSynchronize());
  B200FT_CUDA(cudaMemcpyFromSymbol(log_out, g_dlog, sizeof(long long) * 2 * 4096 * 2));
  B200FT_CUDA(cudaMemcpyFromSymbol(counts_out, g_dcnt, sizeof(int) * 2));
  return 0;
}
#endif
