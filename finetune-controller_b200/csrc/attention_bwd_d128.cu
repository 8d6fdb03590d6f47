// Causal flash-attention BACKWARD on tcgen05 for d_head = 128 (Llama-3-8B: BASELINE.json configs[3], [4]).
//
// At this head size a [128 x 128] bf16 tile is 32 KB and an fp32 accumulator tile 128 TMEM columns, so the d_head-64
// design (attention_bwd.cu: P and dS staged in shared memory, dQ partials beside S, dP, dK, dV in TMEM) does not
// fit.  Here P and dS never leave tensor memory: the softmax warps write them as bf16 over the S / dP columns they
// were computed from and the gradient UMMAs take them as the A operand straight from TMEM.  That fixes the
// orientation of every product — the A operand's rows are TMEM lanes — and gives two kernels of ONE template:
//
//   MODE 0  dK | dV   work item (kv tile j, batch, kv head): K_j, V_j resident; streams (Q_i, dO_i), i >= j, over the
//           q heads of the group.  S^T = K_j Q_i^T and dP^T = V_j dO_i^T (lanes = keys, columns = queries);
//           dV_j += P^T dO_i, dK_j += dS^T Q_i.  lse / D are per COLUMN: staged per pair in shared memory.
//   MODE 1  dQ        work item (q tile i, batch, head): Q_i, dO_i resident; streams (K_j, V_j), j <= i.
//           S = Q_i K_j^T, dP = dO_i V_j^T (lanes = queries); dQ_i += dS K_j.  lse / D are per thread.
//
// In both: product A operand = resident tile (K-major; MODE 1 keeps Q_i, dO_i in the TMEM columns of the second
// accumulator, which that kernel does not need, and reads them from there), B = streamed tile (K-major); gradient A = bf16 P / dS in
// TMEM, B = streamed tile read MN-major ([128 rows = reduction index] x [d_head], two 64-column boxes).  No atomics,
// no fp32 scratch: results are bit-reproducible.  TMEM: S 2 x 64 columns at 0 | dP 2 x 64 at 128 | acc0 256..383 |
// acc1 384..511; a thread of slice `part` owns columns 16*part..+15 of an S / dP buffer and writes its 16 bf16 results
// into the first 8 of them, so nobody overwrites scores another thread still has to read; the k-th gradient UMMA
// (16 reduction indices) reads its A operand at column 16*k of the buffer.
//
// 640 threads: warp 0 TMA producer, warp 1 UMMA issue, warp 2 TMEM allocation, warps 4-19 softmax / dS (four per
// TMEM lane quarter, one 16-column slice each).  Shared memory: resident pair 64 KB + two stream stages of 64 KB.
#include "attn_common.cuh"

namespace {
using namespace b200ft_attn;

constexpr int kDH = 128;
constexpr int kB = 128;                       // tile edge in queries and in keys
constexpr int kThreads = 640;
constexpr int kBox = kB * 64 * 2;             // 16 KB: one TMA box (128 rows x 64 columns, 128B swizzle)
constexpr int kTile = 2 * kBox;               // 32 KB: one Q / dO / K / V tile
constexpr int kRing = 2;
constexpr int kOffF = 0;                      // resident pair
constexpr int kOffS = 2 * kTile;              // ring stages of (tile0, tile1)
constexpr int kOffStat = kOffS + kRing * 2 * kTile;   // MODE 0: [2 buffers][lse | D][128] fp32
constexpr int kOffBar = kOffStat + 2 * 2 * kB * 4;
constexpr int kSmemBytes = kOffBar + 256;     // 194.25 KB
constexpr uint32_t kSub = 64;                 // streamed rows per sub-pair: S / dP are [128 x 64] and double-buffered
constexpr uint32_t kColS = 0, kColDP = 128, kColAcc0 = 256, kColAcc1 = 384;   // S, dP: two buffers of 64 columns each

#ifdef B200FT_ATTN_TIMING   // per-phase clock readings of CTA 0 (tools/attn_timing.py); never defined in the shipped build
__device__ long long g_blog[2][2][4096][2];   // [MODE][role]
__device__ int g_bcnt[2][2];
#define BLOG(role, tag) do { if (blockIdx.x == 0) { int i__ = g_bcnt[MODE][role]++; if (i__ < 4096) { g_blog[MODE][role][i__][0] = (tag); g_blog[MODE][role][i__][1] = clock64(); } } } while (0)
#else
#define BLOG(role, tag) do { } while (0)
#endif

struct Params {
  int batch, seq, n_head, n_kv_head;
  int q_off, k_off, v_off;
  int n_tiles;
  float scale, scale_log2;
  const float* lse2;
  const float* delta;
  __nv_bfloat16* dqkv; int64_t ld_d;
  const __nv_bfloat16* qkv; int64_t ld_qkv;
  const __nv_bfloat16* dout; int64_t ld_do;
};

// Work item of a CTA.  MODE 0: (kv tile ascending, batch, kv head); MODE 1: (q tile descending, batch, head):
// heaviest first.  Every warp role walks the same item sequence w = blockIdx.x, += gridDim.x.
struct Item {
  int fixed_tile, b, hx, hk, per_head, n_iter;
};
template <int MODE>
__device__ __forceinline__ Item decode_item(const Params& p, int w, int G) {
  Item it;
  const int heads = MODE == 0 ? p.n_kv_head : p.n_head;
  const int per_tile = p.batch * heads;
  const int tile_ord = w / per_tile, rest = w % per_tile;
  it.fixed_tile = MODE == 0 ? tile_ord : p.n_tiles - 1 - tile_ord;
  it.b = rest / heads;
  it.hx = rest % heads;                       // kv head (MODE 0) or q head (MODE 1)
  it.hk = MODE == 0 ? it.hx : it.hx / G;
  it.per_head = MODE == 0 ? p.n_tiles - it.fixed_tile : it.fixed_tile + 1;   // streamed tiles per head
  it.n_iter = MODE == 0 ? G * it.per_head : it.per_head;
  return it;
}

static __device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Persistent: one CTA per SM walks work items.  Barrier parities run on two counters shared by all roles:
// `g` = SUB-PAIRS processed so far by this CTA — a sub-pair is one half (64 rows) of a streamed tile pair against the
// resident pair; stream ring stage (g / 2) % 2, S / dP buffer g & 1 — and `n` = items processed so far (resident
// tiles, accumulators).  Working on 64 streamed rows at a time makes S and dP [128 x 64]: two buffers of each fit the
// 256 TMEM columns the accumulators leave, so the tensor core computes the products of sub-pair g+2 and the gradients
// of sub-pair g while the softmax warps exponentiate sub-pair g+1 (with full 128-row S / dP tiles, single-buffered,
// the two alternated: tensor pipe 25-32 % active).
// Tensor-core order: S,dP(0) S,dP(1) | grads(0) S,dP(2) | grads(1) S,dP(3) | ...  The products of g+2 overwrite the
// columns P(g), dS(g) live in; UMMAs execute in issue order, so they land after the gradient UMMAs have read them.
template <int MODE>
__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_d128_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                     const Params p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  if (raw & 1023u) __trap();  // dynamic shared memory starts 1024-aligned (no static smem in this kernel)
  const uint32_t sF0 = raw + kOffF, sF1 = sF0 + kTile;
  auto sS0 = [&](uint32_t st) { return raw + kOffS + st * 2u * kTile; };
  auto sS1 = [&](uint32_t st) { return raw + kOffS + st * 2u * kTile + kTile; };
  float* stat = reinterpret_cast<float*>(smem_raw + kOffStat);      // [buffer][lse | D][128]
  const uint32_t bars = raw + kOffBar;
  const uint32_t f_full = bars, f_empty = bars + 8u;
  auto st_full = [&](uint32_t s) { return bars + 8u * (2 + s); };
  auto st_empty = [&](uint32_t s) { return bars + 8u * (4 + s); };
  auto sdp_full = [&](uint32_t bf) { return bars + 8u * (6 + bf); };
  // one hand-off barrier per S / dP buffer: a softmax warp may run a sub-pair ahead of the slowest one (the products are
  // issued two sub-pairs ahead), and on a single barrier its arrival for g+1 would be counted towards phase g
  auto pds_full = [&](uint32_t bf) { return bars + 8u * (8 + bf); };
  const uint32_t acc_full = bars + 8u * 10, acc_free = bars + 8u * 11;
  const uint32_t tmem_slot = bars + 8u * 12;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = p.n_head / p.n_kv_head;
  const int num_items = p.n_tiles * p.batch * (MODE == 0 ? p.n_kv_head : p.n_head);

  if (warp == 1 && lane == 0) {
    mbar_init(f_full, MODE == 1 ? 16 : 1); mbar_init(f_empty, 1);   // MODE 1: the sixteen softmax warps fill the resident operands
    for (uint32_t s = 0; s < kRing; ++s) { mbar_init(st_full(s), 1); mbar_init(st_empty(s), 1); }
    mbar_init(sdp_full(0), 1); mbar_init(sdp_full(1), 1); mbar_init(pds_full(0), 8); mbar_init(pds_full(1), 8);
    mbar_init(acc_full, 1); mbar_init(acc_free, 16);
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
  // (The plain round-robin b, b+grid, ... gave the first CTAs 45 tile pairs and the last 14 at 1 x 4096 x 32/8: 63 %.)
  auto item_at = [&](uint32_t round) -> int {
    const int i = (round & 1u) ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x;
    return (int)(round * gridDim.x) + i;
  };

  // a [128 x 128] tile = two boxes of 64 columns
  auto load_tile = [&](uint32_t dst, const CUtensorMap* map, int col, int row, int b, uint32_t bar) {
    tma_load_3d(dst, map, col, row, b, bar);
    tma_load_3d(dst + kBox, map, col + 64, row, b, bar);
  };

  if (warp == 0) {
    // ================================ TMA producer ================================
    const bool issuer = elect_one();
    uint32_t g = 0, n = 0;
    for (int w = item_at(0); w < num_items; w = item_at(++n)) {
      const Item it = decode_item<MODE>(p, w, G);
      if (MODE == 0) {   // (MODE 1 keeps its resident Q_i, dO_i in TMEM: written by the softmax warps)
        if (n > 0) mbar_wait(f_empty, (n - 1) & 1u);   // the previous item's last dP UMMAs have read the resident pair
        if (issuer) {
          mbar_expect_tx(f_full, 2 * kTile);
          load_tile(sF0, &tm_qkv, p.k_off + it.hk * kDH, it.fixed_tile * kB, it.b, f_full);
          load_tile(sF1, &tm_qkv, p.v_off + it.hk * kDH, it.fixed_tile * kB, it.b, f_full);
        }
        __syncwarp();
      }
      for (int t = 0; t < it.n_iter; ++t, ++g) {
        const uint32_t st = g % kRing;
        mbar_wait(st_empty(st), ((g / kRing) & 1u) ^ 1u);
        if (issuer) {
          mbar_expect_tx(st_full(st), 2 * kTile);
          if (MODE == 0) {
            const int hq = it.hk * G + t / it.per_head, qt = it.fixed_tile + t % it.per_head;
            load_tile(sS0(st), &tm_qkv, p.q_off + hq * kDH, qt * kB, it.b, st_full(st));
            load_tile(sS1(st), &tm_do, hq * kDH, qt * kB, it.b, st_full(st));
          } else {
            load_tile(sS0(st), &tm_qkv, p.k_off + it.hk * kDH, t * kB, it.b, st_full(st));
            load_tile(sS1(st), &tm_qkv, p.v_off + it.hk * kDH, t * kB, it.b, st_full(st));
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ================================ UMMA issuer ================================
    // Sub-pair g = 64 streamed rows (half h = g & 1 of a streamed tile pair); its products go to buffer g & 1.
    const bool issuer = elect_one();
    const uint32_t id_kk = idesc(kSub, 0, 0);  // S, dP: 128 x 64 x 128, both operands K-major (d_head contiguous)
    const uint32_t id_g = idesc(kDH, 0, 1);    // gradients: 128 x 128 (d_head) x 64, A from TMEM, B MN-major
    // Shared-memory descriptors are built ONCE per tile; a UMMA's descriptor is that base plus a compile-time offset in
    // the 16-byte units of its address field (the tiles lie below 256 KB, so the add never carries out of the field).
    // The single issuing thread is on the critical path of every sub-pair: 24 UMMAs with two freshly assembled
    // descriptors each measured ~3100 cycles per sub-pair with the tensor pipe 32 % active.
    // K-major operand, k-th step of 16 d_head elements: box k/4, 32 bytes per step inside the 128-byte swizzle row;
    // MN-major B operand, k-th step of 16 rows (reduction index): 2 KB per step, the second 64 columns one box on.
    const uint64_t kF0 = sdesc(sF0, 16, 1024), kF1 = sdesc(sF1, 16, 1024);
    uint64_t kS0[kRing], kS1[kRing], mS0[kRing], mS1[kRing];
#pragma unroll
    for (uint32_t st = 0; st < kRing; ++st) {
      kS0[st] = sdesc(sS0(st), 16, 1024); kS1[st] = sdesc(sS1(st), 16, 1024);
      mS0[st] = sdesc(sS0(st), kBox, 1024); mS1[st] = sdesc(sS1(st), kBox, 1024);
    }
    auto koff = [](int k) { return (uint64_t)((((uint32_t)(k >> 2) * kBox) + 32u * (uint32_t)(k & 3)) >> 4); };
    auto moff = [](int k) { return (uint64_t)((2048u * (uint32_t)k) >> 4); };
    auto issue_products = [&](uint32_t gg, bool last_of_item) {
      const uint32_t pr = gg >> 1, h = gg & 1u, st = pr % kRing, bf = gg & 1u;
      mbar_wait(st_full(st), (pr / kRing) & 1u);   // (passes at once for the second half of a tile pair)
      tc_fence_after();
      // Buffer bf was last used by sub-pair gg-2: its softmax is over (pds_full was waited for before its gradient
      // UMMAs were issued) and those UMMAs, which read P / dS from it, precede these in issue order.
      const uint64_t rows = (uint64_t)(h * ((kSub * 128u) >> 4));   // 64 rows further into each box of the streamed tiles
      const uint64_t b0 = (st ? kS0[1] : kS0[0]) + rows, b1 = (st ? kS1[1] : kS1[0]) + rows;
      if (issuer) {
        if (MODE == 0) {
#pragma unroll
          for (int k = 0; k < kDH / 16; ++k) umma(tmem + kColS + kSub * bf, kF0 + koff(k), b0 + koff(k), id_kk, k > 0);
#pragma unroll
          for (int k = 0; k < kDH / 16; ++k) umma(tmem + kColDP + kSub * bf, kF1 + koff(k), b1 + koff(k), id_kk, k > 0);
        } else {
          // A operands Q_i, dO_i from TMEM (the second accumulator's columns, free in this kernel): a [128 x 64 x 16]
          // UMMA with both operands in shared memory reads 6 KB per 32 cycles of tensor time — 1.5 x what an SM's
          // shared memory delivers (measured: 75 cycles per UMMA); with A in TMEM it reads 2 KB
#pragma unroll
          for (int k = 0; k < kDH / 16; ++k) umma_ts(tmem + kColS + kSub * bf, tmem + kColAcc1 + 8u * (uint32_t)k, b0 + koff(k), id_kk, k > 0);
#pragma unroll
          for (int k = 0; k < kDH / 16; ++k) umma_ts(tmem + kColDP + kSub * bf, tmem + kColAcc1 + 64u + 8u * (uint32_t)k, b1 + koff(k), id_kk, k > 0);
        }
        umma_commit(sdp_full(bf));
        if (last_of_item) umma_commit(f_empty);   // nothing of this item reads the resident pair any more
      }
      __syncwarp();
    };
    auto issue_grads = [&](uint32_t gg, bool first) {
      const uint32_t pr = gg >> 1, h = gg & 1u, st = pr % kRing, bf = gg & 1u;
      const uint64_t half = (uint64_t)h * moff(4);
      const uint64_t m0 = (st ? mS0[1] : mS0[0]) + half, m1 = (st ? mS1[1] : mS1[0]) + half;
      if (issuer) {
        // a thread packed its 32 columns into the first 16 of them: reduction step k (16 rows) reads its A operand at
        // column 32*(k/2) + 8*(k%2) of the buffer
        if (MODE == 0) {
#pragma unroll
          for (int k = 0; k < (int)kSub / 16; ++k)   // dV += P^T dO: 16 queries per step
            umma_ts(tmem + kColAcc0, tmem + kColS + kSub * bf + 32u * (uint32_t)(k >> 1) + 8u * (uint32_t)(k & 1), m1 + moff(k), id_g, !(first && k == 0));
#pragma unroll
          for (int k = 0; k < (int)kSub / 16; ++k)   // dK += dS^T Q
            umma_ts(tmem + kColAcc1, tmem + kColDP + kSub * bf + 32u * (uint32_t)(k >> 1) + 8u * (uint32_t)(k & 1), m0 + moff(k), id_g, !(first && k == 0));
        } else {
#pragma unroll
          for (int k = 0; k < (int)kSub / 16; ++k)   // dQ += dS K: 16 keys per step
            umma_ts(tmem + kColAcc0, tmem + kColDP + kSub * bf + 32u * (uint32_t)(k >> 1) + 8u * (uint32_t)(k & 1), m0 + moff(k), id_g, !(first && k == 0));
        }
        if (h == 1u) umma_commit(st_empty(st));   // both halves consumed: the streamed pair's slot may be refilled
      }
      __syncwarp();
    };
    uint32_t g = 0, n = 0;
    for (int w = item_at(0); w < num_items; w = item_at(++n)) {
      const Item it = decode_item<MODE>(p, w, G);
      const int N2 = 2 * it.n_iter;           // sub-pairs of this item (>= 2)
      mbar_wait(f_full, n & 1u);
      issue_products(g, false);
      issue_products(g + 1, N2 == 2);
      for (int t2 = 0; t2 < N2; ++t2, ++g) {
        if (issuer) BLOG(1, 10);
        mbar_wait(pds_full(g & 1u), (g >> 1) & 1u);   // softmax warps are done with sub-pair g: P, dS are in TMEM
        tc_fence_after();
        if (t2 == 0 && n > 0) { mbar_wait(acc_free, (n - 1) & 1u); tc_fence_after(); }  // previous accumulators read out
        if (issuer) BLOG(1, 11);
        issue_grads(g, t2 == 0);
        if (issuer) BLOG(1, 12);
        // the products of sub-pair g+2 replace P(g), dS(g) — in issue order after the UMMAs that read them — and run,
        // with the gradients of g, while the softmax warps work on sub-pair g+1
        if (t2 + 2 < N2) issue_products(g + 2, t2 + 3 == N2);
        if (issuer) BLOG(1, 13);
      }
      if (issuer) umma_commit(acc_full);
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ================================ softmax / dS ================================
    // Two COHORTS of eight warps: cohort c works on the sub-pairs with g & 1 == c, i.e. on S / dP buffer c, always the
    // same half (rows 64c .. 64c+63) of every streamed tile pair.  tcgen05.ld moves 64 bytes per clock per SM, so reading
    // the [128 x 64] fp32 S and dP of a sub-pair costs 1024 cycles — as long as its UMMAs.  With all sixteen warps on the
    // same sub-pair they read together and then compute together (2140 cycles per sub-pair measured); staggered, one
    // cohort's reads run under the other's exponentials.  Inside a cohort: two warps per TMEM lane quarter; a thread
    // owns one lane L (a key in MODE 0, a query in MODE 1) and the 32 columns 32*part .. +31 of its buffer.
    const int idx = warp - 4;
    const int cohort = idx >> 3;
    const int quarter = warp & 3;
    const int part = (idx >> 2) & 1;
    const int part4 = idx >> 2;                            // 0..3: the accumulator slice this warp writes out per item
    const int L = quarter * 32 + lane;
    const int co_tid = (idx & 7) * 32 + lane;              // 0..255 inside the cohort
    const uint32_t t_lane = tmem + ((uint32_t)(quarter * 32) << 16);
    const uint32_t bf = (uint32_t)cohort;
    const uint32_t t_s = t_lane + kColS + kSub * bf + 32u * (uint32_t)part;
    const uint32_t t_dp = t_lane + kColDP + kSub * bf + 32u * (uint32_t)part;
    const f32x2_t sc2 = pk2(p.scale_log2, p.scale_log2);
    uint32_t gp = 0, n = 0;                                // gp: tile PAIRS done so far (this cohort's sub-pair is 2*gp + cohort)
    for (int w = item_at(0); w < num_items; w = item_at(++n)) {
      const Item it = decode_item<MODE>(p, w, G);
      // log-sum-exp and D of a query row; rows past the end of the sequence get lse = +inf: P = dS = 0
      auto row_stats = [&](int hq, int qt_, int r, float& lse_o, float& dl_o) {
        const int qr = qt_ * kB + r;
        lse_o = INFINITY; dl_o = 0.f;
        if (qr < p.seq) {
          const int64_t idx2 = ((int64_t)it.b * p.n_head + hq) * p.seq + qr;
          lse_o = __ldg(p.lse2 + idx2); dl_o = __ldg(p.delta + idx2);
        }
      };
      // MODE 0: the statistics belong to the COLUMNS and change with every streamed tile pair.  Threads 0..63 of a
      // cohort fetch the values of ITS 64 query rows one pair ahead (global-memory latency off the critical path) and
      // publish them in shared memory, double-buffered, when the pair starts (one 256-thread named barrier per cohort).
      float lse = INFINITY, dl = 0.f, lse_next = INFINITY, dl_next = 0.f;
      // MODE 1, resident operands -> TMEM: this thread's lane is query row L; slices 0, 1 carry the two halves of its Q
      // row, slices 2, 3 those of its dO row (64 elements = 32 packed columns each), straight from global memory.  Item
      // n + 1's operands are written at the END of item n — as soon as its last products have read theirs, two
      // sub-pairs before its gradients finish — so the tensor core goes from item to item without waiting for them.
      auto fill_operands = [&](const Item& nx, uint32_t n_of) {
        const int qr = nx.fixed_tile * kB + L;
        uint32_t a[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) a[i] = 0u;
        if (qr < p.seq) {
          const __nv_bfloat16* src = part4 < 2
              ? p.qkv + ((int64_t)nx.b * p.seq + qr) * p.ld_qkv + p.q_off + nx.hx * kDH + 64 * part4
              : p.dout + ((int64_t)nx.b * p.seq + qr) * p.ld_do + nx.hx * kDH + 64 * (part4 - 2);
#pragma unroll
          for (int v = 0; v < 8; ++v) {
            const uint4 q4 = __ldg(reinterpret_cast<const uint4*>(src) + v);
            a[4 * v] = q4.x; a[4 * v + 1] = q4.y; a[4 * v + 2] = q4.z; a[4 * v + 3] = q4.w;
          }
        }
        if (n_of > 0) { mbar_wait(f_empty, (n_of - 1) & 1u); tc_fence_after(); }   // the previous item's products have read theirs
        tmem_st16(t_lane + kColAcc1 + 32u * (uint32_t)part4, a);
        tmem_st16(t_lane + kColAcc1 + 32u * (uint32_t)part4 + 16u, a + 16);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(f_full);
      };
      if (MODE == 1) {
        if (n == 0) fill_operands(it, 0);
        row_stats(it.hx, it.fixed_tile, L, lse, dl);
      }
      else if (co_tid < 64) row_stats(it.hk * G, it.fixed_tile, 64 * cohort + co_tid, lse_next, dl_next);
      int tq = 0, th = 0;   // MODE 0: t % per_head, t / per_head as counters (a division per pair in each of 512 threads otherwise)
      for (int t = 0; t < it.n_iter; ++t, ++gp) {
        int qt, kt;
        if (MODE == 0) { qt = it.fixed_tile + tq; kt = it.fixed_tile; }
        else           { qt = it.fixed_tile; kt = t; }
        int tqn = tq + 1, thn = th;
        if (tqn == it.per_head) { tqn = 0; ++thn; }
        const bool diag = qt == kt;
        float* st_lse = stat + (gp & 1u) * 2 * kB + 64 * cohort;     // this cohort's 64 columns
        float* st_dl = st_lse + kB;
        if (MODE == 0) {
          if (co_tid < 64) {
            st_lse[co_tid] = lse_next;
            st_dl[co_tid] = dl_next;
            if (t + 1 < it.n_iter)
              row_stats(it.hk * G + thn, it.fixed_tile + tqn, 64 * cohort + co_tid, lse_next, dl_next);
          }
          named_bar_sync(1 + cohort, 256);   // the cohort's eight warps: this pair's statistics are visible
        }
        const int c0 = 64 * cohort + 32 * part;          // first of this thread's columns, as a row of the streamed tile
        // ---- phase A: P = exp2(S*c - lse) for this thread's 32 columns (kept in registers for phase B) ----
        f32x2_t P2[16];
        const bool tl = (warp == 4 && lane == 0);
        if (tl) BLOG(0, 0);
        mbar_wait(sdp_full(bf), gp & 1u);
        tc_fence_after();
        if (tl) BLOG(0, 1);
        {
          uint32_t rs[32];
          tmem_ld32(t_s, rs);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const int c = c0 + i;
            f32x2_t nl2;
            if (MODE == 0) { const float2 a = *reinterpret_cast<const float2*>(st_lse + 32 * part + i); nl2 = pk2(-a.x, -a.y); }
            else nl2 = pk2(-lse, -lse);
            float e0, e1;
            upk2(fma2(pk2(__uint_as_float(rs[i]), __uint_as_float(rs[i + 1])), sc2, nl2), e0, e1);
            float p0 = ex2(e0), p1 = ex2(e1);   // lse = +inf (row past the sequence): exp2(-inf) = 0
            if (diag) {                        // warp-uniform; causal: key index > query index contributes nothing
              const bool m0 = MODE == 0 ? (L > c) : (c > L);
              const bool m1 = MODE == 0 ? (L > c + 1) : (c + 1 > L);
              if (m0) p0 = 0.f;
              if (m1) p1 = 0.f;
            }
            P2[i >> 1] = pk2(p0, p1);
          }
        }
        if (tl) BLOG(0, 2);
        if (MODE == 0) {   // P^T as bf16 over this thread's own first 16 score columns: A operand of dV += P^T dO
          uint32_t wp[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) wp[i] = f32x2_to_bf16x2(P2[i]);
          tmem_st16(t_s, wp);
        }
        // ---- phase B: dS' = P * (dP - D) (the softmax scale multiplies the dK / dQ accumulators once, at the end) ----
        {
          uint32_t rd[32];
          tmem_ld32(t_dp, rd);
          tmem_wait_ld();
          uint32_t wd[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            f32x2_t dl2;
            if (MODE == 0) { const float2 a = *reinterpret_cast<const float2*>(st_dl + 32 * part + 2 * i); dl2 = pk2(a.x, a.y); }
            else dl2 = pk2(dl, dl);
            wd[i] = f32x2_to_bf16x2(mul2(P2[i], sub2(pk2(__uint_as_float(rd[2 * i]), __uint_as_float(rd[2 * i + 1])), dl2)));
          }
          tmem_st16(t_dp, wd);
        }
        if (tl) BLOG(0, 3);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(pds_full(bf));
        if (tl) BLOG(0, 4);
        tq = tqn; th = thn;
      }
      if (MODE == 1) {
        const int w_next = item_at(n + 1);
        if (w_next < num_items) fill_operands(decode_item<MODE>(p, w_next, G), n + 1);
      }
      // ---- accumulators -> (x softmax scale for dK, dQ) -> bf16 -> the gradient buffer ----
      mbar_wait(acc_full, n & 1u);
      tc_fence_after();
      const int orow = it.fixed_tile * kB + L;   // key row (MODE 0) or query row (MODE 1)
      if (MODE == 0) {
        // slices 0, 1 write the two 64-column halves of dV (accumulator 0), slices 2, 3 those of dK (accumulator 1)
        const uint32_t t_acc = t_lane + ((part4 >> 1) == 0 ? kColAcc0 : kColAcc1) + 64u * (uint32_t)(part4 & 1);
        const int col = ((part4 >> 1) == 0 ? p.v_off : p.k_off) + it.hk * kDH + 64 * (part4 & 1);
        const float oscale = (part4 >> 1) == 0 ? 1.f : p.scale;
        uint32_t r[64];
        tmem_ld32(t_acc, r);
        tmem_ld32(t_acc + 32, r + 32);
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_free);    // the next item's gradient UMMAs may overwrite the accumulators
        if (orow < p.seq) {
          __nv_bfloat16* dst = p.dqkv + ((int64_t)it.b * p.seq + orow) * p.ld_d + col;
#pragma unroll
          for (int v = 0; v < 8; ++v) {
            float fv[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) fv[i] = __uint_as_float(r[8 * v + i]) * oscale;
            *reinterpret_cast<uint4*>(dst + 8 * v) = pack8(fv);
          }
        }
      } else {
        // dQ: the four slices write 32 columns each
        uint32_t r[32];
        tmem_ld32(t_lane + kColAcc0 + 32u * (uint32_t)part4, r);
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_free);
        if (orow < p.seq) {
          __nv_bfloat16* dst = p.dqkv + ((int64_t)it.b * p.seq + orow) * p.ld_d + p.q_off + it.hx * kDH + 32 * part4;
#pragma unroll
          for (int v = 0; v < 4; ++v) {
            float fv[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) fv[i] = __uint_as_float(r[8 * v + i]) * p.scale;
            *reinterpret_cast<uint4*>(dst + 8 * v) = pack8(fv);
          }
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

}  // namespace

int b200ft_attn::launch_bwd_d128(const CUtensorMap& tm_qkv, const CUtensorMap& tm_do, const Bwd128Args& a,
                                 cudaStream_t st) {
  Params p;
  p.batch = a.batch; p.seq = a.seq; p.n_head = a.n_head; p.n_kv_head = a.n_kv_head;
  p.q_off = a.q_off; p.k_off = a.k_off; p.v_off = a.v_off;
  p.n_tiles = (a.seq + kB - 1) / kB;
  p.scale = a.scale; p.scale_log2 = a.scale * 1.4426950408889634f;
  p.lse2 = a.lse2; p.delta = a.delta; p.dqkv = a.dqkv; p.ld_d = a.ld_d;
  p.qkv = a.qkv; p.ld_qkv = a.ld_qkv; p.dout = a.dout; p.ld_do = a.ld_dout;
  static thread_local int attr_dev = -1;
  int dev = -1;
  B200FT_CUDA(cudaGetDevice(&dev));
  if (dev != attr_dev) {
    B200FT_CUDA(cudaFuncSetAttribute(attn_bwd_d128_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    B200FT_CUDA(cudaFuncSetAttribute(attn_bwd_d128_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_dev = dev;
  }
  // persistent: one CTA per SM walks the work items (heaviest first, round-robin)
  const int items0 = p.n_tiles * a.batch * a.n_kv_head, items1 = p.n_tiles * a.batch * a.n_head;
  B200FT_CUDA(launch_k(attn_bwd_d128_kernel<0>, (unsigned)(items0 < B200FT_NUM_SMS ? items0 : B200FT_NUM_SMS), kThreads,
                       kSmemBytes, st, tm_qkv, tm_do, p));
  B200FT_CUDA(launch_k(attn_bwd_d128_kernel<1>, (unsigned)(items1 < B200FT_NUM_SMS ? items1 : B200FT_NUM_SMS), kThreads,
                       kSmemBytes, st, tm_qkv, tm_do, p));
  return 0;
}

#ifdef B200FT_ATTN_TIMING
extern "C" int b200ft_debug_attn_bwd128_timing(long long* log_out, int* counts_out, int reset) {
  if (reset) {
    int z[4] = {0, 0, 0, 0};
    B200FT_CUDA(cudaMemcpyToSymbol(g_bcnt, z, sizeof(z)));
    return 0;
  }
  B200FT_CUDA(cudaDeviceSynchronize());
  B200FT_CUDA(cudaMemcpyFromSymbol(log_out, g_blog, sizeof(long long) * 2 * 2 * 4096 * 2));
  B200FT_CUDA(cudaMemcpyFromSymbol(counts_out, g_bcnt, sizeof(int) * 4));
  return 0;
}
#endif
