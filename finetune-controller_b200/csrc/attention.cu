// Causal flash-attention forward on tcgen05 for d_head = 64 (GPT-2, Llama-3.2-1B), reading q, k, v
// straight out of the fused [tokens, q|k|v] projection buffer (grouped-query attention = several
// q heads sharing one kv head) and writing o [tokens, heads*64] plus the per-row log-sum-exp.
//
// Work item = (batch, head, PAIR of 128-row query tiles A and B), heaviest pairs first; persistent CTAs (one per
// SM) walk the items; 384 threads:
//   warp 0    TMA producer: Q_A, Q_B once, then K_j / V_j into 2-stage rings — loaded once, used by both tiles
//   warp 1    MMA issuer:   S_g = Q_g K_j^T (128x128x64, 4 UMMAs) and PV_g = P_g V_j (128x64x128, 8 UMMAs),
//             software-pipelined  PV_A(j), S_A(j+1), PV_B(j), S_B(j+1)  so that while one tile's
//             softmax runs the tensor core works for the other (ping-pong)
//   warp 2    TMEM allocator (512 columns: S_A, S_B 128 each; PV_A, PV_B 64 each; P_A, P_B 64 each)
//   warps 4-7 / 8-11  softmax of tile A / B, one thread per query row: ONE sweep over S in TMEM —
//             p = exp2(s*scale - m_ref) against the running reference max, row sum, bf16 P written back to
//             TMEM with tcgen05.st (the PV UMMA takes its A operand from TMEM: no smem round trip for P,
//             whose 6 KB of reads per 32-cycle UMMA would exceed the shared-memory port) — repeated only if
//             some row's max grew by more than 2^8 over its reference (lazy rescaling; exact after the final
//             normalisation).  The output tile O accumulates in TMEM over the kv tiles (the P V UMMAs add into it);
//             the softmax warps rescale it in place only when a reference max moves, and read it once at the end.
// V_j is consumed as an MN-major B operand ([keys, d_head] rows as TMA delivers them): nothing is
// transposed.  Registers are the per-scheduler resource that decides the shape: 12 warps x 168.
//
// Work per (batch, head): T^2/2 * 64 * 4 FLOP (causal); per 128x128 tile the tensor pipe needs 512
// cycles, the special-function unit 1024 (exp2 at 16/clk/SM): the kernel is bound by MUFU.
#include <stdlib.h>
#include "attn_common.cuh"

namespace {
using namespace b200ft_attn;

constexpr int kBQ = 128, kBK = 128;
constexpr int kThreads = 640;                  // warps 0-2 TMA / UMMA / TMEM, 4-19 softmax (two threads per query row)
constexpr int kStages = 2;
constexpr int kBoxBytes = 128 * 64 * 2;         // one TMA box: 128 rows x 64 columns, 128B-swizzled = 16 KB
constexpr int kTmemCols = 512;
// Per head size.  A [128 x DH] tile is DH/64 boxes side by side.  TMEM: S_g at 128*g; O_g at 256 + DH*g; P_g (bf16
// pairs, 64 columns) in its own columns for DH = 64 and ON TOP OF S_g for DH = 128, where S (256) + O (256) fill
// the 512 columns: the exponentials then overwrite the scores they came from, the row maximum is found in a
// read-only pass first, and the issue order per query tile becomes P V (j) before S (j+1).
#ifndef B200FT_ATTN_POLY_MASK
#define B200FT_ATTN_POLY_MASK 3   // d_head 128: pairs whose index has these bits set take exp2 on the FMA pipe (3: one in four)
#endif
template <int DH> struct Cfg {
  static_assert(DH == 64 || DH == 128, "head sizes built: 64, 128");
  static constexpr int kBoxes = DH / 64;
  static constexpr int kQBytes = kBQ * DH * 2;          // 16 / 32 KB per query tile
  static constexpr int kKVBytes = kBK * DH * 2;
  static constexpr int kTiles = 2 * kQBytes + 2 * kStages * kKVBytes;   // 96 / 192 KB
  static constexpr int kOffXch = kTiles + 256;          // after the barriers: row statistics exchanged between the two
                                                        // threads of a query row, [slot][group][half][128] fp32
  static constexpr int kSmemBytes = kOffXch + 2 * 2 * 2 * 128 * 4;   // no alignment slack (the kernel checks the base)
  static constexpr bool kAliasP = DH == 128;
  // exp2 on the FMA pipe for part of the row.  Measured (us, 25 % of the pairs): 1 x 4096 x 32/8 d 128 149.6 -> 139.8, 2 x 4096
  // 280.8 -> 273.7; d_head 64, where a tile has half the tensor work to hide the same exponentials and the softmax warps
  // are short of issue slots: 4 x 2048 x 32/8 136 -> 140, 8 x 4096 x 12 331 -> 352.  So: d_head 128 only.
  static constexpr int kPolyMask = DH == 128 ? B200FT_ATTN_POLY_MASK : 0xff;
  static constexpr uint32_t kColS = 0, kColPV = 256;
  static constexpr uint32_t kColP = kAliasP ? 0u : 384u;
  static constexpr uint32_t kStrideP = kAliasP ? 128u : 64u;   // column distance between P_A and P_B
  // P holds 128 keys as 64 packed columns.  In its own region they are contiguous (keys 64h.. of half h at column 32h).
  // Written over S, half h of a row (the thread that owns score columns 64h .. 64h+63) keeps its 32 packed columns
  // inside its own score columns, at 64h, so that no thread overwrites scores its partner has yet to read.
  static constexpr uint32_t kHalfP = kAliasP ? 64u : 32u;      // column distance between the two halves of a P row
};
constexpr float kRescaleThreshold = 8.0f;

#ifdef B200FT_ATTN_TIMING   // per-phase clock readings of CTA 0 (tools/attn_timing.py); never defined in the shipped build
__device__ long long g_tlog[2][4096][2];
__device__ int g_tcnt[2];
#define TLOG(role, tag) do { if (blockIdx.x == 0) { int i__ = g_tcnt[role]++; if (i__ < 4096) { g_tlog[role][i__][0] = (tag); g_tlog[role][i__][1] = clock64(); } } } while (0)
#else
#define TLOG(role, tag) do { } while (0)
#endif       // log2 units: p stays below 2^8

struct AttnParams {
  int batch, seq, n_head, n_kv_head;
  int q_off, k_off, v_off;  // column offsets of head 0 of q / k / v in the fused buffer
  int nq_tiles, nq_pairs;
  float scale_log2;         // softmax scale * log2(e)
  __nv_bfloat16* out; int64_t ld_out;
  float* lse2;              // [batch*n_head, seq]: reference max + log2(row sum), log2 domain incl. scale
};



// Work item: (pair of query tiles, batch, head), heaviest pairs (most kv tiles) first.
struct FwdItem {
  int b, h, hk, qtA, qtB, nA, nB, n_kv;
  bool hasB;
};
__device__ __forceinline__ FwdItem decode_item(const AttnParams& p, int w) {
  FwdItem it;
  const int bh_count = p.batch * p.n_head;
  const int pair = p.nq_pairs - 1 - w / bh_count;
  const int bh = w % bh_count;
  it.b = bh / p.n_head; it.h = bh % p.n_head;
  it.hk = it.h / (p.n_head / p.n_kv_head);
  it.qtA = 2 * pair; it.qtB = 2 * pair + 1;
  it.hasB = it.qtB < p.nq_tiles;
  it.nA = it.qtA + 1; it.nB = it.hasB ? it.qtB + 1 : 0;   // causal: tile g attends kv tiles 0 .. qt_g
  it.n_kv = it.nA > it.nB ? it.nA : it.nB;
  return it;
}

// Persistent: one CTA per SM walks the work items w = blockIdx.x, += gridDim.x.  Barrier parities run on counters
// every role advances identically: kc = kv tiles streamed so far, cA / cB = tiles each query group has processed,
// n = items.  The producer requests the next item's query tiles as soon as the last UMMAs of the current one retire
// and keeps the K / V ring running across the boundary, so an item's prologue hides behind its predecessor's tail.
template <int DH>
__global__ void __launch_bounds__(kThreads, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tm, const AttnParams p) {
  using C = Cfg<DH>;
  constexpr int kDH = DH, kQBytes = C::kQBytes, kKVBytes = C::kKVBytes, kTiles = C::kTiles;
  constexpr uint32_t kColS = C::kColS, kColPV = C::kColPV, kColP = C::kColP;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  if (raw & 1023u) __trap();  // dynamic shared memory starts 1024-aligned (no static smem in this kernel)
  const uint32_t sQ = raw;                               // Q_A, Q_B
  const uint32_t sK = sQ + 2 * kQBytes;                  // K ring
  const uint32_t sV = sK + kStages * kKVBytes;           // V ring
  const uint32_t bars = raw + kTiles;
  const uint32_t q_full = bars;
  auto k_full = [&](int s) { return bars + 8u * (1 + s); };
  auto k_empty = [&](int s) { return bars + 8u * (3 + s); };
  auto v_full = [&](int s) { return bars + 8u * (5 + s); };
  auto v_empty = [&](int s) { return bars + 8u * (7 + s); };
  auto s_full = [&](int g) { return bars + 8u * (9 + g); };
  auto s_free = [&](int g) { return bars + 8u * (11 + g); };
  auto p_full = [&](int g) { return bars + 8u * (13 + g); };
  auto pv_full = [&](int g) { return bars + 8u * (15 + g); };
  const uint32_t q_empty = bars + 8u * 17;
  const uint32_t tmem_slot = bars + 8u * 18;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_items = p.nq_pairs * p.batch * p.n_head;

  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1); mbar_init(q_empty, 1);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(k_full(s), 1); mbar_init(k_empty(s), 1); mbar_init(v_full(s), 1); mbar_init(v_empty(s), 1);
    }
    for (int g = 0; g < 2; ++g) {
      mbar_init(s_full(g), 1); mbar_init(s_free(g), 8); mbar_init(p_full(g), 8); mbar_init(pv_full(g), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_ptr;
  grid_dep_sync();  // the prologue above overlaps the previous kernel's tail; global memory only from here on

  // Persistent launches deal the items (heaviest first) out in a snake: round n hands item n*grid + b to CTA b when n
  // is even and n*grid + (grid-1-b) when n is odd; with one CTA per item (grid = items) only round 0 exists.
  auto item_at = [&](uint32_t round) -> int {
    const int i = (round & 1u) ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x;
    return (int)(round * gridDim.x) + i;
  };

  if (warp == 0) {
    // ================================ TMA producer ================================
    const bool issuer = elect_one();
    uint32_t kc = 0, n = 0;
    for (int w = item_at(0); w < num_items; w = item_at(++n)) {
    const FwdItem it = decode_item(p, w);
    const int b = it.b, h = it.h, hk = it.hk;
    if (n > 0) mbar_wait(q_empty, (n - 1) & 1u);   // the previous item's UMMAs are done with the query tiles
    if (issuer) {
      mbar_expect_tx(q_full, it.hasB ? 2 * kQBytes : kQBytes);
#pragma unroll
      for (int bx = 0; bx < C::kBoxes; ++bx) {
        tma_load_3d(sQ + bx * kBoxBytes, &tm, p.q_off + h * kDH + 64 * bx, it.qtA * kBQ, b, q_full);
        if (it.hasB) tma_load_3d(sQ + kQBytes + bx * kBoxBytes, &tm, p.q_off + h * kDH + 64 * bx, it.qtB * kBQ, b, q_full);
      }
    }
    __syncwarp();
    for (int j = 0; j < it.n_kv; ++j, ++kc) {
      const int st = (int)(kc & 1u);
      const uint32_t ph = (kc >> 1) & 1u;
      mbar_wait(k_empty(st), ph ^ 1u);
      if (issuer) {
        mbar_expect_tx(k_full(st), kKVBytes);
#pragma unroll
        for (int bx = 0; bx < C::kBoxes; ++bx)
          tma_load_3d(sK + st * kKVBytes + bx * kBoxBytes, &tm, p.k_off + hk * kDH + 64 * bx, j * kBK, b, k_full(st));
      }
      __syncwarp();
      mbar_wait(v_empty(st), ph ^ 1u);
      if (issuer) {
        mbar_expect_tx(v_full(st), kKVBytes);
#pragma unroll
        for (int bx = 0; bx < C::kBoxes; ++bx)
          tma_load_3d(sV + st * kKVBytes + bx * kBoxBytes, &tm, p.v_off + hk * kDH + 64 * bx, j * kBK, b, v_full(st));
      }
      __syncwarp();
    }
    }  // items
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    const bool issuer = elect_one();
    const uint32_t id_s = idesc(kBK, 0, 0);   // S: N = 128 keys, K-major B (K_j rows = keys, d_head contiguous)
    const uint32_t id_pv = idesc(kDH, 0, 1);  // PV: N = d_head, MN-major B (V_j rows = keys = reduction index)
    // Shared-memory descriptors are built once per tile; a UMMA's descriptor is that base plus a compile-time offset in
    // the 16-byte units of its address field (all tiles lie below 256 KB: the add never carries out of the field) —
    // the single issuing thread sits on the critical path of every kv tile.
    const uint64_t dQ[2] = {sdesc(sQ, 16, 1024), sdesc(sQ + kQBytes, 16, 1024)};
    const uint64_t dK[2] = {sdesc(sK, 16, 1024), sdesc(sK + kKVBytes, 16, 1024)};
    const uint64_t dV[2] = {sdesc(sV, kBoxBytes, 1024), sdesc(sV + kKVBytes, kBoxBytes, 1024)};
    auto koff = [](int k) { return (uint64_t)((((uint32_t)(k >> 2) * kBoxBytes) + 32u * (uint32_t)(k & 3)) >> 4); };
    auto issue_s = [&](int g, int st) {    // S_g = Q_g K^T
      const uint64_t a = g ? dQ[1] : dQ[0], b = st ? dK[1] : dK[0];
      if (issuer) {
#pragma unroll
        for (int k = 0; k < kDH / 16; ++k) umma(tmem + kColS + 128u * g, a + koff(k), b + koff(k), id_s, k > 0);
        umma_commit(s_full(g));
      }
      __syncwarp();
    };
    // O_g (+)= P_g V: reduction over keys; P from TMEM (8 columns = 16 keys per step), V rows 16 per step.  The output
    // tile accumulates in TMEM over all kv tiles of the item (`first` starts it): with lazy rescaling the softmax
    // warps have to touch it only when a row's reference max moves.
    auto issue_pv = [&](int g, int st, bool first) {
      if (issuer) {
#ifndef B200FT_EXP_NO_PV   // bottleneck experiment (tools/attn_exp.py): skip the P V UMMAs, wrong numbers
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k)
          umma_ts(tmem + kColPV + (uint32_t)kDH * g,
                  tmem + kColP + C::kStrideP * g + C::kHalfP * (uint32_t)(k >> 2) + 8u * (uint32_t)(k & 3),
                  (st ? dV[1] : dV[0]) + (uint64_t)(128 * k), id_pv, !(first && k == 0));
#endif
        umma_commit(pv_full(g));
      }
      __syncwarp();
    };
    uint32_t kc = 0, cA = 0, cB = 0, n = 0;
    for (int w = item_at(0); w < num_items; w = item_at(++n)) {
    const FwdItem it = decode_item(p, w);
    const int nA = it.nA, nB = it.nB, n_kv = it.n_kv;
    const bool hasB = it.hasB;
    mbar_wait(q_full, n & 1u);
    mbar_wait(k_full((int)(kc & 1u)), (kc >> 1) & 1u);
    // the previous item's last score tiles have been drained by the softmax warps
    if (cA > 0) mbar_wait(s_free(0), (cA - 1) & 1u);
    if (hasB && cB > 0) mbar_wait(s_free(1), (cB - 1) & 1u);
    tc_fence_after();
    issue_s(0, (int)(kc & 1u));
    if (hasB) issue_s(1, (int)(kc & 1u));
    if (issuer) umma_commit(k_empty((int)(kc & 1u)));
    __syncwarp();
    for (int j = 0; j < n_kv; ++j) {
      const uint32_t kj = kc + (uint32_t)j;
      const int st = (int)(kj & 1u), nst = st ^ 1;
      const uint32_t ph = (kj >> 1) & 1u, nph = ((kj + 1) >> 1) & 1u;
      const uint32_t parA = (cA + (uint32_t)j) & 1u, parB = (cB + (uint32_t)j) & 1u;
      bool k_next_ready = false;
      mbar_wait(v_full(st), ph);
      if (!C::kAliasP) {
        // per tile g: the next scores S_g(j+1) first (the softmax is waiting for them), then P_g V(j)
        if (j + 1 < nA) {
          mbar_wait(k_full(nst), nph); k_next_ready = true;
          mbar_wait(s_free(0), parA);      // softmax A has drained S_A(j)
          tc_fence_after();
          issue_s(0, nst);
        }
        if (j < nA) {
          mbar_wait(p_full(0), parA);      // P_A(j) is in TMEM (and O_A rescaled if its reference moved)
          tc_fence_after();
          issue_pv(0, st, j == 0);
        }
        if (j + 1 < nB) {
          if (!k_next_ready) mbar_wait(k_full(nst), nph);
          mbar_wait(s_free(1), parB);
          tc_fence_after();
          issue_s(1, nst);
        }
        if (j < nB) {
          mbar_wait(p_full(1), parB);
          tc_fence_after();
          issue_pv(1, st, j == 0);
        }
      } else {
        // P_g lives on top of S_g: P_g V(j) first, then S_g(j+1) into the same columns (UMMAs execute in issue
        // order, so the new scores land after P has been read); while tile A's softmax works on S_A(j+1) the
        // tensor core runs P_B V(j) and S_B(j+1)
        if (j < nA) {
          if (issuer) TLOG(1, 10);
          mbar_wait(p_full(0), parA);
          tc_fence_after();
          if (issuer) TLOG(1, 11);
          issue_pv(0, st, j == 0);
        }
        if (j + 1 < nA) {
          mbar_wait(k_full(nst), nph); k_next_ready = true;
          mbar_wait(s_free(0), parA);
          tc_fence_after();
          if (issuer) TLOG(1, 12);
          issue_s(0, nst);
        }
        if (j < nB) {
          if (issuer) TLOG(1, 13);
          mbar_wait(p_full(1), parB);
          tc_fence_after();
          if (issuer) TLOG(1, 14);
          issue_pv(1, st, j == 0);
        }
        if (j + 1 < nB) {
          if (!k_next_ready) mbar_wait(k_full(nst), nph);
          mbar_wait(s_free(1), parB);
          tc_fence_after();
          if (issuer) TLOG(1, 15);
          issue_s(1, nst);
          if (issuer) TLOG(1, 16);
        }
      }
      if (issuer) umma_commit(v_empty(st));
      __syncwarp();
      if (j + 1 < n_kv) {
        if (issuer) umma_commit(k_empty(nst));
        __syncwarp();
      }
    }
    if (issuer) umma_commit(q_empty);   // every UMMA of this item has retired: the query tiles may be replaced
    __syncwarp();
    kc += (uint32_t)n_kv; cA += (uint32_t)nA; cB += (uint32_t)nB;
    }  // items
  } else if (warp >= 4) {
    // ================================ softmax, TWO threads per query row ================================
    // 16 warps: group g (query tile A / B) x half (score columns 64*half .. +63 of the row) x TMEM lane quarter.  Four
    // softmax warps per scheduler hide each other's tcgen05.ld / MUFU / tcgen05.st latencies (with one thread per
    // row — two warps per scheduler — a kv tile took ~2500 cycles against 1024 of MUFU work).  The two threads of a row
    // keep separate partial row sums (added once per item) and exchange their half-row maxima through shared memory,
    // double-buffered, with one 64-thread named barrier per exchange, so both work against the same reference.
    const int idx = warp - 4;
    const int g = (idx >> 2) & 1;                 // 0 = tile A, 1 = tile B
    const int half = idx >> 3;                    // which 64 score columns of the row
    const int quarter = warp & 3;                 // TMEM lanes this warp may touch
    const int row = quarter * 32 + lane;          // row within the query tile
    const uint32_t t_lane = tmem + ((uint32_t)(quarter * 32) << 16);
    const uint32_t t_s = t_lane + kColS + 128u * g + 64u * half;                        // my 64 score columns
    const uint32_t t_p = t_lane + kColP + C::kStrideP * g + C::kHalfP * half;           // my 32 packed P columns
    const uint32_t t_pv = t_lane + kColPV + (uint32_t)kDH * g + (uint32_t)(kDH / 2) * half;   // my half of the O columns
    float* xch = reinterpret_cast<float*>(smem_raw + C::kOffXch);
    const int bar_id = 1 + g * 4 + quarter;       // the two warps that share these 32 rows
    uint32_t xn = 0;                              // exchanges done (slot = xn & 1)
    // value of the other half of my row for v (max or sum): write mine, meet the partner warp, read theirs.  Slot
    // xn & 1 is rewritten two exchanges later, after a barrier the partner reaches only past its read of this one.
    auto exchange = [&](float v) -> float {
      float* slot = xch + (xn & 1u) * 512 + g * 256;
      slot[half * 128 + row] = v;
      asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
      const float o = slot[(half ^ 1) * 128 + row];
      ++xn;
      return o;
    };
    uint32_t cg = 0;                              // tiles this group has processed so far
    uint32_t n = 0;
    for (int w = item_at(0); w < num_items; w = item_at(++n)) {
    const FwdItem it = decode_item(p, w);
    const int b = it.b, h = it.h;
    const int bh = b * p.n_head + h;
    const int qt = g ? it.qtB : it.qtA;
    const int n_mine = g ? it.nB : it.nA;
    const int qrow = qt * kBQ + row;              // position in the sequence
    float m_ref = -INFINITY, l = 0.f;             // l: partial row sum over my columns
    for (int j = 0; j < n_mine; ++j) {
      const uint32_t par = (cg + (uint32_t)j) & 1u;
      const bool tl = (warp == 4 && lane == 0);
      if (tl) TLOG(0, 0);
      mbar_wait(s_full(g), par);
      tc_fence_after();
      if (tl) TLOG(0, 1);
      const bool diag = (j == qt);
      const int kbase = j * kBK + 64 * half;      // first key of my columns
      // rescale my columns of this warp's rows of O in TMEM by alpha (rows whose reference stays pass alpha = 1).
      // Only legal while no P V UMMA of this query tile is in flight: those of tile j-1 have retired, those of tile
      // j are not issued before p_full.
      auto rescale_o = [&](float alpha) {
#pragma unroll
        for (int c = 0; c < kDH / 32; ++c) {     // 16 columns per round trip: the 64 scores stay in registers meanwhile
          uint32_t o16[16];
          tmem_ld16(t_pv + 16 * c, o16);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 16; ++i) o16[i] = __float_as_uint(__uint_as_float(o16[i]) * alpha);
          tmem_st16(t_pv + 16 * c, o16);
        }
        tmem_wait_st();
      };
      // ---- my 64 scores -> registers, ONCE.  tcgen05.ld moves 64 bytes per clock per SM: reading a 128 x 128 fp32
      // score tile costs 1024 cycles, as much as its exponentials, so a second read (a separate row-maximum pass, or a
      // repeated sweep) doubles the kernel's longest resource ----
      uint32_t r[64];
      tmem_ld32(t_s, r);
      tmem_ld32(t_s + 32, r + 32);
      tmem_wait_ld();
      if (tl) TLOG(0, 2);
      if (!C::kAliasP) {
        // the scores are in registers: S_g(j+1) may be issued while this tile is exponentiated
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free(g));
      }
      float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // independent chains
      if (diag) {            // warp-uniform: only the diagonal tile carries masking code
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          if (kbase + i > qrow) r[i] = 0xff800000u;   // -inf: exp2 gives 0
          m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(r[i]));
        }
      } else {
#pragma unroll
        for (int i = 0; i < 64; ++i) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(r[i]));
      }
      float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      if (tl) TLOG(0, 3);
      mx = fmaxf(mx, exchange(mx));                // whole-row maximum
      if (tl) TLOG(0, 4);
      if (!C::kAliasP && j > 0) {
        // P has its own columns here: the P V UMMAs of tile j-1 must have read them (and left O quiescent) before P
        // is rewritten or O rescaled.  (Written over S, both follow from s_full(j): it was committed after P V (j-1).)
        mbar_wait(pv_full(g), par ^ 1u);
        tc_fence_after();
      }
      if (j == 0) {
        m_ref = mx * p.scale_log2;
      } else {
        // lazy rescaling: keep the old reference unless this row's max outgrew it by 2^8 (p stays below 2^8)
        const bool grow = mx * p.scale_log2 > m_ref + kRescaleThreshold;
        if (__any_sync(0xffffffffu, grow)) {       // same rows, same maxima: the partner warp decides alike
          float alpha = 1.f;
          if (grow) {
            const float m_new = mx * p.scale_log2;
            alpha = ex2(m_ref - m_new);
            l *= alpha;
            m_ref = m_new;
          }
          rescale_o(alpha);
        }
      }
      // ---- p = exp2(s*scale - m_ref): bf16 P into TMEM (A operand of the P V UMMA), partial row sum ----
      {
        const f32x2_t sc2 = pk2(p.scale_log2, p.scale_log2);
        const f32x2_t nm2 = pk2(-m_ref, -m_ref);
        f32x2_t sa[2] = {pk2(0.f, 0.f), pk2(0.f, 0.f)};
#pragma unroll
        for (int c = 0; c < 4; ++c) {              // 16 keys = 8 packed columns per store
          uint32_t w[8];
#pragma unroll
          for (int i = 0; i < 16; i += 2) {
            const f32x2_t e2 = fma2(pk2(__uint_as_float(r[16 * c + i]), __uint_as_float(r[16 * c + i + 1])), sc2, nm2);
            float p0, p1;
            if (((i >> 1) & C::kPolyMask) == C::kPolyMask) {   // compile-time: one pair in four on the FMA pipe (d_head 128)
              upk2(exp2_fma2(e2), p0, p1);
            } else {
              float e0, e1;
              upk2(e2, e0, e1);
              p0 = ex2(e0); p1 = ex2(e1);               // exp2(-inf) = 0 for masked keys
            }
            sa[(i >> 1) & 1] = add2(sa[(i >> 1) & 1], pk2(p0, p1));
            w[i >> 1] = pack_bf16x2(p0, p1);
          }
          tmem_st8(t_p + 8 * c, w);
        }
        float s_lo, s_hi;
        upk2(add2(sa[0], sa[1]), s_lo, s_hi);
        l += s_lo + s_hi;
      }
      if (tl) TLOG(0, 5);
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (C::kAliasP) mbar_arrive(s_free(g));
        mbar_arrive(p_full(g));
      }
      if (tl) TLOG(0, 6);
    }
    if (n_mine > 0) {
      // ---- all P V UMMAs of the item have retired: normalise my half of O, write it and the log-sum-exp ----
      mbar_wait(pv_full(g), (cg + (uint32_t)n_mine - 1u) & 1u);
      tc_fence_after();
      l += exchange(l);                            // whole-row sum
      const float inv_l = 1.f / l;
      __nv_bfloat16* dst = p.out + ((int64_t)b * p.seq + qrow) * p.ld_out + h * kDH + (kDH / 2) * half;
#pragma unroll
      for (int c = 0; c < kDH / 64; ++c) {   // 32 columns (64 bytes of output) per TMEM round trip
        uint32_t r[32];
        float o[32];
        tmem_ld32(t_pv + 32 * c, r);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = __uint_as_float(r[i]) * inv_l;
        if (qrow < p.seq) {
#pragma unroll
          for (int v = 0; v < 4; ++v) *reinterpret_cast<uint4*>(dst + 32 * c + 8 * v) = pack8(o + 8 * v);
        }
      }
      if (half == 0 && qrow < p.seq) p.lse2[(int64_t)bh * p.seq + qrow] = m_ref + log2f(l);
    }
    cg += (uint32_t)n_mine;
    }  // items
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
  }
}

}  // namespace

extern "C" int b200ft_attn_fwd(const void* qkv, int64_t ld, int q_off, int k_off, int v_off, void* out,
                               int64_t ld_out, float* lse2, int batch, int seq, int n_head, int n_kv_head,
                               int d_head, float scale, void* stream) {
  B200FT_CHECK_ARG(qkv && out && lse2, "attn_fwd: null buffer");
  if (d_head != 64 && d_head != 128) {
    b200ft_set_error("attn_fwd: d_head=%d is not built (64 and 128 are)", d_head);
    return B200FT_ERR_UNSUPPORTED;
  }
  const int kDH = d_head;
  B200FT_CHECK_ARG(batch > 0 && seq > 0 && n_head > 0 && n_kv_head > 0 && n_head % n_kv_head == 0,
                   "attn_fwd: batch=%d seq=%d heads=%d kv_heads=%d", batch, seq, n_head, n_kv_head);
  B200FT_CHECK_ARG(ld % 8 == 0 && ld_out % 8 == 0 && q_off % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0 &&
                       ((uintptr_t)qkv | (uintptr_t)out) % 16 == 0 && ld_out >= (int64_t)n_head * kDH,
                   "attn_fwd: alignment (ld=%lld ld_out=%lld offsets %d/%d/%d)", (long long)ld, (long long)ld_out,
                   q_off, k_off, v_off);
  EncodeTiledFn enc = encode_fn();
  if (!enc) { b200ft_set_error("attn_fwd: cuTensorMapEncodeTiled unavailable"); return B200FT_ERR_UNSUPPORTED; }
  // the fused buffer as [batch][seq][columns]: rows past the end of a sequence are zero-filled by TMA
  CUtensorMap tm;
  if (int rc = make_map(enc, &tm, qkv, ld, seq, batch)) return rc;
  AttnParams p;
  p.batch = batch; p.seq = seq; p.n_head = n_head; p.n_kv_head = n_kv_head;
  p.q_off = q_off; p.k_off = k_off; p.v_off = v_off;
  p.nq_tiles = (seq + kBQ - 1) / kBQ;
  p.nq_pairs = (p.nq_tiles + 1) / 2;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.out = (__nv_bfloat16*)out; p.ld_out = ld_out; p.lse2 = lse2;
  static thread_local int attr_dev = -1;
  int dev = -1;
  B200FT_CUDA(cudaGetDevice(&dev));
  if (dev != attr_dev) {
    B200FT_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<64>::kSmemBytes));
    B200FT_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<128>::kSmemBytes));
    attr_dev = dev;
  }
  const int items = p.nq_pairs * batch * n_head;
  // Short sequences (few kv tiles per item): persistent CTAs, so that an item's prologue hides behind its
  // predecessor (T = 256 .. 1024: 6-15 % faster).  Long ones: one CTA per item, where the hardware's dynamic CTA
  // placement balances the very unequal items better than the static walk (T >= 2048: 2-3 % faster).
  const bool persistent = p.nq_tiles <= 8 && items > B200FT_NUM_SMS;
  const unsigned grid = (unsigned)(persistent ? B200FT_NUM_SMS : items);
  if (d_head == 64) B200FT_CUDA(launch_k(attn_fwd_kernel<64>, grid, kThreads, Cfg<64>::kSmemBytes, (cudaStream_t)stream, tm, p));
  else              B200FT_CUDA(launch_k(attn_fwd_kernel<128>, grid, kThreads, Cfg<128>::kSmemBytes, (cudaStream_t)stream, tm, p));
  return 0;
}

#ifdef B200FT_ATTN_TIMING
extern "C" int b200ft_debug_attn_timing(long long* log_out, int* counts_out, int reset) {
  if (reset) {
    int z[2] = {0, 0};
    B200FT_CUDA(cudaMemcpyToSymbol(g_tcnt, z, sizeof(z)));
    return 0;
  }
  B200FT_CUDA(cudaDeviceSynchronize());
  B200FT_CUDA(cudaMemcpyFromSymbol(log_out, g_tlog, sizeof(long long) * 2 * 4096 * 2));
  B200FT_CUDA(cudaMemcpyFromSymbol(counts_out, g_tcnt, sizeof(int) * 2));
  return 0;
}
#endif
