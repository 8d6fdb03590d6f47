// bf16 GEMM on Blackwell 5th-gen tensor cores: TMA -> shared memory (128B swizzle) ->
// tcgen05.mma (fp32 accumulators in TMEM) -> tcgen05.ld epilogue.
//
// Two instantiations of one kernel template:
//   CG = 1  one CTA per 128 x BN tile (cta_group::1), 4-stage ring.  Handles every layout
//           (K-major / MN-major operands), split-K, tiny N.
//   CG = 2  a cluster of two CTAs (one TPC) per 256 x BN tile (cta_group::2): each CTA stages
//           its own 128 rows of A and HALF of the B tile, the leader CTA issues 256 x BN x 16
//           UMMAs that read both CTAs' shared memory and write both CTAs' TMEM.  Per FLOP this
//           moves 2/3 of the operand bytes of CG = 1 from L2 into the SMs — and the L2->SM
//           fabric (~42 B/clk/SM), not the tensor pipe, is what bounds CG = 1 (measured:
//           profiles/README.md) — and the halved B stage allows a 6-stage ring.
//
// Persistent, warp-specialised, one CTA per SM (256 threads):
//   warp 0 : TMA producer (one lane)          smem ring of kStages {A 128x64, B BNx64} slabs
//   warp 1 : MMA issuer (one lane)            128 x BN x 16 UMMAs, 4 per K slab
//   warp 2 : TMEM allocator (512 columns = two 128 x 256 fp32 accumulators)
//   warps 4-11 : epilogue, two warps per TMEM lane quarter (32 rows), taking alternate
//               32-column chunks; runs on accumulator `a` while the MMA warp fills accumulator
//               `a^1` for the next tile.  Per chunk: tcgen05.ld -> registers -> (alpha, bias from
//               a per-tile smem copy, residual / gelu / dgelu) -> bf16 into a 64B-swizzled 32x32
//               smem tile -> one TMA store (coalesced, asynchronous, clipped at the matrix edge).
//               The residual / dgelu operand arrives by TMA load into the same tile one chunk
//               ahead and is updated in place.  (Two warps per scheduler, a 32-column body: the
//               first version — one warp per quarter, 64-column chunks, bias from global — was
//               bound by instruction issue, i-cache misses and LDG latency; profiles/README.md.)
// Three mbarrier pipelines: smem full/empty (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue).
//
// The K loop runs over up to two operand segments.  Segment 2 is how a LoRA branch is
// fused: A2 = s*(X*A_lora) [M, r_pad], B2 = B_lora^T [N, r_pad] accumulate into the same
// TMEM tile as the frozen projection, and only ceil(r_pad/16) UMMAs are issued for it.
//
// Operand layouts.  K-major (reduction index contiguous): one TMA box of {64 k, rows} per slab,
// UMMA descriptor SW128 / SBO = 1024 B, K advance = 32 B inside the swizzle row.  MN-major
// (M or N index contiguous): TMA boxes of {64 mn, 64 k} (8 KB each), UMMA descriptor SW128 /
// LBO = 8192 B between 64-wide MN chunks / SBO = 1024 B between 8-row K groups, K advance =
// 2048 B.  Descriptor bit layouts follow the PTX ISA "tcgen05 matrix / instruction descriptor"
// tables.
#include <cuda.h>
#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // 64 bf16 = 128 B = one swizzle row
constexpr int kMaxBN = 256;
constexpr int kAStageBytes = kBlockM * kBlockK * 2;  // 16 KB
constexpr int kTileBytes = 160 * 1024;  // operand ring
// Ring depth and the B slab's size are per launch (GemmParams::n_stages, b_stage_bytes): a slab of B is BN / CG rows of
// 128 bytes (K-major) or 8 KB per 64 columns (MN-major), so the 160 KB hold 3 stages at BN = 256 on one CTA, 5 on a
// pair, and 6-8 for the N = 16 / 32 adapter GEMMs and gradient contractions — which stream [M x K] once and are bound
// by the bytes they keep in flight (3 stages = 48 KB of A per SM measured 2.5-3.9 TB/s).
constexpr int kMaxStages = 8;
constexpr int kMaxSeg = 4;
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // clears the CTA-rank bit of a shared::cluster address -> CTA 0
#ifndef B200FT_GEMM_EPI_WPQ
#define B200FT_GEMM_EPI_WPQ 2
#endif
constexpr int kWPQ = B200FT_GEMM_EPI_WPQ;              // epilogue warps per TMEM lane quarter (= per scheduler)
constexpr int kEpiWarps = 4 * kWPQ;
constexpr int kEpiBufBytes = 32 * 32 * 2;              // one 32-row x 32-col bf16 staging tile (2 KB)
constexpr int kEpiBufs = 8 / kWPQ;                     // staging ring per epilogue warp (64 KB in total either way)
constexpr int kEpiBufShift = kEpiBufs == 4 ? 2 : 1;
static_assert(kWPQ == 2 || kWPQ == 4, "epilogue warps per quarter");
constexpr int kEpiBytes = kEpiWarps * kEpiBufs * kEpiBufBytes;  // 64 KB
constexpr int kBiasBytes = kMaxBN * 4;                 // the tile's bias slice, fp32
constexpr int kBarrierBytes = 512;
constexpr int kSmemBytes = kTileBytes + kEpiBytes + kBiasBytes + kBarrierBytes + 1024;  // + slack to align to 1024
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
constexpr int kThreads = 128 + 32 * kEpiWarps;         // 4 control warps + 8 epilogue warps
constexpr int kTmemCols = 512;
constexpr int kAccStride = 256;  // TMEM columns per accumulator stage
constexpr long long kWaitLimitCycles = 4000000000ll;  // ~2 s: deadlock guard, trap instead of hanging the GPU

struct TmapPack {  // one __grid_constant__ block: TMA reads descriptors straight from param space
  CUtensorMap a[kMaxSeg], b[kMaxSeg];
  CUtensorMap c, c2, aux;
};

#ifdef B200FT_ATTN_TIMING   // per-phase clock readings of CTA 0 (tools/gemm_timing.py); never defined in the shipped build
__device__ long long g_glog[2][8192][2];
__device__ int g_gcnt[2];
#define GLOG(role, tag) do { if (blockIdx.x == 0) { int i__ = g_gcnt[role]++; if (i__ < 8192) { g_glog[role][i__][0] = (tag); g_glog[role][i__][1] = clock64(); } } } while (0)
#ifdef B200FT_GEMM_TIMING_FINE   // also per k-block and per epilogue chunk (each reading costs ~250 cycles itself)
#define GLOG2(role, tag) GLOG(role, tag)
#else
#define GLOG2(role, tag) do { } while (0)
#endif
#else
#define GLOG(role, tag) do { } while (0)
#define GLOG2(role, tag) do { } while (0)
#endif

struct GemmParams {
  int M, N, BN;
  int nseg;                    // K segments: acc = sum_s A_s . B_s^T
  int k[kMaxSeg];              // reduction length of each
  int kb_prefix[kMaxSeg + 1];  // first 64-wide K slab of each (prefix sums)
  int a_mn, b_mn;
  int tiles_m, tiles_n;
  int raster_m;  // 1: consecutive work items walk the M tiles of one N tile (B streamed once, A resident in L2)
  int splits, kb_per_split;  // split-K: work item w -> (tile = w / splits, K slabs of split w % splits)
  int epilogue;
  int n_stages, b_stage_bytes;   // operand ring: depth and bytes per B slab (A slabs are 16 KB)
  int cw;  // epilogue chunk width in columns: 64 (swizzled staging tile), 32 or 16 (dense)
  float alpha;
  const float* bias;
  const __nv_bfloat16* aux; int64_t ld_aux;
  __nv_bfloat16* c; int64_t ldc;
  __nv_bfloat16* c2; int64_t ldc2;
  float* cf;  // ATOMIC_F32_T: fp32 [N, ldc], cf[n, m] += alpha * acc[m, n]
};

__device__ __forceinline__ int tile_m_index(const GemmParams& p, int tile) {
  return p.raster_m ? tile % p.tiles_m : tile / p.tiles_n;
}
__device__ __forceinline__ int tile_n_index(const GemmParams& p, int tile) {
  return p.raster_m ? tile / p.tiles_m : tile % p.tiles_n;
}

// ---------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
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
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > kWaitLimitCycles) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
// cta_group::2 forms -------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 remote;\n\t"
      "mapa.shared::cluster.u32 remote, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [remote];\n\t}"
      ::"r"(bar), "r"(cta)
      : "memory");
}
// TMA load issued by either CTA of a pair; completion bytes are credited to CTA 0's barrier
__device__ __forceinline__ void tma_load_2d_cg2(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar & kPeerBitMask)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_cg2(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// commit: arrive on the barrier at this offset in BOTH CTAs of the pair when the MMAs retire
__device__ __forceinline__ void umma_commit_cg2(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(bar), "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_cg2(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_cg2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(c0), "r"(c1)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ uint4 lds16(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts16(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
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
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// The descriptor's upper word is constant per layout; only the start-address field of the lower
// word moves (per stage, per K step), so the issue loop adds small integers to `lo`.
__device__ __forceinline__ uint32_t smem_desc_hi(uint32_t sbo_bytes) {
  return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
}
__device__ __forceinline__ uint32_t smem_desc_lo(uint32_t addr, uint32_t lbo_bytes) {
  return ((addr & 0x3FFFFu) >> 4) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
__device__ __forceinline__ uint64_t desc64(uint32_t lo, uint32_t hi) { return ((uint64_t)hi << 32) | lo; }
// one lane of a converged warp (the same lane every time for a full mask)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}

// instruction descriptor (kind::f16): D fp32 [4,6)=1, A bf16 [7,10)=1, B bf16 [10,13)=1,
// A major [15], B major [16] (1 = MN-major), N>>3 [17,23), M>>4 [24,29)
__device__ __forceinline__ uint32_t make_idesc(int m, int n, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}


// ---------------------------------------------------------------------------------------
// epilogue of one warp: TMEM lane quarter q (rows m0 + 32q .. +31) of every tile this CTA owns
// ---------------------------------------------------------------------------------------
// address of the 16-byte vector `v` of row `lane` inside a staging tile of CW columns
template <int CW>
__device__ __forceinline__ uint32_t epi_vec_addr(uint32_t buf, int lane, int v) {
  if constexpr (CW == 32) return buf + lane * 64 + ((v ^ ((lane >> 1) & 3)) << 4);  // TMA SWIZZLE_64B pattern
  else return buf + lane * (CW * 2) + (v << 4);                                     // dense rows (SWIZZLE_NONE)
}
__device__ __forceinline__ void epi_bar_sync() {  // the 8 epilogue warps only
  asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

// q = TMEM lane quarter, h = which of the quarter's two warps (takes chunks h, h+2, ...),
// et = index among the 256 epilogue threads.
//
// Staging ring: kEpiBufs 2 KB tiles per warp, used in sequence (n & 3).  For the epilogues with a
// second operand (residual / gelu') chunk n's operand is TMA-loaded into buffer n & 3 THREE chunks
// ahead of its use — across tile boundaries — so that every warp keeps 6 KB (48 KB per SM) of
// that stream in flight; the result is written over it in place and leaves by TMA store.
template <int CW, int CG, int EPI>
__device__ __forceinline__ void epilogue_loop(const GemmParams& p, const CUtensorMap* tmC, const CUtensorMap* tmC2,
                                              const CUtensorMap* tmAux, int q, int h, int lane, int et,
                                              uint32_t tmem_base, uint32_t ebuf, uint32_t sbias, uint32_t abar,
                                              uint32_t tfull0, uint32_t tempty0, int num_work, int w_begin,
                                              int w_stride, int cta_rank) {
  const int BN = p.BN;
  const int nchunks = BN / CW;
  const int cpw = nchunks > h ? (nchunks - h + kWPQ - 1) / kWPQ : 0;  // chunks of a tile this warp owns
  // The epilogue kind is a template parameter and alpha / bias are applied unconditionally (one FFMA2 per pair, the
  // bias slice being zeros when there is none): run-time branches inside the per-vector arithmetic cut it into basic
  // blocks the compiler cannot interleave, and with two epilogue warps per scheduler that dependent chain was the
  // epilogue's time (ISETP / BRA / SEL were 17 % of the x GELU' kernel's instructions).
  constexpr bool has_aux = (EPI == B200FT_EPI_RESIDUAL) || (EPI == B200FT_EPI_DGELU);
  constexpr bool atomic_t = (EPI == B200FT_EPI_ATOMIC_F32_T);
  const bool has_bias = p.bias != nullptr;
  if (!has_bias && !atomic_t) {   // zeros once: the arithmetic below always adds the slice
    if (et < kMaxBN) asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias + 4u * et), "f"(0.f) : "memory");
    epi_bar_sync();
  }
  const uint32_t aux_bytes = 32u * CW * 2u;
  constexpr uint32_t kDepth = kEpiBufs - 1;  // operand prefetch distance, in chunks

  // operand load of this warp's chunk number n (counted across all its tiles); lane 0 only.  The chunks are requested
  // strictly in order, so a cursor (work item, chunk within the tile, the tile's row / column origin) replaces the
  // divisions that turn n into coordinates: done per chunk they cost the issuing lane ~800 cycles, a quarter of the
  // chunk's time, and made the epilogues with a second operand slower than the K = 768 main loop they hide behind.
  int pf_w = w_begin, pf_ci = 0, pf_r0 = 0, pf_c0 = 0;
  bool pf_valid = false;
  auto pf_tile = [&]() {
    pf_valid = pf_w < num_work;
    if (pf_valid) {
      const int tile2 = pf_w / p.splits;
      pf_r0 = tile_m_index(p, tile2) * (kBlockM * CG) + cta_rank * kBlockM + q * 32;
      pf_c0 = tile_n_index(p, tile2) * BN + h * CW;
    }
  };
  auto issue_aux = [&](uint32_t n) {
    if (!pf_valid) return;
    const uint32_t b = n & (kEpiBufs - 1);
    mbar_expect_tx(abar + 8u * b, aux_bytes);
    tma_load_2d(ebuf + b * kEpiBufBytes, tmAux, pf_c0 + kWPQ * CW * pf_ci, pf_r0, abar + 8u * b);
    if (++pf_ci == cpw) { pf_ci = 0; pf_w += w_stride; pf_tile(); }
  };
  if (has_aux && cpw > 0 && lane == 0) pf_tile();
  if (has_aux && cpw > 0 && lane == 0) {
    for (uint32_t n = 0; n < kDepth; ++n) issue_aux(n);
  }

  uint32_t seq = 0;  // staged tiles so far: buffer = seq & 3 (== chunk number when has_aux)
  int acc = 0;
  uint32_t acc_phase = 0;
  for (int w = w_begin; w < num_work; w += w_stride) {
    const int tile = w / p.splits;
    const int m0 = tile_m_index(p, tile) * (kBlockM * CG) + cta_rank * kBlockM;
    const int n0 = tile_n_index(p, tile) * BN;
    const int row0 = m0 + q * 32;
    if (has_bias) {  // this tile's bias slice -> smem, once, while the MMAs of the tile still run
      epi_bar_sync();  // everyone is done with the previous tile's slice
      if (et < BN) {
        const int col = n0 + et;
        const float bv = col < p.N ? __ldg(p.bias + col) : 0.f;
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias + 4u * et), "f"(bv) : "memory");
      }
      epi_bar_sync();
    }
    if (et == 0) GLOG(0, 30);
    mbar_wait(tfull0 + 8u * acc, acc_phase);
    tc_fence_after();
    if (et == 0) GLOG(0, 31);
    const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * kAccStride);
    // One chunk: its accumulator columns are already in registers (r).
    auto process_chunk = [&](const int c, uint32_t (&r)[CW]) {
      const uint32_t b = seq & (kEpiBufs - 1);
      const uint32_t buf = ebuf + b * kEpiBufBytes;
      const int colbase = n0 + c * CW;
      if (atomic_t) {
        // split-K partial: fp32 reduction into the transposed output (lanes = consecutive m,
        // so every red.global.add of a warp hits one 128-byte line)
        const int mrow = row0 + lane;
        if (mrow < p.M) {
#pragma unroll
          for (int j = 0; j < CW; ++j) {
            if (colbase + j < p.N)
              atomicAdd(p.cf + (int64_t)(colbase + j) * p.ldc + mrow, __uint_as_float(r[j]) * p.alpha);
          }
        }
        return;
      }
      if (et == 0) GLOG2(0, 33);
      if (has_aux) {
        mbar_wait(abar + 8u * b, (seq >> kEpiBufShift) & 1u);  // operand of this chunk has landed
        if (et == 0) GLOG2(0, 34);
      } else {
        if (lane == 0) tma_store_wait_read<kEpiBufs - 1>();  // the store 4 tiles back has drained this buffer
        __syncwarp();
      }
      if (et == 0) GLOG2(0, 35);
      // the arithmetic below runs on packed fp32 pairs (FFMA2/FMUL2/FADD2): half the issue slots
      const f32x2_t alpha2 = pk2(p.alpha, p.alpha);
#pragma unroll
      for (int v = 0; v < CW / 8; ++v) {
        f32x2_t f[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
          f[i] = pk2(__uint_as_float(r[v * 8 + 2 * i]), __uint_as_float(r[v * 8 + 2 * i + 1]));
        {
          const float4 b0 = lds_f4(sbias + 4u * (c * CW + v * 8));
          const float4 b1 = lds_f4(sbias + 4u * (c * CW + v * 8 + 4));
          f[0] = fma2(f[0], alpha2, pk2(b0.x, b0.y)); f[1] = fma2(f[1], alpha2, pk2(b0.z, b0.w));
          f[2] = fma2(f[2], alpha2, pk2(b1.x, b1.y)); f[3] = fma2(f[3], alpha2, pk2(b1.z, b1.w));
        }
        const uint32_t addr = epi_vec_addr<CW>(buf, lane, v);
        if (has_aux) {
          const uint4 aw = lds16(addr);
          const uint32_t a4[4] = {aw.x, aw.y, aw.z, aw.w};
          if constexpr (EPI == B200FT_EPI_RESIDUAL) {
#pragma unroll
            for (int i = 0; i < 4; ++i) f[i] = add2(f[i], bf16x2_to_f32x2(a4[i]));
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) f[i] = mul2(f[i], gelu_new_grad2(bf16x2_to_f32x2(a4[i])));
          }
        }
        uint4 packed;
        packed.x = f32x2_to_bf16x2(f[0]); packed.y = f32x2_to_bf16x2(f[1]);
        packed.z = f32x2_to_bf16x2(f[2]); packed.w = f32x2_to_bf16x2(f[3]);
        sts16(addr, packed);
        if constexpr (EPI == B200FT_EPI_GELU) {  // keep the bf16-rounded pre-activation for pass 2
          r[v * 8 + 0] = packed.x; r[v * 8 + 1] = packed.y; r[v * 8 + 2] = packed.z; r[v * 8 + 3] = packed.w;
        }
      }
      if (et == 0) GLOG2(0, 36);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) tma_store_2d(tmC, buf, colbase, row0);
      if (et == 0) GLOG2(0, 37);
      ++seq;
      if constexpr (EPI == B200FT_EPI_GELU) {
        // second output: gelu of the bf16-ROUNDED pre-activation, so that backward (which only
        // has the stored bf16 value) differentiates exactly the function applied here
        const uint32_t buf2 = ebuf + (seq & (kEpiBufs - 1)) * kEpiBufBytes;
        if (lane == 0) tma_store_wait_read<kEpiBufs - 1>();
        __syncwarp();
#pragma unroll
        for (int v = 0; v < CW / 8; ++v) {
          uint4 gq;
          gq.x = f32x2_to_bf16x2(gelu_new2(bf16x2_to_f32x2(r[v * 8 + 0])));
          gq.y = f32x2_to_bf16x2(gelu_new2(bf16x2_to_f32x2(r[v * 8 + 1])));
          gq.z = f32x2_to_bf16x2(gelu_new2(bf16x2_to_f32x2(r[v * 8 + 2])));
          gq.w = f32x2_to_bf16x2(gelu_new2(bf16x2_to_f32x2(r[v * 8 + 3])));
          sts16(epi_vec_addr<CW>(buf2, lane, v), gq);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) tma_store_2d(tmC2, buf2, colbase, row0);
        ++seq;
      }
      if (has_aux && lane == 0) {
        // chunk seq-1 was just stored; chunk seq+2 reuses the buffer of chunk seq-2, whose store
        // is older than the one just issued: allow only that newest store to be unread
        tma_store_wait_read<1>();
        if (et == 0) GLOG2(0, 38);
        issue_aux(seq - 1 + kDepth);
        if (et == 0) GLOG2(0, 39);
      }
    };
    // Software pipeline over this warp's chunks (h, h + kWPQ, ...): the tcgen05.ld of the next chunk is issued before
    // the current one is processed, into the other register set (measured: ~1070 cycles per plain chunk of which the
    // load's round trip was a third; the epilogues with arithmetic were slower than the K = 768 main loop they hide
    // behind).  tcgen05.wait::ld waits for every outstanding load, hence the strict alternation.
    auto load_chunk = [&](const int c, uint32_t (&r)[CW]) {
      if constexpr (CW == 32) tmem_ld32(t_row + c * CW, r);
      else tmem_ld16(t_row + c * CW, r);
    };
    uint32_t rA[CW], rB[CW];
    int c = h;
    if (c < nchunks) load_chunk(c, rA);
    while (c < nchunks) {
      tmem_wait_ld();
      const int c1 = c + kWPQ;
      if (c1 < nchunks) load_chunk(c1, rB);
      process_chunk(c, rA);
      if (c1 >= nchunks) break;
      tmem_wait_ld();
      const int c2 = c1 + kWPQ;
      if (c2 < nchunks) load_chunk(c2, rA);
      process_chunk(c1, rB);
      c = c2;
    }
    if (et == 0) GLOG(0, 32);
    tc_fence_before();
    __syncwarp();
    if (lane == 0) {  // accumulator drained: tell the MMA issuer (it lives in CTA 0 of the pair)
      if constexpr (CG == 1) mbar_arrive(tempty0 + 8u * acc);
      else mbar_arrive_cluster(tempty0 + 8u * acc, 0);
    }
    acc ^= 1;
    if (acc == 0) acc_phase ^= 1u;
  }
  if (lane == 0) tma_store_wait_all();  // smem must outlive the bulk stores reading it
  __syncwarp();
}

// ---------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------
template <int CG>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ TmapPack maps, const GemmParams p) {
  const int kStages = p.n_stages;
  const uint32_t kBStageBytes = (uint32_t)p.b_stage_bytes;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;  // identical in both CTAs of a pair (same kernel, same layout)
  const uint32_t sA = base;
  const uint32_t sB = base + kStages * kAStageBytes;
  const uint32_t sEpi = base + kTileBytes;
  const uint32_t sBias = sEpi + kEpiBytes;
  const uint32_t bars = sBias + kBiasBytes;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int a) { return bars + 8u * (2 * kMaxStages + a); };
  auto tempty_bar = [&](int a) { return bars + 8u * (2 * kMaxStages + 2 + a); };
  auto aux_bar = [&](int ew, int b) { return bars + 8u * (2 * kMaxStages + 4 + kEpiBufs * ew + b); };
  const uint32_t tmem_slot = bars + 8u * (2 * kMaxStages + 4 + kEpiBufs * kEpiWarps);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(
      smem_raw + (base - raw) + kTileBytes + kEpiBytes + kBiasBytes + 8 * (2 * kMaxStages + 4 + kEpiBufs * kEpiWarps));
  static_assert(8 * (2 * kMaxStages + 4 + kEpiBufs * kEpiWarps) + 4 <= kBarrierBytes, "barrier area");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int cta_rank = CG == 1 ? 0 : (int)cluster_ctarank();
  const int w_begin = blockIdx.x / CG;   // work items are handed to CTA pairs when CG == 2
  const int w_stride = gridDim.x / CG;

  if (warp == 0 && lane == 0) {
    for (int sgi = 0; sgi < p.nseg; ++sgi) { tma_prefetch_desc(&maps.a[sgi]); tma_prefetch_desc(&maps.b[sgi]); }
    tma_prefetch_desc(&maps.c);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full_bar(s), CG); mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), kEpiWarps * CG); }
    for (int ew = 0; ew < kEpiWarps; ++ew)
      for (int b = 0; b < kEpiBufs; ++b) mbar_init(aux_bar(ew, b), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    if constexpr (CG == 1) tmem_alloc(tmem_slot, kTmemCols);
    else tmem_alloc_cg2(tmem_slot, kTmemCols);
  }
  tc_fence_before();
  if constexpr (CG == 1) __syncthreads();
  else cluster_sync_all();  // the peer's barriers must be initialised before anything remote touches them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  grid_dep_sync();  // everything above ran under the previous kernel's tail; global memory is touched only below

  const int num_work = p.tiles_m * p.tiles_n * p.splits;
  const int kb_total = p.kb_prefix[p.nseg];
  const int BN = p.BN;

  if (warp == 0) {
    // =============================== TMA producer (every CTA) ===============================
    // The whole warp walks the loop (warp-uniform control flow keeps addresses and coordinates in
    // uniform registers); one elected lane arms the barrier and issues the copies.
    const bool issuer = elect_one();
    const uint32_t b_rows = (uint32_t)BN / CG;  // CG == 2: this CTA stages half of the B tile
    // MN-major B is staged in 64-column boxes of 8 KB (CG == 2: BN / 2 is a multiple of 64, host-checked)
    const uint32_t b_bytes = p.b_mn ? (uint32_t)((b_rows + 63) / 64) * 8192u : b_rows * 128u;
    const uint32_t b2_bytes = b_rows * 128u;
    const bool a_mn = p.a_mn, b_mn = p.b_mn;
    int stage = 0;
    uint32_t phase = 0;
    for (int w = w_begin; w < num_work; w += w_stride) {
      const int tile = w / p.splits;
      const int m0 = tile_m_index(p, tile) * (kBlockM * CG) + cta_rank * kBlockM;
      const int n0 = tile_n_index(p, tile) * BN + cta_rank * (int)b_rows;
      const int kb_begin = (w % p.splits) * p.kb_per_split;
      const int kb_end = min(kb_total, kb_begin + p.kb_per_split);
      int seg = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        while (kb >= p.kb_prefix[seg + 1]) ++seg;
        const bool seg2 = seg > 0;  // only segment 0 may be MN-major
        const int kk = (kb - p.kb_prefix[seg]) * kBlockK;
        const CUtensorMap* tma_a = &maps.a[seg];
        const CUtensorMap* tma_b = &maps.b[seg];
        const uint32_t a_dst = sA + stage * kAStageBytes;
        const uint32_t b_dst = sB + stage * kBStageBytes;
        const uint32_t fb = full_bar(stage);
        const uint32_t bytes = (uint32_t)kAStageBytes + (seg2 ? b2_bytes : b_bytes);
        if (issuer) {
          if constexpr (CG == 1) {
            mbar_expect_tx(fb, bytes);
            if (seg2) {
              tma_load_2d(a_dst, tma_a, kk, m0, fb);
              tma_load_2d(b_dst, tma_b, kk, n0, fb);
            } else {
              if (a_mn) {
                tma_load_2d(a_dst, tma_a, m0, kk, fb);
                tma_load_2d(a_dst + 8192u, tma_a, m0 + 64, kk, fb);
              } else {
                tma_load_2d(a_dst, tma_a, kk, m0, fb);
              }
              if (b_mn) {
                for (int j = 0; j * 64 < BN; ++j) tma_load_2d(b_dst + 8192u * j, tma_b, n0 + 64 * j, kk, fb);
              } else {
                tma_load_2d(b_dst, tma_b, kk, n0, fb);
              }
            }
          } else {
            // both CTAs' bytes land on CTA 0's barrier: the leader arms it with the pair's total,
            // the peer adds its (byte-less) arrival
            if (cta_rank == 0) mbar_expect_tx(fb, 2u * bytes);
            else mbar_arrive_cluster(fb, 0);
            if (a_mn && !seg2) {
              tma_load_2d_cg2(a_dst, tma_a, m0, kk, fb);
              tma_load_2d_cg2(a_dst + 8192u, tma_a, m0 + 64, kk, fb);
            } else {
              tma_load_2d_cg2(a_dst, tma_a, kk, m0, fb);
            }
            if (b_mn && !seg2) {
              for (int j = 0; j * 64 < (int)b_rows; ++j) tma_load_2d_cg2(b_dst + 8192u * j, tma_b, n0 + 64 * j, kk, fb);
            } else {
              tma_load_2d_cg2(b_dst, tma_b, kk, n0, fb);
            }
          }
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer (CTA 0 of a pair) ===============================
    // Warp-uniform loop; one elected lane issues every tcgen05.mma / tcgen05.commit (a commit only
    // tracks the MMAs of the thread that executes it).
    if (cta_rank == 0) {
      const bool issuer = elect_one();
      const bool a_mn = p.a_mn, b_mn = p.b_mn;
      const uint32_t idesc1 = make_idesc(kBlockM * CG, BN, a_mn, b_mn);
      const uint32_t idesc2 = make_idesc(kBlockM * CG, BN, 0, 0);
      // descriptor words: K-major SW128 (LBO 16, SBO 1024, +32 B per K step) or MN-major SW128
      // (LBO 8192, SBO 1024, +2048 B per K step)
      const uint32_t hi = smem_desc_hi(1024u);
      const uint32_t a_lo1 = smem_desc_lo(sA, a_mn ? 8192u : 16u), a_step1 = a_mn ? 128u : 2u;
      const uint32_t b_lo1 = smem_desc_lo(sB, b_mn ? 8192u : 16u), b_step1 = b_mn ? 128u : 2u;
      const uint32_t a_lo2 = smem_desc_lo(sA, 16u), b_lo2 = smem_desc_lo(sB, 16u);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int w = w_begin; w < num_work; w += w_stride) {
        if (issuer) GLOG(1, 20);
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        if (issuer) GLOG(1, 21);
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * kAccStride);
        const int kb_begin = (w % p.splits) * p.kb_per_split;
        const int kb_end = min(kb_total, kb_begin + p.kb_per_split);
        int seg = 0;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          if (issuer) GLOG2(1, 22);
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          if (issuer) GLOG2(1, 23);
          while (kb >= p.kb_prefix[seg + 1]) ++seg;
          const bool seg2 = seg > 0;
          const int kleft = p.k[seg] - (kb - p.kb_prefix[seg]) * kBlockK;
          const uint32_t a_lo = (seg2 ? a_lo2 : a_lo1) + (uint32_t)stage * (kAStageBytes >> 4);
          const uint32_t b_lo = (seg2 ? b_lo2 : b_lo1) + (uint32_t)stage * (kBStageBytes >> 4);
          const uint32_t a_step = seg2 ? 2u : a_step1, b_step = seg2 ? 2u : b_step1;
          const uint32_t idesc = seg2 ? idesc2 : idesc1;
          const uint32_t first = kb > kb_begin ? 1u : 0u;
          if (issuer) {
            if (kleft >= kBlockK) {  // the common case: four K steps, no per-step predicate
              if constexpr (CG == 1) {
                umma_bf16(d_tmem, desc64(a_lo, hi), desc64(b_lo, hi), idesc, first);
                umma_bf16(d_tmem, desc64(a_lo + a_step, hi), desc64(b_lo + b_step, hi), idesc, 1u);
                umma_bf16(d_tmem, desc64(a_lo + 2 * a_step, hi), desc64(b_lo + 2 * b_step, hi), idesc, 1u);
                umma_bf16(d_tmem, desc64(a_lo + 3 * a_step, hi), desc64(b_lo + 3 * b_step, hi), idesc, 1u);
              } else {
                umma_bf16_cg2(d_tmem, desc64(a_lo, hi), desc64(b_lo, hi), idesc, first);
                umma_bf16_cg2(d_tmem, desc64(a_lo + a_step, hi), desc64(b_lo + b_step, hi), idesc, 1u);
                umma_bf16_cg2(d_tmem, desc64(a_lo + 2 * a_step, hi), desc64(b_lo + 2 * b_step, hi), idesc, 1u);
                umma_bf16_cg2(d_tmem, desc64(a_lo + 3 * a_step, hi), desc64(b_lo + 3 * b_step, hi), idesc, 1u);
              }
            } else {                 // K tail / the LoRA slab: ceil(kleft / 16) steps
              const int nmma = (kleft + 15) / 16;
              for (int j = 0; j < nmma; ++j) {
                const uint32_t accf = (j > 0) ? 1u : first;
                if constexpr (CG == 1)
                  umma_bf16(d_tmem, desc64(a_lo + j * a_step, hi), desc64(b_lo + j * b_step, hi), idesc, accf);
                else
                  umma_bf16_cg2(d_tmem, desc64(a_lo + j * a_step, hi), desc64(b_lo + j * b_step, hi), idesc, accf);
              }
            }
            // frees the smem slot (in both CTAs when CG == 2) once these MMAs retire
            if constexpr (CG == 1) umma_commit(empty_bar(stage));
            else umma_commit_cg2(empty_bar(stage));
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        if (issuer) {  // accumulator complete -> epilogue warps (of both CTAs)
          if constexpr (CG == 1) umma_commit(tfull_bar(acc));
          else umma_commit_cg2(tfull_bar(acc));
        }
        __syncwarp();
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else if (warp >= 4) {
    // =============================== epilogue (every CTA: its own 128 rows) ===============================
    const int ew = warp - 4;      // 0 .. kEpiWarps-1
    const int q = warp & 3;       // TMEM lane quarter this warp may touch (== warp % 4)
    const int h = ew >> 2;        // which of the quarter's warps
    const int et = threadIdx.x - 128;
    const uint32_t eb = sEpi + ew * kEpiBufs * kEpiBufBytes;
#define B200FT_EPI_CALL(CWV, E)                                                                                       \
  epilogue_loop<CWV, CG, E>(p, &maps.c, &maps.c2, &maps.aux, q, h, lane, et, tmem_base, eb, sBias, aux_bar(ew, 0),    \
                            tfull_bar(0), tempty_bar(0), num_work, w_begin, w_stride, cta_rank)
    if (p.cw == 32) {
      switch (p.epilogue) {
        case B200FT_EPI_RESIDUAL: B200FT_EPI_CALL(32, B200FT_EPI_RESIDUAL); break;
        case B200FT_EPI_GELU: B200FT_EPI_CALL(32, B200FT_EPI_GELU); break;
        case B200FT_EPI_DGELU: B200FT_EPI_CALL(32, B200FT_EPI_DGELU); break;
        case B200FT_EPI_ATOMIC_F32_T: B200FT_EPI_CALL(32, B200FT_EPI_ATOMIC_F32_T); break;
        default: B200FT_EPI_CALL(32, B200FT_EPI_NONE); break;
      }
    } else {   // 16-column chunks: the N = 16 adapter GEMMs and the split-K adapter gradients
      switch (p.epilogue) {
        case B200FT_EPI_RESIDUAL: B200FT_EPI_CALL(16, B200FT_EPI_RESIDUAL); break;
        case B200FT_EPI_GELU: B200FT_EPI_CALL(16, B200FT_EPI_GELU); break;
        case B200FT_EPI_DGELU: B200FT_EPI_CALL(16, B200FT_EPI_DGELU); break;
        case B200FT_EPI_ATOMIC_F32_T: B200FT_EPI_CALL(16, B200FT_EPI_ATOMIC_F32_T); break;
        default: B200FT_EPI_CALL(16, B200FT_EPI_NONE); break;
      }
    }
#undef B200FT_EPI_CALL
  }

  tc_fence_before();
  if constexpr (CG == 1) __syncthreads();
  else cluster_sync_all();  // neither CTA may free TMEM / exit while the other still reads its smem
  if (warp == 2) {
    tc_fence_after();
    if constexpr (CG == 1) tmem_dealloc(tmem_base, kTmemCols);
    else tmem_dealloc_cg2(tmem_base, kTmemCols);
  }
}

// ---------------------------------------------------------------------------------------
// host side: TMA descriptors (cached), tile-shape choice, launch
// ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)f;
  });
  return fn;
}

using MapKey = std::tuple<const void*, uint64_t, uint64_t, uint64_t, uint32_t, uint32_t, int>;
std::mutex g_map_mu;
std::map<MapKey, CUtensorMap> g_maps;

// 2-D bf16 tensor {inner, outer} with row pitch ld (elements), box {box_inner, box_outer}, SW128
int get_tensor_map(const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                   uint32_t box_outer, CUtensorMap* out, int swizzle = 2 /* 0 none, 1 = 64B, 2 = 128B */) {
  MapKey key{ptr, inner, outer, ld, box_inner, box_outer, swizzle};
  {
    std::lock_guard<std::mutex> lk(g_map_mu);
    auto it = g_maps.find(key);
    if (it != g_maps.end()) { *out = it->second; return 0; }
  }
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { b200ft_set_error("gemm: cuTensorMapEncodeTiled unavailable (no CUDA driver?)"); return B200FT_ERR_UNSUPPORTED; }
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle == 2 ? CU_TENSOR_MAP_SWIZZLE_128B : (swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE),
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b200ft_set_error("gemm: cuTensorMapEncodeTiled failed (%d) ptr=%p dims=(%llu,%llu) ld=%llu box=(%u,%u)", (int)r,
                     ptr, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld, box_inner,
                     box_outer);
    return B200FT_ERR_ARG;
  }
  std::lock_guard<std::mutex> lk(g_map_mu);
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps[key] = *out;
  return 0;
}

// units = CTAs (CG == 1) or CTA pairs (CG == 2) working concurrently; tile_m = rows per unit tile
int choose_block_n(int M, int N, int units, int tile_m) {
  if (N <= kMaxBN) return (N + 15) / 16 * 16;
  const int tiles_m = (M + tile_m - 1) / tile_m;
  const int ctas = units < 1 ? 1 : units;
  int best = 256;
  double best_cost = -1.0;
  // time ~ waves x per-tile time; per FLOP a narrower tile re-reads more of A from L2 and keeps the
  // tensor pipe less busy (measured on the step's shapes, profiles/README.md): relative efficiency
  // 1.0 / 0.82 / 0.68 for BN = 256 / 192 / 128.  A narrower tile only wins when it saves more than a whole wave:
  // at 32768 x 768 (384 tiles of BN 256 = 5.2 waves of 74 pairs, 512 of BN 192 = 6.9) BN 256 is still the faster one
  // (K = 768 / 2320 / 3072: 38.0 / 98.1 / 120.1 us against 41.8 / 109.0 / 129.4 with BN 192; profiles/r02_gemm_probe_bn.log).
  const int bns[3] = {256, 192, 128};
  const double eff[3] = {1.0, 0.82, 0.68};
  for (int i = 0; i < 3; ++i) {
    const int bn = bns[i];
    long tiles = (long)tiles_m * ((N + bn - 1) / bn);
    long waves = (tiles + ctas - 1) / ctas;
    double cost = (double)waves * (bn / eff[i] + 16.0);  // + fixed per-tile overhead (pipeline fill, epilogue tail)
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = bn; }
  }
  return best;
}

}  // namespace

extern "C" int b200ft_gemm_bf16(const b200ft_gemm_args* a, void* stream) {
  B200FT_CHECK_ARG(a != nullptr, "gemm: null args");
  B200FT_CHECK_ARG(a->m > 0 && a->n > 0 && a->k1 > 0, "gemm: m=%d n=%d k1=%d", a->m, a->n, a->k1);
  B200FT_CHECK_ARG(a->a1 && a->b1 && a->c, "gemm: null operand");
  const bool atomic_t = a->epilogue == B200FT_EPI_ATOMIC_F32_T;
  if (atomic_t)
    B200FT_CHECK_ARG(a->n % 8 == 0 && a->ldc >= a->m, "gemm: ATOMIC_F32_T needs n %% 8 == 0 and ldc >= m (n=%d m=%d ldc=%lld)",
                     a->n, a->m, (long long)a->ldc);
  else
    B200FT_CHECK_ARG(a->n % 8 == 0 && a->ldc % 8 == 0 && a->ldc >= a->n, "gemm: n=%d ldc=%lld must be multiples of 8", a->n,
                     (long long)a->ldc);
  B200FT_CHECK_ARG(a->lda1 % 8 == 0 && a->ldb1 % 8 == 0, "gemm: lda1=%lld ldb1=%lld must be multiples of 8",
                   (long long)a->lda1, (long long)a->ldb1);
  B200FT_CHECK_ARG(((uintptr_t)a->a1 | (uintptr_t)a->b1 | (uintptr_t)a->c) % 16 == 0, "gemm: 16-byte alignment required");
  // K segments 2..4 (all K-major): a NULL a_i ends the list
  const void* seg_a[kMaxSeg] = {a->a1, a->a2, a->a3, a->a4};
  const void* seg_b[kMaxSeg] = {a->b1, a->b2, a->b3, a->b4};
  const int64_t seg_lda[kMaxSeg] = {a->lda1, a->lda2, a->lda3, a->lda4};
  const int64_t seg_ldb[kMaxSeg] = {a->ldb1, a->ldb2, a->ldb3, a->ldb4};
  const int seg_k[kMaxSeg] = {a->k1, a->k2, a->k3, a->k4};
  int nseg = 1;
  while (nseg < kMaxSeg && seg_a[nseg] != nullptr) ++nseg;
  for (int i = nseg; i < kMaxSeg; ++i)
    B200FT_CHECK_ARG(seg_a[i] == nullptr, "gemm: segment %d given after an empty segment", i + 1);
  for (int i = 1; i < nseg; ++i) {
    B200FT_CHECK_ARG(seg_b[i] && seg_k[i] > 0 && seg_lda[i] % 8 == 0 && seg_ldb[i] % 8 == 0 &&
                         ((uintptr_t)seg_a[i] | (uintptr_t)seg_b[i]) % 16 == 0,
                     "gemm: bad segment %d (k=%d lda=%lld ldb=%lld)", i + 1, seg_k[i], (long long)seg_lda[i],
                     (long long)seg_ldb[i]);
  }
  B200FT_CHECK_ARG(a->epilogue >= 0 && a->epilogue <= 4, "gemm: epilogue=%d", a->epilogue);
  B200FT_CHECK_ARG(a->split_k >= 0 && (a->split_k <= 1 || atomic_t), "gemm: split_k=%d needs the ATOMIC_F32_T epilogue",
                   a->split_k);
  if (a->epilogue == B200FT_EPI_RESIDUAL || a->epilogue == B200FT_EPI_DGELU)
    B200FT_CHECK_ARG(a->aux && a->ld_aux % 8 == 0 && (uintptr_t)a->aux % 16 == 0, "gemm: epilogue needs aux");
  if (a->epilogue == B200FT_EPI_GELU)
    B200FT_CHECK_ARG(a->c2 && a->ldc2 % 8 == 0 && (uintptr_t)a->c2 % 16 == 0, "gemm: GELU epilogue needs c2");
  if (a->bias) B200FT_CHECK_ARG((uintptr_t)a->bias % 16 == 0, "gemm: bias must be 16-byte aligned");

  const int ctas_max = a->max_ctas > 0 ? a->max_ctas : B200FT_NUM_SMS;
  const bool kmajor = (a->layout & (B200FT_A_MN_MAJOR | B200FT_B_MN_MAJOR)) == 0;
  B200FT_CHECK_ARG(a->cta_group >= 0 && a->cta_group <= 2, "gemm: cta_group=%d", a->cta_group);
  int cg = a->cta_group;
  const bool b_mn_flag = (a->layout & B200FT_B_MN_MAJOR) != 0;
  (void)kmajor;
  if (cg == 0) cg = (!atomic_t && a->m > kBlockM && a->n >= 128 && ctas_max >= 2) ? 2 : 1;
  B200FT_CHECK_ARG(cg == 1 || (!atomic_t && ctas_max >= 2),
                   "gemm: cta_group=2 needs a non-atomic epilogue and >= 2 CTAs");
  int BN = a->block_n > 0 ? a->block_n : choose_block_n(a->m, a->n, ctas_max / cg, kBlockM * cg);
  // an MN-major B operand of a CTA pair is staged in 64-column boxes per CTA: BN / 2 must be a multiple of 64
  if (cg == 2 && b_mn_flag && BN % 128 != 0) BN = a->block_n > 0 ? BN : (a->n > 128 ? 256 : 128);
  B200FT_CHECK_ARG(BN % 16 == 0 && BN >= 16 && BN <= kMaxBN, "gemm: block_n=%d", BN);
  B200FT_CHECK_ARG(!(cg == 2 && b_mn_flag) || BN % 128 == 0, "gemm: cta_group=2 with MN-major B needs block_n 128 or 256 (got %d)", BN);

  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = a->m; p.N = a->n; p.BN = BN;
  p.nseg = nseg;
  p.kb_prefix[0] = 0;
  for (int i = 0; i < kMaxSeg; ++i) {
    p.k[i] = i < nseg ? seg_k[i] : 0;
    p.kb_prefix[i + 1] = p.kb_prefix[i] + (p.k[i] + kBlockK - 1) / kBlockK;
  }
  p.a_mn = (a->layout & B200FT_A_MN_MAJOR) ? 1 : 0;
  p.b_mn = (a->layout & B200FT_B_MN_MAJOR) ? 1 : 0;
  p.tiles_m = (a->m + kBlockM * cg - 1) / (kBlockM * cg);
  p.tiles_n = (a->n + BN - 1) / BN;
  {
    // Rasterisation.  By default consecutive work items share an M tile and walk N: each A tile is fetched
    // once, by all CTAs of a wave at the same time, and B is cycled through L2 once per M tile.  That is
    // right while B is the smaller operand.  When A is the smaller one (4096-token Llama steps against
    // 100-500 MB weight matrices; the lm_head against the vocabulary) walk M instead, so that every B
    // tile is fetched once and A stays resident.  ncu on 4096 x 28672 x 4096 with the N walk: 3.73 GB of
    // DRAM reads for 0.27 GB of operands; with the M walk the same GEMM goes from 1168 to 1490 TFLOP/s.
    int64_t k_total = 0;
    for (int i = 0; i < nseg; ++i) k_total += seg_k[i];
    const int64_t a_bytes = (int64_t)a->m * k_total * 2, b_bytes = (int64_t)a->n * k_total * 2;
    p.raster_m = (b_bytes > ((int64_t)32 << 20) && a_bytes < b_bytes) ? 1 : 0;
    static const char* force = getenv("B200FT_GEMM_RASTER");  // "m" / "n": measurement aid only
    if (force && (force[0] == 'm' || force[0] == 'n')) p.raster_m = force[0] == 'm';
  }
  p.epilogue = a->epilogue;
  {
    const int b_rows = BN / cg;
    int b_stage = b_rows * kBlockK * 2;                                  // K-major slab (every segment after the first)
    if (p.b_mn) { const int mn = (b_rows + 63) / 64 * 8192; if (mn > b_stage) b_stage = mn; }
    b_stage = (b_stage + 1023) / 1024 * 1024;                            // 128B-swizzle atoms are 1024-byte aligned
    int st = kTileBytes / (kAStageBytes + b_stage);
    p.n_stages = st < 2 ? 2 : (st > kMaxStages ? kMaxStages : st);
    p.b_stage_bytes = b_stage;
    B200FT_CHECK_ARG(p.n_stages * (kAStageBytes + b_stage) <= kTileBytes, "gemm: operand ring does not fit (block_n=%d)", BN);
  }
  p.alpha = a->alpha;
  p.bias = a->bias;
  p.aux = (const __nv_bfloat16*)a->aux; p.ld_aux = a->ld_aux;
  p.c = (__nv_bfloat16*)a->c; p.ldc = a->ldc;
  p.c2 = (__nv_bfloat16*)a->c2; p.ldc2 = a->ldc2;
  p.cf = (float*)a->c;
  {
    const int kb_total = p.kb_prefix[nseg];
    int splits = a->split_k > 0 ? a->split_k : 1;
    if (atomic_t && a->split_k == 0) {  // fill the machine: tiles * splits ~ one wave of CTAs
      const int tiles = p.tiles_m * p.tiles_n;
      splits = ctas_max / (tiles < 1 ? 1 : tiles);
    }
    if (splits < 1) splits = 1;
    if (splits > kb_total) splits = kb_total;
    p.kb_per_split = (kb_total + splits - 1) / splits;
    p.splits = (kb_total + p.kb_per_split - 1) / p.kb_per_split;  // no empty split
  }

  p.cw = (BN % 32 == 0) ? 32 : 16;
  const int esw = p.cw == 32 ? 1 : 0;  // staging tile swizzle: 64B for the 32-column tile, none for 16
  TmapPack maps;
  int rc;
  if (p.a_mn) rc = get_tensor_map(a->a1, (uint64_t)a->m, (uint64_t)a->k1, (uint64_t)a->lda1, 64, 64, &maps.a[0]);
  else        rc = get_tensor_map(a->a1, (uint64_t)a->k1, (uint64_t)a->m, (uint64_t)a->lda1, 64, kBlockM, &maps.a[0]);
  if (rc) return rc;
  if (p.b_mn) rc = get_tensor_map(a->b1, (uint64_t)a->n, (uint64_t)a->k1, (uint64_t)a->ldb1, 64, 64, &maps.b[0]);
  else        rc = get_tensor_map(a->b1, (uint64_t)a->k1, (uint64_t)a->n, (uint64_t)a->ldb1, 64, (uint32_t)(BN / cg), &maps.b[0]);
  if (rc) return rc;
  for (int i = 1; i < kMaxSeg; ++i) {
    if (i < nseg) {
      rc = get_tensor_map(seg_a[i], (uint64_t)seg_k[i], (uint64_t)a->m, (uint64_t)seg_lda[i], 64, kBlockM, &maps.a[i]);
      if (rc) return rc;
      rc = get_tensor_map(seg_b[i], (uint64_t)seg_k[i], (uint64_t)a->n, (uint64_t)seg_ldb[i], 64, (uint32_t)(BN / cg), &maps.b[i]);
      if (rc) return rc;
    } else {
      maps.a[i] = maps.a[0]; maps.b[i] = maps.b[0];  // never dereferenced; keep the block initialised
    }
  }
  // epilogue staging tiles: {cw columns, 32 rows}
  if (atomic_t) {
    maps.c = maps.a[0];  // unused by this epilogue; any valid descriptor
  } else {
    rc = get_tensor_map(a->c, (uint64_t)a->n, (uint64_t)a->m, (uint64_t)a->ldc, (uint32_t)p.cw, 32, &maps.c, esw);
    if (rc) return rc;
  }
  maps.c2 = maps.c; maps.aux = maps.c;
  if (a->epilogue == B200FT_EPI_GELU) {
    rc = get_tensor_map(a->c2, (uint64_t)a->n, (uint64_t)a->m, (uint64_t)a->ldc2, (uint32_t)p.cw, 32, &maps.c2, esw);
    if (rc) return rc;
  }
  if (a->epilogue == B200FT_EPI_RESIDUAL || a->epilogue == B200FT_EPI_DGELU) {
    rc = get_tensor_map(a->aux, (uint64_t)a->n, (uint64_t)a->m, (uint64_t)a->ld_aux, (uint32_t)p.cw, 32, &maps.aux, esw);
    if (rc) return rc;
  }

  // the opt-in shared-memory size is a per-device function attribute
  static thread_local int attr_dev = -1;
  int dev = -1;
  B200FT_CUDA(cudaGetDevice(&dev));
  if (dev != attr_dev) {
    B200FT_CUDA(cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    B200FT_CUDA(cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_dev = dev;
  }
  const int num_work = p.tiles_m * p.tiles_n * p.splits;
  if (cg == 1) {
    const int grid = num_work < ctas_max ? num_work : ctas_max;
    B200FT_CUDA(launch_k(gemm_bf16_tcgen05_kernel<1>, grid, kThreads, kSmemBytes, (cudaStream_t)stream, maps, p));
    return 0;
  }
  // cta_group::2: clusters of two CTAs (one TPC); work items go to pairs
  const int pairs_max = ctas_max / 2;
  const int pairs = num_work < pairs_max ? num_work : pairs_max;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemBytes;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = b200ft_pdl_enabled() ? 2 : 1;
  B200FT_CUDA(cudaLaunchKernelEx(&cfg, gemm_bf16_tcgen05_kernel<2>, maps, p));
  return 0;
}

#ifdef B200FT_ATTN_TIMING
extern "C" int b200ft_debug_gemm_timing(long long* log_out, int* counts_out, int reset) {
  if (reset) {
    int z[2] = {0, 0};
    B200FT_CUDA(cudaMemcpyToSymbol(g_gcnt, z, sizeof(z)));
    return 0;
  }
  B200FT_CUDA(cudaDeviceSynchronize());
  B200FT_CUDA(cudaMemcpyFromSymbol(log_out, g_glog, sizeof(long long) * 2 * 8192 * 2));
  B200FT_CUDA(cudaMemcpyFromSymbol(counts_out, g_gcnt, sizeof(int) * 2));
  return 0;
}
#endif
