// Causal flash-attention forward on tcgen05 for d_head = 64 (GPT-2, Llama-3.2-1B), reading q, k, v
// straight out of the fused [tokens, q|k|v] projection buffer (grouped-query attention = several
// q heads sharing one kv head) and writing o [tokens, heads*64] plus the per-row log-sum-exp.
//
// One CTA per (batch, head, 128-row query tile), heaviest tiles first; 192 threads:
//   warp 0   TMA producer: Q once, then K_j / V_j into a 2-stage ring (and TMEM alloc/dealloc)
//   warp 1   MMA issuer:   S = Q K_j^T (128x128x64, 4 UMMAs) into TMEM columns [0,128);
//                          PV = P V_j (128x64x128, 8 UMMAs) into TMEM columns [128,192)
//   warps 2-5 softmax, one thread per query row: two sweeps over S in TMEM (row max; then
//            exp2, row sum, bf16 P into 128B-swizzled smem = the A operand of the PV UMMA), and the
//            running output O kept in registers: O = O*alpha + PV_{j-1} folded in one tile later.
// V_j is consumed as an MN-major B operand ([keys, d_head] rows as TMA delivers them), so nothing
// is transposed.  Shared memory 113 KB and 256 TMEM columns per CTA: two CTAs share an SM, and one
// CTA's softmax (MUFU-bound: 128x128 exp2 per tile) overlaps the other's UMMAs.
//
// Work per (batch, head): T^2/2 * 64 * 4 FLOP (causal); tensor-pipe time 512 cycles per 128x128
// tile vs 1024 cycles of exp2 at 16/clk/SM: the kernel is bound by the special-function unit.
#include <cuda.h>
#include <mutex>
#include <stdlib.h>
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

constexpr int kDH = 64;
constexpr int kBQ = 128, kBK = 128;
constexpr int kThreads = 192;
constexpr int kQBytes = kBQ * kDH * 2;          // 16 KB
constexpr int kKVBytes = kBK * kDH * 2;         // 16 KB each
constexpr int kPBytes = kBQ * kBK * 2;          // 32 KB: two 64-key slabs of [128 rows x 128 B]
constexpr int kKStages = 2;                     // K_j double-buffered; V_j single (it is needed a whole softmax later)
constexpr int kTiles = kQBytes + (kKStages + 1) * kKVBytes + kPBytes;  // 96 KB: two CTAs per SM
constexpr int kSmemBytes = kTiles + 256;        // + barriers; no alignment slack (the kernel checks the base)
constexpr int kTmemCols = 256;
constexpr uint32_t kColS = 0, kColPV = 128;
constexpr long long kWaitLimit = 4000000000ll;

struct AttnParams {
  int batch, seq, n_head, n_kv_head;
  int q_off, k_off, v_off;  // column offsets of head 0 of q / k / v in the fused buffer
  int nq_tiles;
  float scale_log2;         // softmax scale * log2(e)
  __nv_bfloat16* out; int64_t ld_out;
  float* lse2;              // [batch*n_head, seq]: row max + log2(row sum), log2 domain incl. scale
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
// smem matrix descriptor words (SW128): lo = start>>4 | LBO>>4 << 16, hi = SBO>>4 | version 1 | swizzle 128B
__device__ __forceinline__ uint64_t sdesc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  const uint32_t lo = ((addr & 0x3FFFFu) >> 4) | (((lbo >> 4) & 0x3FFFu) << 16);
  const uint32_t hi = ((sbo >> 4) & 0x3FFFu) | (1u << 14) | (2u << 29);
  return ((uint64_t)hi << 32) | lo;
}
// kind::f16 instruction descriptor: fp32 accumulate, bf16 x bf16, M = 128
__device__ __forceinline__ uint32_t idesc(int n, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn << 16) | ((uint32_t)(n >> 3) << 17) | (8u << 24);
}

// 160 registers: 6 warps x 160 x 32 = 30720 per CTA, two CTAs per SM (168 rounds up past the file)
__global__ void __maxnreg__(160)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tm, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  if (raw & 1023u) __trap();  // dynamic shared memory starts 1024-aligned (no static smem in this kernel)
  const uint32_t base = raw;
  const uint32_t sQ = base;
  const uint32_t sK = sQ + kQBytes;
  const uint32_t sV = sK + kKStages * kKVBytes;
  const uint32_t sP = sV + kKVBytes;
  const uint32_t bars = base + kTiles;
  const uint32_t q_full = bars, k_full = bars + 8, k_empty = bars + 24, s_full = bars + 40, s_free = bars + 48,
                 p_full = bars + 56, pv_full = bars + 64, v_full = bars + 72, v_empty = bars + 80, tmem_slot = bars + 88;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heaviest query tiles (most kv tiles) first
  const int bh_count = p.batch * p.n_head;
  const int qt = p.nq_tiles - 1 - (int)(blockIdx.x / bh_count);
  const int bh = (int)(blockIdx.x % bh_count);
  const int b = bh / p.n_head, h = bh % p.n_head;
  const int hk = h / (p.n_head / p.n_kv_head);
  const int q0 = qt * kBQ;
  const int n_kv = qt + 1;  // causal: kv tiles 0..qt

  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < kKStages; ++s) { mbar_init(k_full + 8 * s, 1); mbar_init(k_empty + 8 * s, 1); }
    mbar_init(v_full, 1); mbar_init(v_empty, 1);
    mbar_init(s_full, 1); mbar_init(s_free, 4); mbar_init(p_full, 4); mbar_init(pv_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(kTmemCols) : "memory");
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
      mbar_expect_tx(q_full, kQBytes);
      tma_load_3d(sQ, &tm, p.q_off + h * kDH, q0, b, q_full);
    }
    int stage = 0;
    uint32_t phase = 0;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(k_empty + 8 * stage, phase ^ 1u);
      if (issuer) {
        mbar_expect_tx(k_full + 8 * stage, kKVBytes);
        tma_load_3d(sK + stage * kKVBytes, &tm, p.k_off + hk * kDH, j * kBK, b, k_full + 8 * stage);
      }
      __syncwarp();
      mbar_wait(v_empty, (uint32_t)(j & 1) ^ 1u);  // P V_{j-1} has consumed the V buffer
      if (issuer) {
        mbar_expect_tx(v_full, kKVBytes);
        tma_load_3d(sV, &tm, p.v_off + hk * kDH, j * kBK, b, v_full);
      }
      __syncwarp();
      if (++stage == kKStages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    const bool issuer = elect_one();
    const uint32_t id_s = idesc(kBK, 0);   // S: N = 128 keys, K-major B (K_j rows = keys, d_head contiguous)
    const uint32_t id_pv = idesc(kDH, 1);  // PV: N = 64, MN-major B (V_j rows = keys = reduction index)
    mbar_wait(q_full, 0);
    int stage = 0;
    uint32_t phase = 0;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(k_full + 8 * stage, phase);
      mbar_wait(s_free, (uint32_t)(j & 1) ^ 1u);  // softmax has drained S_{j-1}
      tc_fence_after();
      if (issuer) {
#pragma unroll
        for (int k = 0; k < kDH / 16; ++k)
          umma(tmem + kColS, sdesc(sQ + 32 * k, 16, 1024), sdesc(sK + stage * kKVBytes + 32 * k, 16, 1024), id_s, k > 0);
        umma_commit(s_full);
        umma_commit(k_empty + 8 * stage);
      }
      __syncwarp();
      mbar_wait(v_full, (uint32_t)(j & 1));
      mbar_wait(p_full, (uint32_t)(j & 1));  // P_j is in smem and PV_{j-1} has been read out of TMEM
      tc_fence_after();
      if (issuer) {
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k)  // reduction over keys: P slab k/4, 32 B per step; V rows 16 per step
          umma(tmem + kColPV, sdesc(sP + (k >> 2) * (kBQ * 128) + 32 * (k & 3), 16, 1024),
               sdesc(sV + 2048 * k, 8192, 1024), id_pv, k > 0);
        umma_commit(pv_full);
        umma_commit(v_empty);
      }
      __syncwarp();
      if (++stage == kKStages) { stage = 0; phase ^= 1u; }
    }
  } else {
    // ================================ softmax, one thread per query row ================================
    const int quarter = warp & 3;                 // TMEM lanes this warp may touch
    const int row = quarter * 32 + lane;          // row within the query tile
    const uint32_t t_lane = tmem + ((uint32_t)(quarter * 32) << 16);
    const int qrow = q0 + row;                    // position in the sequence
    float o[kDH];
#pragma unroll
    for (int i = 0; i < kDH; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(s_full, (uint32_t)(j & 1));
      tc_fence_after();
      const bool diag = (j == qt);
      const int kbase = j * kBK;
      // ---- sweep 1: row max of the raw scores ----
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < kBK / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(t_lane + kColS + 32 * c, r);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float s = __uint_as_float(r[i]);
          mx = fmaxf(mx, (diag && kbase + 32 * c + i > qrow) ? -INFINITY : s);
        }
      }
      const float m_new = fmaxf(m, mx * p.scale_log2);
      const float alpha = ex2(m - m_new);  // exp2(-inf) = 0 on the first tile
      // ---- fold the previous tile's P V into O (it was computed against the old max), then rescale ----
      if (j > 0) {
        mbar_wait(pv_full, (uint32_t)((j - 1) & 1));
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < kDH / 32; ++c) {
          uint32_t r[32];
          tmem_ld32(t_lane + kColPV + 32 * c, r);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[32 * c + i] = (o[32 * c + i] + __uint_as_float(r[i])) * alpha;
        }
      }
      l *= alpha;
      m = m_new;
      // ---- sweep 2: p = exp2(s*scale - m), row sum, bf16 P into the swizzled A-operand tile ----
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < kBK / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(t_lane + kColS + 32 * c, r);
        tmem_wait_ld();
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = ex2(fmaf(__uint_as_float(r[i]), p.scale_log2, -m));
          float p1 = ex2(fmaf(__uint_as_float(r[i + 1]), p.scale_log2, -m));
          if (diag) {
            if (kbase + 32 * c + i > qrow) p0 = 0.f;
            if (kbase + 32 * c + i + 1 > qrow) p1 = 0.f;
          }
          sum += p0 + p1;
          w[i >> 1] = pack_bf16x2(p0, p1);
        }
        // keys 32c..32c+31 of this row: slab = c/2, 16-byte chunks 4*(c&1) .. +3 of the 128-byte row
        const uint32_t rowaddr = sP + (uint32_t)(c >> 1) * (kBQ * 128) + (uint32_t)row * 128;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const uint32_t chunk = (uint32_t)(4 * (c & 1) + v);
          asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(rowaddr + ((chunk ^ (uint32_t)(row & 7)) << 4)),
                       "r"(w[4 * v]), "r"(w[4 * v + 1]), "r"(w[4 * v + 2]), "r"(w[4 * v + 3]) : "memory");
        }
      }
      l += sum;
      tc_fence_before();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) { mbar_arrive(s_free); mbar_arrive(p_full); }
    }
    // ---- last tile's P V, normalise, write o and the log-sum-exp ----
    mbar_wait(pv_full, (uint32_t)((n_kv - 1) & 1));
    tc_fence_after();
    const float inv_l = 1.f / l;
#pragma unroll
    for (int c = 0; c < kDH / 32; ++c) {
      uint32_t r[32];
      tmem_ld32(t_lane + kColPV + 32 * c, r);
      tmem_wait_ld();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[32 * c + i] = (o[32 * c + i] + __uint_as_float(r[i])) * inv_l;
    }
    if (qrow < p.seq) {
      __nv_bfloat16* dst = p.out + ((int64_t)b * p.seq + qrow) * p.ld_out + h * kDH;
#pragma unroll
      for (int v = 0; v < kDH / 8; ++v) *reinterpret_cast<uint4*>(dst + 8 * v) = pack8(o + 8 * v);
      p.lse2[(int64_t)bh * p.seq + qrow] = m + log2f(l);
    }
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
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

}  // namespace

extern "C" int b200ft_attn_fwd(const void* qkv, int64_t ld, int q_off, int k_off, int v_off, void* out,
                               int64_t ld_out, float* lse2, int batch, int seq, int n_head, int n_kv_head,
                               int d_head, float scale, void* stream) {
  B200FT_CHECK_ARG(qkv && out && lse2, "attn_fwd: null buffer");
  if (d_head != kDH) {
    b200ft_set_error("attn_fwd: d_head=%d is not built (only 64)", d_head);
    return B200FT_ERR_UNSUPPORTED;
  }
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
  cuuint64_t dims[3] = {(cuuint64_t)ld, (cuuint64_t)seq, (cuuint64_t)batch};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)ld * 2 * (cuuint64_t)seq};
  cuuint32_t box[3] = {kDH, kBQ, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(qkv), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { b200ft_set_error("attn_fwd: cuTensorMapEncodeTiled failed (%d)", (int)r); return B200FT_ERR_ARG; }
  AttnParams p;
  p.batch = batch; p.seq = seq; p.n_head = n_head; p.n_kv_head = n_kv_head;
  p.q_off = q_off; p.k_off = k_off; p.v_off = v_off;
  p.nq_tiles = (seq + kBQ - 1) / kBQ;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.out = (__nv_bfloat16*)out; p.ld_out = ld_out; p.lse2 = lse2;
  static thread_local int attr_dev = -1;
  int dev = -1;
  B200FT_CUDA(cudaGetDevice(&dev));
  if (dev != attr_dev) {
    B200FT_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_dev = dev;
  }
  static int occ_reported = 0;
  if (!occ_reported && getenv("B200FT_DEBUG")) {
    int nb = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, attn_fwd_kernel, kThreads, kSmemBytes);
    fprintf(stderr, "[b200ft] attn_fwd_kernel: %d CTAs/SM (smem %d B)\n", nb, kSmemBytes);
    occ_reported = 1;
  }
  const unsigned grid = (unsigned)(p.nq_tiles * batch * n_head);
  attn_fwd_kernel<<<grid, kThreads, kSmemBytes, (cudaStream_t)stream>>>(tm, p);
  B200FT_LAUNCH_CHECK();
  return 0;
}
