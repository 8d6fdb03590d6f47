// Kernels only the full-parameter fine-tune needs (BASELINE.json configs[4]; SURVEY.md §8 a2.5
// "full-FT adds dW" and a2.9): the RMSNorm gain gradient and the embedding-table gradient.
// Both are HBM-bound column/row reductions; the weight gradients proper are tcgen05 GEMMs.
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

// dgamma[c] += sum_r dy[r,c] * x[r,c] * rstd[r].  A block owns 256 columns x one row chunk: a lane
// carries 8 columns (one 16-byte load per operand and row, 512 contiguous bytes per warp), the 8 warps
// stride the rows, partial sums meet in shared memory and leave as one fp32 atomic per column and block.
__global__ void __launch_bounds__(256)
rmsnorm_dgamma_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                      const float* __restrict__ rstd, float* __restrict__ dgamma, int64_t rows, int d,
                      int64_t rows_per_block) {
  grid_dep_sync();
  __shared__ float part[8][256 + 8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 256 + 8 * lane;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_block;
  const int64_t r1 = min(rows, r0 + rows_per_block);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c < d) {
#pragma unroll 4
    for (int64_t r = r0 + warp; r < r1; r += 8) {
      const float rs = __ldg(rstd + r);
      float a[8], b[8];
      unpack8(ld_stream16(dy + r * d + c), a);
      unpack8(ld_stream16(x + r * d + c), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = fmaf(a[j] * b[j], rs, acc[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[warp][8 * lane + j] = acc[j];
  __syncthreads();
  const int cc = blockIdx.x * 256 + threadIdx.x;
  if (cc < d) {
    float s_ = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s_ += part[w][threadIdx.x];
    atomicAdd(dgamma + cc, s_);
  }
}

// dwte[ids[r], :] += dx[r, :] without atomics and with fp32 accumulation.  One CTA per token row r.  The CTA first
// looks for an EARLIER row with the same id: if there is one, that row's CTA (the id's "leader") does the work and this
// one leaves.  A leader gathers every later row with its id (ascending), adds their dx rows in fp32 registers — each
// thread owns 8 columns per pass — and adds the sum to the table row once, in bf16: one rounding however often a
// token occurs (real text repeats some ids hundreds of times per step; bf16 atomics would drop the small addends and
// depend on arrival order), and the same bits every run.  Scanning the ids costs rows^2 * 8 bytes of L2 reads in
// total: 134 MB at the 4096 rows of a full-parameter micro-batch.
constexpr int kEmbThreads = 256;
constexpr int kEmbMaxDup = 2048;   // duplicates of one id gathered per pass of the list (more: further passes)
constexpr int kEmbVec = 4;         // 16-byte vectors per thread: d <= 256 * 4 * 8 = 8192

__global__ void __launch_bounds__(kEmbThreads)
embed_bwd_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ dx,
                 __nv_bfloat16* __restrict__ dwte, int64_t rows, int d, int vocab) {
  grid_dep_sync();
  __shared__ int s_found;
  __shared__ int s_cnt;
  __shared__ int s_list[kEmbMaxDup];
  __shared__ int s_warp[kEmbThreads / 32];
  const int64_t row = blockIdx.x;
  const int tid = threadIdx.x;
  auto clamp = [&](int64_t id) { return id < 0 ? (int64_t)0 : (id >= vocab ? (int64_t)vocab - 1 : id); };  // as the forward gather
  const int64_t id = clamp(ids[row]);
  if (tid == 0) { s_found = 0; s_cnt = 0; }
  __syncthreads();
  // ---- an earlier row with this id?  (chunked so that the CTA can stop at the first hit)
  for (int64_t base = 0; base < row; base += kEmbThreads * 8) {
    int hit = 0;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int64_t r = base + (int64_t)u * kEmbThreads + tid;
      if (r < row && clamp(ids[r]) == id) hit = 1;
    }
    if (hit) s_found = 1;
    __syncthreads();
    if (s_found) return;
  }
  // ---- leader: this row, then every later row with the same id, kEmbMaxDup of them per pass of the list, in
  // ascending order; the fp32 sums stay in registers across the passes (a thread owns up to kEmbVec 16-byte vectors)
  const int nvec = d >> 3;
  float acc[kEmbVec][8];
#pragma unroll
  for (int vi = 0; vi < kEmbVec; ++vi)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[vi][j] = 0.f;
  int64_t next = row;            // first row not yet looked at
  while (next < rows) {
    __syncthreads();
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    // gather matches in [next, rows) until the list is full; a chunk is taken whole or not at all, so order is kept
    int64_t base = next;
    for (; base < rows; base += kEmbThreads) {
      if (s_cnt + kEmbThreads > kEmbMaxDup) break;
      const int64_t r = base + tid;
      const bool m = r < rows && (r == row || clamp(ids[r]) == id);
      // ordered compaction inside the chunk: warp ballots + a prefix over the 8 warps
      const unsigned bal = __ballot_sync(0xffffffffu, m);
      if ((tid & 31) == 0) s_warp[tid >> 5] = __popc(bal);
      __syncthreads();
      int off = s_cnt;
      for (int w = 0; w < (tid >> 5); ++w) off += s_warp[w];
      if (m) s_list[off + __popc(bal & ((1u << (tid & 31)) - 1u))] = (int)(r - row);
      __syncthreads();
      if (tid == 0) { int t = 0; for (int w = 0; w < kEmbThreads / 32; ++w) t += s_warp[w]; s_cnt += t; }
      __syncthreads();
    }
    next = base;
    const int cnt = s_cnt;
#pragma unroll
    for (int vi = 0; vi < kEmbVec; ++vi) {
      const int v0 = tid + vi * kEmbThreads;
      if (v0 < nvec) {
        for (int k = 0; k < cnt; ++k) {
          float f[8];
          unpack8(ld_stream16(dx + (row + s_list[k]) * d + (v0 << 3)), f);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[vi][j] += f[j];
        }
      }
    }
  }
#pragma unroll
  for (int vi = 0; vi < kEmbVec; ++vi) {
    const int v0 = tid + vi * kEmbThreads;
    if (v0 < nvec) {
      __nv_bfloat16* dst = dwte + id * d + (v0 << 3);
      float t[8];
      unpack8(*reinterpret_cast<const uint4*>(dst), t);   // what is there already (a tied lm_head's gradient)
#pragma unroll
      for (int j = 0; j < 8; ++j) t[j] += acc[vi][j];
      *reinterpret_cast<uint4*>(dst) = pack8(t);
    }
  }
}

}  // namespace

extern "C" int b200ft_rmsnorm_bwd_dgamma(const void* dy, const void* x, const float* rstd, float* dgamma,
                                         int64_t rows, int d, void* stream) {
  B200FT_CHECK_ARG(dy && x && rstd && dgamma && rows > 0 && d > 0 && d % 8 == 0 &&
                       ((uintptr_t)dy | (uintptr_t)x) % 16 == 0,
                   "rmsnorm_bwd_dgamma: rows=%lld d=%d (d must be a multiple of 8, buffers 16-byte aligned)",
                   (long long)rows, d);
  const int col_blocks = (d + 255) / 256;
  // ~4 blocks per SM in total; at least 64 rows per block so the atomics stay negligible
  int64_t chunks = (4LL * B200FT_NUM_SMS + col_blocks - 1) / col_blocks;
  const int64_t max_chunks = (rows + 63) / 64;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  if (chunks > 65535) chunks = 65535;
  const int64_t rpb = (rows + chunks - 1) / chunks;
  dim3 grid(col_blocks, (unsigned)((rows + rpb - 1) / rpb));
  launch_k(rmsnorm_dgamma_kernel, grid, 256, 0, (cudaStream_t)stream,
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, rstd, dgamma, rows, d, rpb);
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_embed_bwd(const int64_t* ids, const void* dx, void* dwte, int64_t rows, int d, int vocab,
                                void* stream) {
  B200FT_CHECK_ARG(ids && dx && dwte && rows > 0 && rows < (1ll << 31) && d > 0 && d % 8 == 0 && d <= 8192 && vocab > 0 &&
                       ((uintptr_t)dx | (uintptr_t)dwte) % 16 == 0,
                   "embed_bwd: rows=%lld d=%d (d must be a multiple of 8 and <= 8192, buffers 16-byte aligned)", (long long)rows, d);
  launch_k(embed_bwd_kernel, (unsigned)rows, kEmbThreads, 0, (cudaStream_t)stream,
      ids, (const __nv_bfloat16*)dx, (__nv_bfloat16*)dwte, rows, d, vocab);
  B200FT_LAUNCH_CHECK();
  return 0;
}
