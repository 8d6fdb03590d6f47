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

// dwte[ids[r], :] += dx[r, :]; one warp per token row, bf16x2 atomics (rows of the same token collide)
__global__ void __launch_bounds__(256)
embed_bwd_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ dx,
                 __nv_bfloat16* __restrict__ dwte, int64_t rows, int d, int vocab) {
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  int64_t id = ids[row];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);  // same clamp as the forward gather
  const __nv_bfloat162* src = reinterpret_cast<const __nv_bfloat162*>(dx + row * d);
  __nv_bfloat162* dst = reinterpret_cast<__nv_bfloat162*>(dwte + id * d);
  for (int i = lane; i < (d >> 1); i += 32) atomicAdd(dst + i, src[i]);
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
  rmsnorm_dgamma_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, rstd, dgamma, rows, d, rpb);
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_embed_bwd(const int64_t* ids, const void* dx, void* dwte, int64_t rows, int d, int vocab,
                                void* stream) {
  B200FT_CHECK_ARG(ids && dx && dwte && rows > 0 && d > 0 && d % 2 == 0 && vocab > 0, "embed_bwd: bad arguments");
  embed_bwd_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
      ids, (const __nv_bfloat16*)dx, (__nv_bfloat16*)dwte, rows, d, vocab);
  B200FT_LAUNCH_CHECK();
  return 0;
}
