// Small HBM-bound helpers around the step: embedding gather, bf16 transpose (weight
// re-layout at load time and the tiny adapter shadows), fp32 -> bf16 cast.
#include "common.cuh"
#include "../../include/b200ft.h"

namespace {

// one warp per token row; 16-byte gathers from wte[id] and wpe[pos]
__global__ void __launch_bounds__(256)
embed_fwd_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ wte,
                 const __nv_bfloat16* __restrict__ wpe, __nv_bfloat16* __restrict__ x, int64_t rows,
                 int seq, int d, int vocab) {
  grid_dep_sync();
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  int64_t id = ids[row];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);  // out-of-range ids are clamped, never read OOB
  const int pos = (int)(row % seq);
  const __nv_bfloat16* te = wte + id * d;
  const __nv_bfloat16* pe = wpe ? wpe + (int64_t)pos * d : nullptr;
  for (int vi = lane; vi < (d >> 3); vi += 32) {
    const uint4 tv = __ldg(reinterpret_cast<const uint4*>(te + (vi << 3)));
    if (pe) {
      float a[8], b[8];
      unpack8(tv, a);
      unpack8(__ldg(reinterpret_cast<const uint4*>(pe + (vi << 3))), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] += b[j];
      st_stream16(x + row * d + (vi << 3), pack8(a));
    } else {
      st_stream16(x + row * d + (vi << 3), tv);
    }
  }
}

// 32x32 tiles through padded shared memory: coalesced on both sides
__global__ void __launch_bounds__(256)
transpose_bf16_kernel(const __nv_bfloat16* __restrict__ in, int64_t ld_in, __nv_bfloat16* __restrict__ out,
                      int64_t ld_out, int rows, int cols) {
  grid_dep_sync();
  __shared__ __nv_bfloat16 tile[32][34];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    int r = r0 + ty + j, c = c0 + tx;
    if (r < rows && c < cols) tile[ty + j][tx] = in[(int64_t)r * ld_in + c];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    int c = c0 + ty + j, r = r0 + tx;
    if (r < rows && c < cols) out[(int64_t)c * ld_out + r] = tile[tx][ty + j];
  }
}

__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, int64_t n) {
  grid_dep_sync();
  const int64_t nvec = n >> 3;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const float4 a = *reinterpret_cast<const float4*>(in + (i << 3));
    const float4 b = *reinterpret_cast<const float4*>(in + (i << 3) + 4);
    const float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    *reinterpret_cast<uint4*>(out + (i << 3)) = pack8(f);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (int64_t e = nvec << 3; e < n; ++e) out[e] = __float2bfloat16_rn(in[e]);
}

}  // namespace

extern "C" int b200ft_embed_fwd(const int64_t* ids, const void* wte, const void* wpe, void* x,
                                int64_t rows, int seq, int d, int vocab, void* stream) {
  B200FT_CHECK_ARG(ids && wte && x, "embed_fwd: null buffer");
  B200FT_CHECK_ARG(rows > 0 && seq > 0 && d > 0 && d % 8 == 0 && vocab > 0,
                   "embed_fwd: rows=%lld seq=%d d=%d vocab=%d", (long long)rows, seq, d, vocab);
  launch_k(embed_fwd_kernel, (unsigned)((rows + 7) / 8), 256, 0, (cudaStream_t)stream,
      ids, (const __nv_bfloat16*)wte, (const __nv_bfloat16*)wpe, (__nv_bfloat16*)x, rows, seq, d, vocab);
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_transpose_bf16(const void* in, int64_t ld_in, void* out, int64_t ld_out,
                                     int rows, int cols, void* stream) {
  B200FT_CHECK_ARG(in && out && rows > 0 && cols > 0 && ld_in >= cols && ld_out >= rows,
                   "transpose_bf16: rows=%d cols=%d ld_in=%lld ld_out=%lld", rows, cols,
                   (long long)ld_in, (long long)ld_out);
  dim3 grid((cols + 31) / 32, (rows + 31) / 32);
  launch_k(transpose_bf16_kernel, grid, 256, 0, (cudaStream_t)stream,
      (const __nv_bfloat16*)in, ld_in, (__nv_bfloat16*)out, ld_out, rows, cols);
  B200FT_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200ft_zero(void* ptr, size_t bytes, void* stream) {
  B200FT_CHECK_ARG(ptr || bytes == 0, "zero: null buffer");
  if (bytes) B200FT_CUDA(cudaMemsetAsync(ptr, 0, bytes, (cudaStream_t)stream));
  return 0;
}

extern "C" int b200ft_cast_f32_to_bf16(const float* in, void* out, int64_t n, void* stream) {
  B200FT_CHECK_ARG(in && out && n >= 0, "cast_f32_to_bf16: bad arguments");
  B200FT_CHECK_ARG(((uintptr_t)in | (uintptr_t)out) % 16 == 0, "cast_f32_to_bf16: 16-byte alignment required");
  if (n == 0) return 0;
  int64_t want = ((n >> 3) + 255) / 256;
  int blocks = (int)(want < 1 ? 1 : (want > B200FT_NUM_SMS * 8 ? B200FT_NUM_SMS * 8 : want));
  launch_k(cast_f32_bf16_kernel, blocks, 256, 0, (cudaStream_t)stream, in, (__nv_bfloat16*)out, n);
  B200FT_LAUNCH_CHECK();
  return 0;
}
