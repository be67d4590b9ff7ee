// Local (single-GPU) optimizer and cast kernels.
//
//  * sgd_momentum   - torch.optim.SGD(momentum) semantics of ref.py:91,211 in one pass over flat
//                     buffers: g' = g * scale;  buf = first ? g' : mu*buf + g';  p -= lr*buf;
//                     optionally refreshes the bf16 shadow weights the GEMMs read.  Used when the
//                     world has one rank, and after the NCCL all-reduce baseline.  The multi-GPU
//                     product path is the fused kernel in allreduce_sgd.cu.
//  * cast_f32_bf16  - one-time input / weight conversion (the reference's per-step .float() cast,
//                     ref.py:159, becomes a load-time cast).
//  * transpose_f32  - dst[c, r] = src[r, c].  The fp32 (kind::tf32) GEMM path needs it: tcgen05 reads
//                     TF32 operands K-major only (measured on B200: an MN-major kind::tf32 MMA leaves
//                     the accumulator untouched), so backward's W^T, dY^T and X^T are materialised -
//                     unlike the bf16 path, which expresses them in the UMMA descriptors.
#include "common.cuh"

namespace b200dp {
namespace {

template <bool GRAD_BF16>
__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ p, const void* __restrict__ g, float* __restrict__ m,
           __nv_bfloat16* __restrict__ shadow, long long n4, float lr, float mu, float scale, int first) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 gv;
    if (GRAD_BF16) {
      const uint2 u = __ldcs(reinterpret_cast<const uint2*>(g) + i);
      gv = make_float4(bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y));
    } else {
      gv = __ldcs(reinterpret_cast<const float4*>(g) + i);
    }
    gv.x *= scale; gv.y *= scale; gv.z *= scale; gv.w *= scale;
    float4 mv;
    if (first) mv = gv;
    else {
      mv = __ldcs(reinterpret_cast<const float4*>(m) + i);
      mv.x = fmaf(mu, mv.x, gv.x); mv.y = fmaf(mu, mv.y, gv.y);
      mv.z = fmaf(mu, mv.z, gv.z); mv.w = fmaf(mu, mv.w, gv.w);
    }
    // streaming (evict-first) accesses: this kernel moves 20 B/parameter through L2 while GEMMs
    // run on the other stream; it must not push their operand panels out of L2
    __stcs(reinterpret_cast<float4*>(m) + i, mv);
    float4 pv = __ldcs(reinterpret_cast<const float4*>(p) + i);
    pv.x -= lr * mv.x; pv.y -= lr * mv.y; pv.z -= lr * mv.z; pv.w -= lr * mv.w;
    __stcs(reinterpret_cast<float4*>(p) + i, pv);
    if (shadow) __stcs(reinterpret_cast<uint2*>(shadow) + i, make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w)));
  }
}

__global__ void __launch_bounds__(256)
cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    reinterpret_cast<uint2*>(dst)[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __float2bfloat16_rn(src[i]);
}

// dst[r, :] = src[idx[r], :] for rows of `row_bytes` bytes (multiple of 16): one warp per row.
__global__ void __launch_bounds__(256)
gather_rows_kernel(const uint4* __restrict__ src, long long src_ld16, const long long* __restrict__ idx, int rows,
                   int row16, uint4* __restrict__ dst) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int r = warp; r < rows; r += nwarps) {
    const uint4* s = src + idx[r] * src_ld16;
    uint4* d = dst + (long long)r * row16;
    for (int c = lane; c < row16; c += 32) d[c] = __ldg(s + c);
  }
}

// 32x32 tiles through padded shared memory: coalesced 128-byte reads along src rows and writes along dst rows
__global__ void __launch_bounds__(256)
transpose_f32_kernel(const float* __restrict__ src, long long lds, int rows, int cols, float* __restrict__ dst,
                     long long ldd) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8 threads
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int r = r0 + ty + j, c = c0 + tx;
    if (r < rows && c < cols) tile[ty + j][tx] = __ldg(src + (long long)r * lds + c);
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int c = c0 + ty + j, r = r0 + tx;
    if (r < rows && c < cols) dst[(long long)c * ldd + r] = tile[tx][ty + j];
  }
}

inline int ew_grid(long long n) {
  long long g = (n + 255) / 256;
  const long long cap = 8ll * sm_count();
  return int(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace
}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_sgd_momentum_ex(float* params, const void* grads, int grads_bf16, float* momentum,
                                      void* shadow_bf16, long long n, float lr, float mu, float grad_scale,
                                      int first_step, int max_blocks, void* stream);

extern "C" int b200dp_sgd_momentum(float* params, const void* grads, int grads_bf16, float* momentum,
                                   void* shadow_bf16, long long n, float lr, float mu, float grad_scale,
                                   int first_step, void* stream) {
  return b200dp_sgd_momentum_ex(params, grads, grads_bf16, momentum, shadow_bf16, n, lr, mu, grad_scale, first_step,
                                0, stream);
}

// max_blocks > 0 caps the grid: an update that runs beside tensor-core GEMMs is deliberately held
// below the full HBM rate so the GEMMs' TMA loads keep their latency (it only has to finish before
// the layer's backward does).
extern "C" int b200dp_sgd_momentum_ex(float* params, const void* grads, int grads_bf16, float* momentum,
                                      void* shadow_bf16, long long n, float lr, float mu, float grad_scale,
                                      int first_step, int max_blocks, void* stream) {
  B200DP_REQUIRE(params && grads && momentum, "sgd: null pointer");
  B200DP_REQUIRE(n % 4 == 0, "sgd: n must be a multiple of 4 (flat buffers are padded)");
  B200DP_REQUIRE(((uintptr_t)params & 15) == 0 && ((uintptr_t)momentum & 15) == 0 &&
                     ((uintptr_t)grads & (grads_bf16 ? 7 : 15)) == 0,
                 "sgd: buffers must be 16-byte aligned");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long n4 = n / 4;
  int grid = ew_grid(n4);
  if (max_blocks > 0 && grid > max_blocks) grid = max_blocks;
  if (grads_bf16)
    sgd_kernel<true><<<grid, 256, 0, s>>>(params, grads, momentum, reinterpret_cast<__nv_bfloat16*>(shadow_bf16),
                                          n4, lr, mu, grad_scale, first_step);
  else
    sgd_kernel<false><<<grid, 256, 0, s>>>(params, grads, momentum, reinterpret_cast<__nv_bfloat16*>(shadow_bf16),
                                           n4, lr, mu, grad_scale, first_step);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

extern "C" int b200dp_cast_f32_bf16(const float* src, void* dst, long long n, void* stream) {
  B200DP_REQUIRE(src && dst, "cast: null pointer");
  B200DP_REQUIRE(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 7) == 0, "cast: misaligned buffers");
  cast_kernel<<<ew_grid(n / 4 + 1), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

// Mini-batch assembly on the device: dst[r,:] = src[idx[r],:].  Replaces the
// DataLoader's index + default_collate on the host (ref.py:146,155) when batch < shard.
extern "C" int b200dp_gather_rows(const void* src, long long src_row_bytes, const long long* idx, int rows,
                                  long long row_bytes, void* dst, void* stream) {
  B200DP_REQUIRE(src && idx && dst, "gather_rows: null pointer");
  B200DP_REQUIRE(row_bytes > 0 && row_bytes % 16 == 0 && src_row_bytes % 16 == 0,
                 "gather_rows: row sizes must be multiples of 16 bytes");
  B200DP_REQUIRE(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0, "gather_rows: misaligned buffers");
  if (rows <= 0) return B200DP_OK;
  int grid = (rows + 7) / 8;
  if (grid > 8 * sm_count()) grid = 8 * sm_count();
  gather_rows_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const uint4*>(src), src_row_bytes / 16, idx, rows, int(row_bytes / 16),
      reinterpret_cast<uint4*>(dst));
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

extern "C" int b200dp_transpose_f32(const float* src, long long lds, int rows, int cols, float* dst, long long ldd,
                                    void* stream) {
  B200DP_REQUIRE(src && dst, "transpose: null pointer");
  B200DP_REQUIRE(rows > 0 && cols > 0 && lds >= cols && ldd >= rows, "transpose: bad shape (%d x %d, lds %lld, ldd %lld)",
                 rows, cols, lds, ldd);
  dim3 grid((cols + 31) / 32, (rows + 31) / 32);
  B200DP_REQUIRE(grid.y <= 65535, "transpose: too many rows (%d)", rows);
  transpose_f32_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(src, lds, rows, cols, dst, ldd);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}
