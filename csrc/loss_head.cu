// Loss and narrow-output ("head") kernels.
//
//  * mse_loss_grad     - fused MSE loss + gradient: one pass over the predictions writes the scalar
//                        loss and dY = 2 (Yhat - Y) / numel  (aten::mse_loss + MseLossBackward,
//                        ref.py:173,176).  Deterministic two-stage reduction through scratch memory.
//  * head_fwd_mse      - last Linear layer when it has very few outputs (the regression head,
//                        e.g. 4096 -> 1): a GEMV per row on CUDA cores (tensor cores cannot be fed by
//                        N = 1), fused with bias, the MSE loss and dY (ref.py:44,170,173,176).
//  * head_bwd          - backward of that layer in ONE pass over its input H: dH = (dY W) * [H>0]
//                        (bf16) and per-row-strip partial sums of dW = dY^T H, db = sum dY and - for the
//                        bias gradient of the layer below - the column sums of dH.
//  * colsum / reduce_partials - bias gradient of the GEMM layers (db = column sums of dY) and the
//                        fixed-order reduction of any [strips][n] partial buffer.
#include "common.cuh"

namespace b200dp {
namespace {

constexpr int kRedThreads = 256;

// 8 consecutive elements of a bf16 (bf16 training) or fp32 (fp32 / TF32 training) row as floats
template <typename T> __device__ __forceinline__ void load8(const T* p, float (&o)[8]);
template <> __device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float (&o)[8]) {
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  o[0] = bf16_lo(v.x); o[1] = bf16_hi(v.x); o[2] = bf16_lo(v.y); o[3] = bf16_hi(v.y);
  o[4] = bf16_lo(v.z); o[5] = bf16_hi(v.z); o[6] = bf16_lo(v.w); o[7] = bf16_hi(v.w);
}
template <> __device__ __forceinline__ void load8<float>(const float* p, float (&o)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}
template <typename T> __device__ __forceinline__ void store8(T* p, const float (&v)[8]);
template <> __device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* p, const float (&v)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                            pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
}
template <> __device__ __forceinline__ void store8<float>(float* p, const float (&v)[8]) {
  reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}

// block-wide sum; result valid in thread 0
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = 0.0f;
  if (threadIdx.x == 0)
    for (int w = 0; w < (blockDim.x >> 5); ++w) s += red[w];
  return s;
}

// Publishes this block's partial, and lets the last block to arrive add all partials in block
// order (bit-reproducible) and write scale * sum to `out`.  scratch = [ticket(u32) pad to 256 B][partials].
__device__ __forceinline__ void finish_scalar(float block_val, float scale, float* out, float* partials,
                                              unsigned int* ticket, bool* is_last_smem) {
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = block_val;
    __threadfence();
    const unsigned int prev = atomicAdd(ticket, 1u);
    *is_last_smem = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (*is_last_smem && threadIdx.x == 0) {
    __threadfence();
    float s = 0.0f;
    for (unsigned int b = 0; b < gridDim.x; ++b) s += partials[b];
    *out = s * scale;
    *ticket = 0;      // the scratch is reusable by the next launch on the stream
  }
}

template <bool PRED_BF16, bool DY_BF16>
__global__ void __launch_bounds__(kRedThreads)
mse_kernel(const void* __restrict__ pred, const float* __restrict__ target, long long n, float inv_n,
           float gscale, void* __restrict__ dy, float* __restrict__ loss, float* partials, unsigned int* ticket) {
  __shared__ float red[kRedThreads / 32];
  __shared__ bool is_last;
  float acc = 0.0f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float p = PRED_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(pred)[i])
                              : reinterpret_cast<const float*>(pred)[i];
    const float d = p - target[i];
    acc = fmaf(d, d, acc);
    const float g = 2.0f * d * gscale;
    if (DY_BF16) reinterpret_cast<__nv_bfloat16*>(dy)[i] = __float2bfloat16_rn(g);
    else reinterpret_cast<float*>(dy)[i] = g;
  }
  const float b = block_sum(acc, red);
  finish_scalar(b, inv_n, loss, partials, ticket, &is_last);
}

// ---- head forward: one warp per row ----
template <typename T, int NOUT>
__global__ void __launch_bounds__(kRedThreads)
head_fwd_kernel(const T* __restrict__ H, long long ldh, const T* __restrict__ W,
                const float* __restrict__ bias, const float* __restrict__ target, int rows, int K,
                float inv_n, float gscale, float* __restrict__ yhat, float* __restrict__ dy,
                float* __restrict__ loss, float* partials, unsigned int* ticket) {
  __shared__ float red[kRedThreads / 32];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int gw = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
  const int nw = gridDim.x * warps_per_block;
  float acc_loss = 0.0f;
  for (int r = gw; r < rows; r += nw) {
    const T* h = H + (long long)r * ldh;
    float s[NOUT];
#pragma unroll
    for (int o = 0; o < NOUT; ++o) s[o] = 0.0f;
#pragma unroll 4
    for (int k = lane * 8; k < K; k += 256) {
      float hv[8];
      load8<T>(h + k, hv);
#pragma unroll
      for (int o = 0; o < NOUT; ++o) {
        float wv[8];
        load8<T>(W + (long long)o * K + k, wv);
#pragma unroll
        for (int q = 0; q < 8; ++q) s[o] = fmaf(hv[q], wv[q], s[o]);
      }
    }
#pragma unroll
    for (int o = 0; o < NOUT; ++o) s[o] = warp_sum(s[o]);
    if (lane == 0) {
#pragma unroll
      for (int o = 0; o < NOUT; ++o) {
        const float p = s[o] + bias[o];
        const float d = p - target[(long long)r * NOUT + o];
        yhat[(long long)r * NOUT + o] = p;
        dy[(long long)r * NOUT + o] = 2.0f * d * gscale;
        acc_loss = fmaf(d, d, acc_loss);
      }
    }
  }
  const float b = block_sum(acc_loss, red);
  finish_scalar(b, inv_n, loss, partials, ticket, &is_last);
}

// ---- head backward: block = (row strip, 2048-column chunk); thread owns 8 columns ----
constexpr int kHeadBwdThreads = 256;
constexpr int kHeadBwdCols = kHeadBwdThreads * 8;

template <typename T, int NOUT>
__global__ void __launch_bounds__(kHeadBwdThreads)
head_bwd_kernel(const float* __restrict__ dy, const T* __restrict__ W,
                const T* __restrict__ H, long long ldh, int rows, int K, int rows_per_strip,
                T* __restrict__ dH, long long lddh, float* __restrict__ partials, int prow) {
  // partials layout: [strip][prow][K + 1], prow = NOUT (+1: column sums of dH); column K of the first
  // NOUT rows holds the bias-gradient partial
  const int strip = blockIdx.x;
  const int k0 = blockIdx.y * kHeadBwdCols + threadIdx.x * 8;
  const int r_begin = strip * rows_per_strip;
  const int r_end = min(rows, r_begin + rows_per_strip);
  const bool active = k0 < K;
  float w[NOUT][8];
  float acc[NOUT][8];
#pragma unroll
  for (int o = 0; o < NOUT; ++o) {
#pragma unroll
    for (int q = 0; q < 8; ++q) w[o][q] = acc[o][q] = 0.0f;
    if (active) load8<T>(W + (long long)o * K + k0, w[o]);
  }
  float bacc[NOUT];
#pragma unroll
  for (int o = 0; o < NOUT; ++o) bacc[o] = 0.0f;
  float dhsum[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) dhsum[j] = 0.0f;

#pragma unroll 4
  for (int r = r_begin; r < r_end; ++r) {
    float g[NOUT];
#pragma unroll
    for (int o = 0; o < NOUT; ++o) g[o] = __ldg(dy + (long long)r * NOUT + o);
    if (blockIdx.y == 0 && threadIdx.x == 0) {
#pragma unroll
      for (int o = 0; o < NOUT; ++o) bacc[o] += g[o];
    }
    if (!active) continue;
    float h[8];
    load8<T>(H + (long long)r * ldh + k0, h);
    float dh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) dh[j] = 0.0f;
#pragma unroll
    for (int o = 0; o < NOUT; ++o) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        dh[j] = fmaf(g[o], w[o][j], dh[j]);
        acc[o][j] = fmaf(g[o], h[j], acc[o][j]);
      }
    }
    if (dH != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        dh[j] = h[j] > 0.0f ? dh[j] : 0.0f;
        dhsum[j] += dh[j];
      }
      store8<T>(dH + (long long)r * lddh + k0, dh);
    }
  }
  float* my = partials + (long long)strip * prow * (K + 1);
  if (active) {
#pragma unroll
    for (int o = 0; o < NOUT; ++o) {
      // (K+1)-float rows are not 16-byte aligned in general: scalar stores
      float* d = my + (long long)o * (K + 1) + k0;
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = acc[o][j];
    }
    if (prow > NOUT) {
      float* d = my + (long long)NOUT * (K + 1) + k0;
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = dhsum[j];
    }
  }
  if (blockIdx.y == 0 && threadIdx.x == 0) {
#pragma unroll
    for (int o = 0; o < NOUT; ++o) my[(long long)o * (K + 1) + K] = bacc[o];
  }
}

// out[j] = sum_s partials[s][j], fp32 or bf16 output.  A block owns 64 columns; its 4 thread groups
// each add every 4th strip (loads independent, unrolled), then the 4 group sums are added in group
// order - a fixed summation tree, so the result is bit-reproducible.
constexpr int kRedCols = 64;
constexpr int kRedGroups = 4;

__device__ __forceinline__ float strided_strip_sum(const float* __restrict__ partials, int strips, long long n,
                                                   long long j, float (*sm)[kRedCols]) {
  const int c = threadIdx.x % kRedCols, g = threadIdx.x / kRedCols;
  float s = 0.0f;
  if (j < n) {
#pragma unroll 8
    for (int t = g; t < strips; t += kRedGroups) s += partials[(long long)t * n + j];
  }
  sm[g][c] = s;
  __syncthreads();
  float tot = 0.0f;
  if (g == 0) {
#pragma unroll
    for (int q = 0; q < kRedGroups; ++q) tot += sm[q][c];
  }
  return tot;
}

template <bool OUT_BF16>
__global__ void __launch_bounds__(kRedCols * kRedGroups)
reduce_partials_kernel(const float* __restrict__ partials, int strips, long long n, void* out) {
  __shared__ float sm[kRedGroups][kRedCols];
  const long long j = (long long)blockIdx.x * kRedCols + threadIdx.x % kRedCols;
  const float s = strided_strip_sum(partials, strips, n, j, sm);
  if (threadIdx.x >= kRedCols || j >= n) return;
  if (OUT_BF16) reinterpret_cast<__nv_bfloat16*>(out)[j] = __float2bfloat16_rn(s);
  else reinterpret_cast<float*>(out)[j] = s;
}

// same reduction; scatters the "[prow][K+1]" head partials into the weight / bias gradient slots (row
// `nout`, when present, holds the column sums of dH = the bias gradient of the layer below)
template <bool OUT_BF16>
__global__ void __launch_bounds__(kRedCols * kRedGroups)
head_reduce_kernel(const float* __restrict__ partials, int strips, int nout, int prow, int K, void* dW, void* db,
                   void* db_prev) {
  __shared__ float sm[kRedGroups][kRedCols];
  const long long n = (long long)prow * (K + 1);
  const long long j = (long long)blockIdx.x * kRedCols + threadIdx.x % kRedCols;
  const float s = strided_strip_sum(partials, strips, n, j, sm);
  if (threadIdx.x >= kRedCols || j >= n) return;
  const int o = int(j / (K + 1)), k = int(j % (K + 1));
  if (o >= nout && k >= K) return;
  void* dst = o >= nout ? db_prev : (k < K ? dW : db);
  const long long idx = o >= nout ? k : (k < K ? (long long)o * K + k : o);
  if (OUT_BF16) reinterpret_cast<__nv_bfloat16*>(dst)[idx] = __float2bfloat16_rn(s);
  else reinterpret_cast<float*>(dst)[idx] = s;
}

// column sums of a bf16 / fp32 [rows, n] matrix over a strip of rows -> partials[strip][n]
template <typename T>
__global__ void __launch_bounds__(256)
colsum_kernel(const T* __restrict__ A, long long lda, int rows, int n, int rows_per_strip,
              float* __restrict__ partials) {
  const int c0 = (blockIdx.y * blockDim.x + threadIdx.x) * 8;
  if (c0 >= n) return;
  const int r_begin = blockIdx.x * rows_per_strip;
  const int r_end = min(rows, r_begin + rows_per_strip);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.0f;
  for (int r = r_begin; r < r_end; ++r) {
    float v[8];
    load8<T>(A + (long long)r * lda + c0, v);
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[q] += v[q];
  }
  float* dst = partials + (long long)blockIdx.x * n + c0;
  *reinterpret_cast<float4*>(dst) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
}

inline int reduce_grid(long long n) {
  long long g = (n + kRedThreads - 1) / kRedThreads;
  const long long cap = 4ll * sm_count();
  return int(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace
}  // namespace b200dp

using namespace b200dp;

extern "C" long long b200dp_mse_scratch_bytes(long long numel) { return 256 + 4ll * reduce_grid(numel); }

extern "C" int b200dp_mse_loss_grad(const void* pred, int pred_bf16, const float* target, long long numel,
                                    long long denom, void* dy, int dy_bf16, float* loss, float grad_scale,
                                    void* scratch, long long scratch_bytes, void* stream) {
  B200DP_REQUIRE(pred && target && dy && loss && scratch, "mse: null pointer");
  B200DP_REQUIRE(numel >= 1, "mse: numel must be >= 1");
  const int grid = reduce_grid(numel);
  B200DP_REQUIRE(scratch_bytes >= 256 + 4ll * grid, "mse: scratch too small (%lld < %lld bytes)", scratch_bytes,
                 256 + 4ll * grid);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  float* partials = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  const float inv_n = 1.0f / float(denom > 0 ? denom : numel);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const float gscale = inv_n * grad_scale;
#define LAUNCH(PB, DB) mse_kernel<PB, DB><<<grid, kRedThreads, 0, s>>>(pred, target, numel, inv_n, gscale, dy, loss, partials, ticket)
  if (pred_bf16 && dy_bf16) LAUNCH(true, true);
  else if (pred_bf16) LAUNCH(true, false);
  else if (dy_bf16) LAUNCH(false, true);
  else LAUNCH(false, false);
#undef LAUNCH
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

extern "C" long long b200dp_head_scratch_bytes(int rows, int K, int nout) {
  const int strips = (rows + 31) / 32;
  const long long fwd = 256 + 4ll * 4 * sm_count();
  return fwd + 4ll * strips * (nout + 1) * (K + 1);
}

extern "C" int b200dp_head_fwd_mse(const void* H, long long ldh, const void* W, int elems_fp32, const float* bias,
                                   const float* target, int rows, int K, int nout, float* yhat, float* dy,
                                   float* loss, float grad_scale, void* scratch, long long scratch_bytes,
                                   void* stream) {
  B200DP_REQUIRE(H && W && bias && target && yhat && dy && loss && scratch, "head_fwd: null pointer");
  B200DP_REQUIRE(nout >= 1 && nout <= 4, "head_fwd: supports 1..4 outputs (got %d)", nout);
  B200DP_REQUIRE(K % 8 == 0 && ldh % 8 == 0, "head_fwd: K and ldh must be multiples of 8");
  B200DP_REQUIRE(rows >= 1, "head_fwd: rows must be >= 1");
  int grid = (rows + (kRedThreads / 32) - 1) / (kRedThreads / 32);
  if (grid > 4 * sm_count()) grid = 4 * sm_count();
  B200DP_REQUIRE(scratch_bytes >= 256 + 4ll * grid, "head_fwd: scratch too small");
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  float* partials = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  const float inv_n = 1.0f / (float(rows) * float(nout));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const float gscale = inv_n * grad_scale;
#define LAUNCH(TT, NO) head_fwd_kernel<TT, NO><<<grid, kRedThreads, 0, s>>>(reinterpret_cast<const TT*>(H), ldh, reinterpret_cast<const TT*>(W), bias, target, rows, K, inv_n, gscale, yhat, dy, loss, partials, ticket)
#define LAUNCH_T(NO) do { if (elems_fp32) LAUNCH(float, NO); else LAUNCH(__nv_bfloat16, NO); } while (0)
  switch (nout) {
    case 1: LAUNCH_T(1); break;
    case 2: LAUNCH_T(2); break;
    case 3: LAUNCH_T(3); break;
    default: LAUNCH_T(4); break;
  }
#undef LAUNCH_T
#undef LAUNCH
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

extern "C" int b200dp_head_bwd(const float* dy, const void* W, const void* H, long long ldh, int elems_fp32,
                               int rows, int K, int nout, void* dH, long long lddh, void* dW, void* db, void* db_prev,
                               int grads_bf16, void* scratch, long long scratch_bytes, void* stream) {
  B200DP_REQUIRE(dy && W && H && dW && db && scratch, "head_bwd: null pointer");
  B200DP_REQUIRE(db_prev == nullptr || dH != nullptr, "head_bwd: db_prev (column sums of dH) needs dH");
  B200DP_REQUIRE(nout >= 1 && nout <= 4, "head_bwd: supports 1..4 outputs (got %d)", nout);
  B200DP_REQUIRE(K % 8 == 0 && ldh % 8 == 0 && (dH == nullptr || lddh % 8 == 0),
                 "head_bwd: K, ldh, lddh must be multiples of 8");
  const int rows_per_strip = 32;
  const int strips = (rows + rows_per_strip - 1) / rows_per_strip;
  const long long fwd = 256 + 4ll * 4 * sm_count();
  const int prow = nout + (db_prev != nullptr ? 1 : 0);
  B200DP_REQUIRE(scratch_bytes >= fwd + 4ll * strips * prow * (K + 1), "head_bwd: scratch too small");
  float* partials = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + fwd);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  dim3 grid(strips, (K + kHeadBwdCols - 1) / kHeadBwdCols);
#define LAUNCH(TT, NO) head_bwd_kernel<TT, NO><<<grid, kHeadBwdThreads, 0, s>>>(dy, reinterpret_cast<const TT*>(W), reinterpret_cast<const TT*>(H), ldh, rows, K, rows_per_strip, reinterpret_cast<TT*>(dH), lddh, partials, prow)
#define LAUNCH_T(NO) do { if (elems_fp32) LAUNCH(float, NO); else LAUNCH(__nv_bfloat16, NO); } while (0)
  switch (nout) {
    case 1: LAUNCH_T(1); break;
    case 2: LAUNCH_T(2); break;
    case 3: LAUNCH_T(3); break;
    default: LAUNCH_T(4); break;
  }
#undef LAUNCH_T
#undef LAUNCH
  B200DP_CUDA_OK(cudaGetLastError());
  const long long n = (long long)prow * (K + 1);
  const int rgrid = int((n + kRedCols - 1) / kRedCols);
  if (grads_bf16)
    head_reduce_kernel<true><<<rgrid, kRedCols * kRedGroups, 0, s>>>(partials, strips, nout, prow, K, dW, db, db_prev);
  else
    head_reduce_kernel<false><<<rgrid, kRedCols * kRedGroups, 0, s>>>(partials, strips, nout, prow, K, dW, db, db_prev);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

extern "C" long long b200dp_colsum_scratch_bytes(int rows, int n) {
  const int strips = (rows + 63) / 64;
  return 4ll * strips * n;
}

extern "C" int b200dp_colsum(const void* A, int a_fp32, long long lda, int rows, int n, void* out, int out_bf16,
                             void* scratch, long long scratch_bytes, void* stream) {
  B200DP_REQUIRE(A && out && scratch, "colsum: null pointer");
  B200DP_REQUIRE(n % 8 == 0 && lda % 8 == 0, "colsum: n and lda must be multiples of 8");
  const int rows_per_strip = 64;
  const int strips = (rows + rows_per_strip - 1) / rows_per_strip;
  B200DP_REQUIRE(scratch_bytes >= 4ll * strips * n, "colsum: scratch too small");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  float* partials = reinterpret_cast<float*>(scratch);
  dim3 grid(strips, (n / 8 + 255) / 256);
  if (a_fp32) colsum_kernel<float><<<grid, 256, 0, s>>>(reinterpret_cast<const float*>(A), lda, rows, n, rows_per_strip, partials);
  else colsum_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(A), lda, rows, n, rows_per_strip, partials);
  B200DP_CUDA_OK(cudaGetLastError());
  const int rgrid = (n + kRedCols - 1) / kRedCols;
  if (out_bf16) reduce_partials_kernel<true><<<rgrid, kRedCols * kRedGroups, 0, s>>>(partials, strips, n, out);
  else reduce_partials_kernel<false><<<rgrid, kRedCols * kRedGroups, 0, s>>>(partials, strips, n, out);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}
