// Whole-network SIMT kernel for MLPs too small for tensor-core tiles (the reference's 2-3-1 net,
// ref.py:41-45): forward, MSE loss, backward and gradient reduction over the batch in ONE launch,
// fp32 throughout.  Replaces the ATen op sequence of ref.py:170-176 (addmm, relu, addmm, mse_loss,
// MseLossBackward, AddmmBackward x2, threshold_backward, AccumulateGrad x4).
//
// Each block walks row tiles; per tile it runs the layers forward with (row, output) parallelism,
// back-propagates deltas the same way, then switches to parameter parallelism: every parameter is
// owned by one thread which sums its per-row contributions in row order - no atomics, so the
// result is bit-reproducible.  Blocks publish partial gradients to scratch memory; the last block
// to finish (ticket counter) adds them in block order and writes the flat gradient buffer + loss.
#include <mutex>

#include "common.cuh"

namespace b200dp {
namespace {

constexpr int kThreads = 256;

struct TinyDesc {
  int num_layers;
  int sizes[B200DP_MAX_LAYERS + 1];
  int act_off[B200DP_MAX_LAYERS + 1];   // offset of layer-l activations inside one row record
  long long w_off[B200DP_MAX_LAYERS];
  long long b_off[B200DP_MAX_LAYERS];
  int row_stride;                         // floats per row record (sum of sizes, padded odd)
  int tile_rows;
  long long flat_numel;
};

__global__ void __launch_bounds__(kThreads)
tiny_mlp_kernel(const TinyDesc d, const float* __restrict__ params, const float* __restrict__ x,
                const float* __restrict__ y, int rows, float inv_count, float* __restrict__ grads,
                float* __restrict__ loss_out, float* __restrict__ part, unsigned int* __restrict__ ticket) {
  extern __shared__ float sm[];
  const int L = d.num_layers;
  const int S = d.row_stride;
  float* act = sm;                                   // [tile_rows][S]   activations a_0..a_L
  float* dlt = act + d.tile_rows * S;                // [tile_rows][S]   deltas dz_1..dz_L (slot l)
  float* gacc = dlt + d.tile_rows * S;               // [flat_numel]     this block's gradient sums
  __shared__ float red[kThreads / 32];
  __shared__ bool is_last;

  const int tid = threadIdx.x;
  const long long P = d.flat_numel;
  for (long long p = tid; p < P; p += kThreads) gacc[p] = 0.0f;
  float loss_local = 0.0f;
  const int nout = d.sizes[L];

  for (int r0 = blockIdx.x * d.tile_rows; r0 < rows; r0 += gridDim.x * d.tile_rows) {
    const int nr = min(d.tile_rows, rows - r0);
    __syncthreads();
    // ---- load inputs ----
    for (int i = tid; i < nr * d.sizes[0]; i += kThreads) {
      const int r = i / d.sizes[0], j = i % d.sizes[0];
      act[r * S + j] = x[(long long)(r0 + r) * d.sizes[0] + j];
    }
    __syncthreads();
    // ---- forward ----
    for (int l = 0; l < L; ++l) {
      const int nin = d.sizes[l], no = d.sizes[l + 1];
      const float* W = params + d.w_off[l];
      const float* b = params + d.b_off[l];
      for (int i = tid; i < nr * no; i += kThreads) {
        const int r = i / no, o = i % no;
        const float* a = act + r * S + d.act_off[l];
        float z = b[o];
        for (int k = 0; k < nin; ++k) z = fmaf(W[o * nin + k], a[k], z);
        if (l < L - 1) z = fmaxf(z, 0.0f);
        act[r * S + d.act_off[l + 1] + o] = z;
      }
      __syncthreads();
    }
    // ---- loss + output delta: dz_L = 2 (yhat - y) / (rows * nout) ----
    for (int i = tid; i < nr * nout; i += kThreads) {
      const int r = i / nout, o = i % nout;
      const float diff = act[r * S + d.act_off[L] + o] - y[(long long)(r0 + r) * nout + o];
      loss_local += diff * diff;
      dlt[r * S + d.act_off[L] + o] = 2.0f * diff * inv_count;
    }
    __syncthreads();
    // ---- backward deltas ----
    for (int l = L - 1; l >= 1; --l) {
      const int nin = d.sizes[l], no = d.sizes[l + 1];
      const float* W = params + d.w_off[l];
      for (int i = tid; i < nr * nin; i += kThreads) {
        const int r = i / nin, k = i % nin;
        const float* dz = dlt + r * S + d.act_off[l + 1];
        float da = 0.0f;
        for (int o = 0; o < no; ++o) da = fmaf(dz[o], W[o * nin + k], da);
        dlt[r * S + d.act_off[l] + k] = act[r * S + d.act_off[l] + k] > 0.0f ? da : 0.0f;
      }
      __syncthreads();
    }
    // ---- parameter-parallel gradient accumulation ----
    for (int l = 0; l < L; ++l) {
      const int nin = d.sizes[l], no = d.sizes[l + 1];
      for (int i = tid; i < no * (nin + 1); i += kThreads) {
        const int o = i / (nin + 1), k = i % (nin + 1);
        float g = 0.0f;
        if (k < nin) {
          for (int r = 0; r < nr; ++r) g = fmaf(dlt[r * S + d.act_off[l + 1] + o], act[r * S + d.act_off[l] + k], g);
          gacc[d.w_off[l] + o * nin + k] += g;
        } else {
          for (int r = 0; r < nr; ++r) g += dlt[r * S + d.act_off[l + 1] + o];
          gacc[d.b_off[l] + o] += g;
        }
      }
    }
  }
  __syncthreads();

  // ---- block loss ----
  float ws = warp_sum(loss_local);
  if ((tid & 31) == 0) red[tid >> 5] = ws;
  __syncthreads();
  float block_loss = 0.0f;
  if (tid == 0) {
    for (int w = 0; w < kThreads / 32; ++w) block_loss += red[w];
  }

  if (gridDim.x == 1) {
    for (long long p = tid; p < P; p += kThreads) grads[p] = gacc[p];
    if (tid == 0) *loss_out = block_loss * inv_count;
    return;
  }
  // ---- publish partials; last block reduces in block order ----
  float* my = part + (long long)blockIdx.x * (P + 1);
  for (long long p = tid; p < P; p += kThreads) my[p] = gacc[p];
  if (tid == 0) my[P] = block_loss;
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned int prev = atomicAdd(ticket, 1u);
    is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  for (long long p = tid; p <= P; p += kThreads) {
    float s = 0.0f;
    for (int b = 0; b < gridDim.x; ++b) s += part[(long long)b * (P + 1) + p];
    if (p < P) grads[p] = s;
    else *loss_out = s * inv_count;
  }
  if (tid == 0) *ticket = 0;    // ready for the next launch
}

int fill_desc(const b200dp_mlp_desc* m, TinyDesc* d, int* smem_bytes) {
  B200DP_REQUIRE(m->num_layers >= 1 && m->num_layers <= B200DP_MAX_LAYERS, "tiny_mlp: bad layer count %d",
                 m->num_layers);
  d->num_layers = m->num_layers;
  int off = 0;
  for (int l = 0; l <= m->num_layers; ++l) {
    B200DP_REQUIRE(m->sizes[l] >= 1, "tiny_mlp: bad layer size");
    d->sizes[l] = m->sizes[l];
    d->act_off[l] = off;
    off += m->sizes[l];
  }
  for (int l = 0; l < m->num_layers; ++l) {
    d->w_off[l] = m->w_off[l];
    d->b_off[l] = m->b_off[l];
  }
  d->row_stride = off | 1;
  d->flat_numel = m->flat_numel;
  const long long budget = 200 * 1024;
  int tile = 128;
  while (tile > 1 && (2ll * tile * d->row_stride + d->flat_numel) * 4 > budget) tile >>= 1;
  B200DP_REQUIRE((2ll * tile * d->row_stride + d->flat_numel) * 4 <= budget,
                 "tiny_mlp: network too large for the fused SIMT kernel (%lld parameters)", d->flat_numel);
  d->tile_rows = tile;
  *smem_bytes = int((2ll * tile * d->row_stride + d->flat_numel) * 4);
  return B200DP_OK;
}

int grid_for(const TinyDesc& d, int rows) {
  int tiles = (rows + d.tile_rows - 1) / d.tile_rows;
  int g = tiles < sm_count() ? tiles : sm_count();
  return g < 1 ? 1 : g;
}

}  // namespace
}  // namespace b200dp

extern "C" long long b200dp_tiny_mlp_scratch_bytes(const b200dp_mlp_desc* m, int rows) {
  using namespace b200dp;
  TinyDesc d;
  int smem;
  if (fill_desc(m, &d, &smem)) return -1;
  const int g = grid_for(d, rows);
  return 256 + (long long)g * (d.flat_numel + 1) * 4;
}

extern "C" int b200dp_tiny_mlp_fwd_bwd(const b200dp_mlp_desc* m, const float* params, const float* x,
                                       const float* y, int rows, float* grads, float* loss, void* scratch,
                                       long long scratch_bytes, void* stream) {
  using namespace b200dp;
  B200DP_REQUIRE(m && params && x && y && grads && loss, "tiny_mlp: null pointer");
  B200DP_REQUIRE(rows >= 1, "tiny_mlp: rows must be >= 1");
  TinyDesc d;
  int smem;
  if (int rc = fill_desc(m, &d, &smem)) return rc;
  const int g = grid_for(d, rows);
  const long long need = g > 1 ? 256 + (long long)g * (d.flat_numel + 1) * 4 : 0;
  B200DP_REQUIRE(g == 1 || (scratch != nullptr && scratch_bytes >= need),
                 "tiny_mlp: scratch too small (%lld < %lld bytes)", scratch_bytes, need);
  static int attr_smem_dev[64] = {0};          // per device: the attribute belongs to the device's context
  int dev_ = 0;
  cudaGetDevice(&dev_);
  int& attr_smem = attr_smem_dev[dev_ & 63];
  if (smem > attr_smem) {
    B200DP_CUDA_OK(cudaFuncSetAttribute(tiny_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_smem = smem;
  }
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  float* part = scratch ? reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256) : nullptr;
  const float inv_count = 1.0f / (float(rows) * float(d.sizes[d.num_layers]));
  tiny_mlp_kernel<<<g, kThreads, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
      d, params, x, y, rows, inv_count, grads, loss, part, ticket);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

// ================================================================================================
// Allocation-free variant of b200dp_tiny_mlp_fwd_bwd.
//
// Who it is for: callers that must not (or cannot) call cudaMalloc at the point where they train -
// e.g. a host thread inside a latency-critical loop where an allocation would serialise the whole
// device, a plugin running inside another framework whose memory pool owns every byte of the GPU,
// or a process that was handed a fixed memory budget at start-up.  Such callers still own the
// inputs and outputs; what they cannot provide is the kernel's cross-block scratch (one ticket
// counter + one partial gradient/loss record per thread block).
//
// That scratch lives here instead, in a file-scope device array of a fixed, modest size.  Every
// GPU that loads the library carries this array until the process exits.
//
// Rules that follow from sharing one array:
//   * a call whose scratch need exceeds B200DP_TINY_NOALLOC_SCRATCH_BYTES is refused with an error;
//   * the array keeps whatever the previous call left in it; the only part that must start at
//     zero is the 4-byte ticket counter, which this entry point clears on the stream before each
//     launch (a memset node, not an allocation);
//   * two calls that can overlap on the same GPU (different streams, different threads) race on
//     the array and produce garbage - serialise them, or use the scratch-taking entry point.
//
// In host code the array's name is only a host-side handle, not a device address, so the address
// the kernel receives is looked up with cudaGetSymbolAddress - once per device, with that device
// current, on the first call made on that device.  Because that look-up talks to the runtime,
// make the first call on each device BEFORE starting any stream capture.  Nothing here calls
// cudaMalloc / cudaFree or any allocating library routine.
// ================================================================================================
namespace b200dp {
namespace {

static __device__ __align__(256) unsigned char g_tiny_noalloc_scratch[B200DP_TINY_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
struct NoAllocDeviceState {
  bool ready;
  void* scratch;       // device address of g_tiny_noalloc_scratch on this device
};
NoAllocDeviceState g_noalloc[kMaxDevices];
std::mutex g_noalloc_mu;

// Per-device set-up; the current device must be the one the caller is about to launch on.
int noalloc_setup(int dev, void** scratch_out) {
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  NoAllocDeviceState& st = g_noalloc[dev];
  if (!st.ready) {
    void* ptr = nullptr;
    B200DP_CUDA_OK(cudaGetSymbolAddress(&ptr, g_tiny_noalloc_scratch));
    B200DP_REQUIRE(ptr != nullptr, "tiny_mlp_noalloc: cudaGetSymbolAddress returned a null address");
    st.scratch = ptr;
    st.ready = true;
  }
  *scratch_out = st.scratch;
  return B200DP_OK;
}

}  // namespace
}  // namespace b200dp

extern "C" int b200dp_tiny_mlp_fwd_bwd_noalloc(const b200dp_mlp_desc* m, const float* params, const float* x,
                                               const float* y, int rows, float* grads, float* loss,
                                               void* stream) {
  using namespace b200dp;
  B200DP_REQUIRE(m && params && x && y && grads && loss, "tiny_mlp_noalloc: null pointer");
  B200DP_REQUIRE(rows >= 1, "tiny_mlp_noalloc: rows must be >= 1");
  TinyDesc d;
  int smem;
  if (int rc = fill_desc(m, &d, &smem)) return rc;
  const int g = grid_for(d, rows);
  const long long need = g > 1 ? 256 + (long long)g * (d.flat_numel + 1) * 4 : 0;
  B200DP_REQUIRE(need <= (long long)B200DP_TINY_NOALLOC_SCRATCH_BYTES,
                 "tiny_mlp_noalloc: this problem needs %lld bytes of scratch (%d blocks x %lld floats) but the "
                 "built-in scratch holds %d bytes; use b200dp_tiny_mlp_fwd_bwd with caller-provided scratch",
                 need, g, d.flat_numel + 1, B200DP_TINY_NOALLOC_SCRATCH_BYTES);
  int dev = 0;
  B200DP_CUDA_OK(cudaGetDevice(&dev));
  B200DP_REQUIRE(dev >= 0 && dev < kMaxDevices, "tiny_mlp_noalloc: device ordinal %d out of range", dev);
  void* scratch = nullptr;
  if (int rc = noalloc_setup(dev, &scratch)) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  // Nets that need more than the default 48 KB of dynamic shared memory: opt in to the kernel's
  // full budget.  Always the maximum, never less, so this cannot shrink a limit that
  // b200dp_tiny_mlp_fwd_bwd set for itself.
  if (smem > 48 * 1024)
    B200DP_CUDA_OK(cudaFuncSetAttribute(tiny_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  // the ticket counter is the only part of the shared array that must start at zero
  B200DP_CUDA_OK(cudaMemsetAsync(scratch, 0, sizeof(unsigned int), s));
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  float* part = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  const float inv_count = 1.0f / (float(rows) * float(d.sizes[d.num_layers]));
  tiny_mlp_kernel<<<g, kThreads, smem, s>>>(d, params, x, y, rows, inv_count, grads, loss, part, ticket);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}
