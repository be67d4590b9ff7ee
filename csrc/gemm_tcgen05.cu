// Linear-layer GEMMs on the 5th-gen tensor cores: tcgen05.mma with TMEM accumulators, operands
// staged by TMA (128B swizzle) through a 4-stage mbarrier ring, persistent over output tiles.
//
//   D[M,N] = epilogue( A[M,K] * B[N,K]^T )        bf16 in, fp32 accumulate
//
// One kernel covers the three products of a Linear layer (the reference's
// aten::addmm forward at ref.py:42-44,170 and the two mm's of AddmmBackward0 at ref.py:176):
//
//   forward   Y  = relu?(X W^T + b)   A = X  [M,K]  K-major     B = W  [N,K]  K-major
//   dgrad     dX = (dY W) * [X > 0]   A = dY [M,K'] K-major     B = W  [K',N] MN-major  (K' = out features)
//   wgrad     dW = dY^T X             A = dY [K',M] MN-major    B = X  [K',N] MN-major  (K' = batch rows)
//
// so no transposed copies of activations or weights are ever materialised: the MN-major cases are
// expressed in the UMMA shared-memory descriptor (instruction-descriptor bits 15/16) and the TMA
// box shape.
//
// Warp roles (192 threads): warp 0 = TMA producer (one lane), warp 1 = TMEM allocator + MMA issuer
// (one lane), warps 2..5 = epilogue (TMEM -> registers -> bias/ReLU/mask -> global).  Two TMEM
// accumulator stages let the epilogue of tile i overlap the main loop of tile i+1.
#include "common.cuh"

namespace b200dp {
namespace {

constexpr int BM = 128;        // UMMA M (cta_group::1)
constexpr int BK = 64;         // one 128-byte swizzle atom of bf16 along K
constexpr int UMMA_K = 16;
constexpr int kStages = 4;
constexpr int kThreads = 192;
constexpr int kEpiWarp0 = 2;

struct GemmParams {
  int M, N, K;
  void* D;
  long long ldd;
  const float* bias;
  const __nv_bfloat16* mask;
  long long ldmask;
  int relu;
  int d_fp32;
  int num_m_blocks, num_n_blocks;
  int group_m;          // rasterisation: tiles are visited in bands of group_m row-blocks
};

// Tile index -> (row block, column block).  Consecutive tiles (= the CTAs of one wave) walk down
// a band of `group_m` row blocks before moving to the next column block, so a wave of 148 CTAs
// touches ~16 row blocks x ~9 column blocks: both operand panels stay L2-resident instead of
// streaming the whole A matrix through L2 once per wave.
__device__ __forceinline__ void tile_coords(const GemmParams& p, int t, int& m_blk, int& n_blk) {
  const int per_group = p.group_m * p.num_n_blocks;
  const int g = t / per_group;
  const int r = t - g * per_group;
  const int m_first = g * p.group_m;
  const int gsize = min(p.group_m, p.num_m_blocks - m_first);
  n_blk = r / gsize;
  m_blk = m_first + (r - n_blk * gsize);
}

template <int BN>
struct Smem {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarrierOffset = kStages * kStageBytes;
  static constexpr int kTotal = kBarrierOffset + 256 + 1024;  // + barriers + alignment slack
};

// UMMA shared-memory descriptor, SWIZZLE_128B (layout type 2), descriptor version 1 (sm_100).
// K-major : rows of 128 B (64 bf16 along K); 8-row groups 1024 B apart (SBO).
// MN-major: k-rows of 128 B (64 bf16 along M/N); 8-k groups 1024 B apart (SBO); the next 64
//           elements along M/N live in the next TMA box, BK*128 B further (LBO).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, bool mn_major) {
  const uint64_t lbo = mn_major ? uint64_t((BK * 128) >> 4) : 1ull;
  const uint64_t sbo = 1024 >> 4;
  return uint64_t((smem_addr & 0x3FFFF) >> 4) | (lbo << 16) | (sbo << 32) | (1ull << 46) | (2ull << 61);
}

template <int BN, bool A_MN, bool B_MN>
__device__ __forceinline__ constexpr uint32_t make_idesc() {
  // c_format F32 (bits 4-5 = 1), a/b format BF16 (bits 7-9 / 10-12 = 1), majors, N>>3, M>>4
  return (1u << 4) | (1u << 7) | (1u << 10) | (uint32_t(A_MN) << 15) | (uint32_t(B_MN) << 16) |
         (uint32_t(BN >> 3) << 17) | (uint32_t(BM >> 4) << 24);
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmParams p) {
  using S = Smem<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarrierOffset);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.num_m_blocks * p.num_n_blocks;
  const int num_kb = (p.K + BK - 1) / BK;
  constexpr uint32_t kTmemCols = 2 * BN;   // two accumulator stages (power of two: 256 or 512)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 4);    // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(p, t, m_blk, n_blk);
        const int m0 = m_blk * BM;
        const int n0 = n_blk * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * S::kStageBytes;
          uint8_t* sb = sa + S::kABytes;
          mbar_expect_tx(&full_bar[stage], S::kStageBytes);
          const int k0 = kb * BK;
          if constexpr (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * (BK * 128), &tmA, &full_bar[stage], m0 + j * 64, k0);
          } else {
            tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);
          }
          if constexpr (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), &tmB, &full_bar[stage], n0 + j * 64, k0);
          } else {
            tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc<BN, A_MN, B_MN>();
      // descriptor start-address step per UMMA_K (16 bf16 along K), in 16-byte units
      constexpr uint32_t a_step = A_MN ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      constexpr uint32_t b_step = B_MN ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * S::kStageBytes);
          const uint64_t adesc = make_smem_desc(sa, A_MN);
          const uint64_t bdesc = make_smem_desc(sa + S::kABytes, B_MN);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            umma_bf16(d_tmem, adesc + uint64_t(k * a_step), bdesc + uint64_t(k * b_step), idesc,
                      (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);          // smem slot reusable once these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tmem_full[as]);               // accumulator complete -> epilogue
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int quad = warp & 3;                     // TMEM lane quadrant this warp may access
    int it = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      int m_blk, n_blk;
      tile_coords(p, t, m_blk, n_blk);
      const int m0 = m_blk * BM;
      const int n0 = n_blk * BN;
      const int row = m0 + quad * 32 + lane;
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t(quad * 32) << 16) + uint32_t(as * BN);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(taddr + c * 32, v);
        tmem_ld_wait();
        const int col0 = n0 + c * 32;
        if (row < p.M && col0 < p.N) {
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if (p.bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (col0 + j < p.N) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
                f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
              }
            }
          }
          if (p.relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
          }
          if (p.mask != nullptr) {
            const __nv_bfloat16* mrow = p.mask + (long long)row * p.ldmask + col0;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              if (col0 + j < p.N) {
                const uint4 m4 = __ldg(reinterpret_cast<const uint4*>(mrow + j));
                const uint32_t mw[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  if (!(bf16_lo(mw[q]) > 0.0f)) f[j + 2 * q] = 0.0f;
                  if (!(bf16_hi(mw[q]) > 0.0f)) f[j + 2 * q + 1] = 0.0f;
                }
              }
            }
          }
          if (p.d_fp32) {
            float* drow = reinterpret_cast<float*>(p.D) + (long long)row * p.ldd + col0;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (col0 + j < p.N)
                *reinterpret_cast<float4*>(drow + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
            }
          } else {
            __nv_bfloat16* drow = reinterpret_cast<__nv_bfloat16*>(p.D) + (long long)row * p.ldd + col0;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              if (col0 + j < p.N) {
                uint4 o;
                o.x = pack_bf16x2(f[j], f[j + 1]);
                o.y = pack_bf16x2(f[j + 2], f[j + 3]);
                o.z = pack_bf16x2(f[j + 4], f[j + 5]);
                o.w = pack_bf16x2(f[j + 6], f[j + 7]);
                *reinterpret_cast<uint4*>(drow + j) = o;
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[as]);
    }
  }

  __syncwarp();                        // re-converge: only lane 0 of the producer / MMA warps ran a role loop
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ----------------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}

// 2-D bf16 tensor map over a row-major [rows, cols] matrix with leading dimension ld (elements),
// box = [box_rows, 64 cols] (64 bf16 = 128 B = one swizzle atom), 128B swizzle, zero OOB fill.
int make_tensor_map(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld,
                    int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return fail("cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {cuuint64_t(cols), cuuint64_t(rows)};
  cuuint64_t strides[1] = {cuuint64_t(ld) * 2};
  cuuint32_t box[2] = {64u, cuuint32_t(box_rows)};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with CUresult %d", int(r));
  return B200DP_OK;
}

template <int BN, bool A_MN, bool B_MN>
int launch(const b200dp_gemm_args* a, cudaStream_t stream) {
  using S = Smem<BN>;
  CUtensorMap tmA, tmB;
  // K-major operand [MN, K] (ld along K): box = [BM or BN rows] x 64 k.
  // MN-major operand stored [K, MN] (ld along MN): box = 64 k-rows x 64 mn.
  int rc;
  if (A_MN) rc = make_tensor_map(&tmA, a->A, a->K, a->M, a->lda, BK);
  else rc = make_tensor_map(&tmA, a->A, a->M, a->K, a->lda, BM);
  if (rc) return rc;
  if (B_MN) rc = make_tensor_map(&tmB, a->B, a->K, a->N, a->ldb, BK);
  else rc = make_tensor_map(&tmB, a->B, a->N, a->K, a->ldb, BN);
  if (rc) return rc;

  GemmParams p;
  p.M = a->M; p.N = a->N; p.K = a->K;
  p.D = a->D; p.ldd = a->ldd;
  p.bias = a->bias;
  p.mask = reinterpret_cast<const __nv_bfloat16*>(a->mask_src);
  p.ldmask = a->ldmask;
  p.relu = a->relu; p.d_fp32 = a->d_fp32;
  p.num_m_blocks = (a->M + BM - 1) / BM;
  p.num_n_blocks = (a->N + BN - 1) / BN;
  p.group_m = 16;
  const int tiles = p.num_m_blocks * p.num_n_blocks;
  int grid = a->max_ctas > 0 ? a->max_ctas : sm_count();
  if (grid > tiles) grid = tiles;

  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN>;
  static bool attr_set_dev[64] = {false};     // per device: the attribute belongs to the device's context
  int dev_ = 0;
  cudaGetDevice(&dev_);
  bool& attr_set = attr_set_dev[dev_ & 63];
  if (!attr_set) {
    B200DP_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal));
    attr_set = true;
  }
  kern<<<grid, kThreads, S::kTotal, stream>>>(tmA, tmB, p);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

template <int BN>
int dispatch_major(const b200dp_gemm_args* a, cudaStream_t s) {
  if (!a->a_mn_major && !a->b_mn_major) return launch<BN, false, false>(a, s);
  if (!a->a_mn_major && a->b_mn_major) return launch<BN, false, true>(a, s);
  if (a->a_mn_major && a->b_mn_major) return launch<BN, true, true>(a, s);
  return launch<BN, true, false>(a, s);
}

}  // namespace

int gemm_pair_dispatch(const b200dp_gemm_args* a, cudaStream_t s);   // gemm_tcgen05_2cta.cu

}  // namespace b200dp

extern "C" int b200dp_gemm_bf16(const b200dp_gemm_args* a, void* stream) {
  using namespace b200dp;
  B200DP_REQUIRE(a != nullptr, "gemm: null args");
  B200DP_REQUIRE(a->M > 0 && a->N > 0 && a->K > 0, "gemm: M, N, K must be positive (got %d %d %d)", a->M, a->N, a->K);
  B200DP_REQUIRE(a->A && a->B && a->D, "gemm: null operand pointer");
  B200DP_REQUIRE(a->lda % 8 == 0 && a->ldb % 8 == 0, "gemm: lda/ldb must be multiples of 8 elements (16 bytes)");
  B200DP_REQUIRE((reinterpret_cast<uintptr_t>(a->A) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->B) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(a->D) & 15) == 0,
                 "gemm: A, B, D must be 16-byte aligned");
  B200DP_REQUIRE(a->N % 8 == 0, "gemm: N must be a multiple of 8 (vectorised epilogue); got %d", a->N);
  B200DP_REQUIRE(a->ldd % (a->d_fp32 ? 4 : 8) == 0, "gemm: ldd must keep rows 16-byte aligned");
  if (a->mask_src) {
    B200DP_REQUIRE(a->ldmask % 8 == 0 && (reinterpret_cast<uintptr_t>(a->mask_src) & 15) == 0,
                   "gemm: mask rows must be 16-byte aligned");
  }
  if (a->bias) B200DP_REQUIRE((reinterpret_cast<uintptr_t>(a->bias) & 15) == 0, "gemm: bias must be 16-byte aligned");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  B200DP_REQUIRE(a->cta_group >= 0 && a->cta_group <= 2, "gemm: cta_group must be 0, 1 or 2");
  // auto: CTA pairs (256x256 tiles, measured 1.50-1.61 PFLOP/s vs 1.32-1.44 for single CTAs at
  // 8192x4096x4096) whenever there are enough 256x256 tiles to occupy most of the 74 SM pairs
  const long long tiles2 = (long long)((a->M + 255) / 256) * ((a->N + 255) / 256);
  const bool pair = a->cta_group == 2 || (a->cta_group == 0 && a->tile_n == 0 && tiles2 >= (3 * (sm_count() / 2)) / 4);
  if (pair) return gemm_pair_dispatch(a, s);
  // 128x256 tiles unless the problem is too narrow to fill the machine with them
  const bool narrow = a->N <= 128 ||
                      ((long long)((a->M + 127) / 128) * ((a->N + 255) / 256) < sm_count() && a->N % 256 != 0);
  if (a->tile_n == 128 || (a->tile_n == 0 && narrow)) return dispatch_major<128>(a, s);
  return dispatch_major<256>(a, s);
}
