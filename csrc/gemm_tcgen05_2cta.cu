// CTA-pair variant of the Linear GEMM (tcgen05.mma.cta_group::2): two SMs of one TPC work on one
// 256x256 output tile.  Each CTA stages its own 128 rows of A and HALF of the B tile (128 of the 256
// columns); the pair's tensor cores read both halves, so per-SM shared-memory and L2->SM traffic
// drop by a third (32 KB instead of 48 KB per 64-deep k-block) and the freed shared memory buys a
// 6-stage TMA ring instead of 4.  Same math, epilogues and operand layouts as gemm_tcgen05.cu
// (forward / dgrad / wgrad of ref.py:42-44,170,176 without transposed copies).
//
// Protocol differences to the single-CTA kernel:
//   * cluster of 2 CTAs; CTA rank 0 (the leader) issues every MMA for the pair;
//   * both producers credit their TMA bytes to the LEADER's "full" barrier (TMA .cta_group::2);
//   * the leader's tcgen05.commit multicasts the "slot free" and "accumulator ready" arrivals to both
//     CTAs; the peer's epilogue warps arrive remotely on the leader's "accumulator drained" barrier;
//   * TMEM is allocated/freed with cta_group::2 by the same warp of both CTAs, bracketed by cluster
//     barriers so neither CTA exits while its partner can still touch its shared memory or TMEM.
#include "common.cuh"

namespace b200dp {
namespace {

constexpr int BM = 128;        // rows per CTA (UMMA M = 256 per pair)
constexpr int BN = 256;        // columns per pair
constexpr int BNH = 128;       // B columns staged by each CTA
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int kStages = 6;
constexpr int kThreads = 192;

constexpr int kABytes = BM * BK * 2;
constexpr int kBBytes = BNH * BK * 2;
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kBarrierOffset = kStages * kStageBytes;
constexpr int kSmemTotal = kBarrierOffset + 256 + 1024;

struct Gemm2Params {
  int M, N, K;
  void* D;
  long long ldd;
  const float* bias;
  const __nv_bfloat16* mask;
  long long ldmask;
  int relu;
  int d_fp32;
  int num_m_blocks, num_n_blocks;   // in 256-row / 256-column tiles
  int group_m;
};

__device__ __forceinline__ void tile_coords2(const Gemm2Params& p, int t, int& m_blk, int& n_blk) {
  const int per_group = p.group_m * p.num_n_blocks;
  const int g = t / per_group;
  const int r = t - g * per_group;
  const int m_first = g * p.group_m;
  const int gsize = min(p.group_m, p.num_m_blocks - m_first);
  n_blk = r / gsize;
  m_blk = m_first + (r - n_blk * gsize);
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_smem_addr` in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion bytes are credited to a barrier given by its shared::cluster address
// (the leader's barrier), as required for operands consumed by a cta_group::2 MMA.
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster_addr,
                                                 int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_result, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
               "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs once the MMAs issued so far retire
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  const uint16_t mask = 0x3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(mask)
      : "memory");
}

__device__ __forceinline__ uint64_t make_smem_desc2(uint32_t smem_addr, bool mn_major) {
  const uint64_t lbo = mn_major ? uint64_t((BK * 128) >> 4) : 1ull;
  const uint64_t sbo = 1024 >> 4;
  return uint64_t((smem_addr & 0x3FFFF) >> 4) | (lbo << 16) | (sbo << 32) | (1ull << 46) | (2ull << 61);
}

template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_bf16_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const Gemm2Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kBarrierOffset);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int num_tiles = p.num_m_blocks * p.num_n_blocks;
  const int num_kb = (p.K + BK - 1) / BK;
  constexpr uint32_t kTmemCols = 2 * BN;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kStages; ++s) {
      // Only the leader's copy is used.  One arrival (the leader's arrive.expect_tx for the bytes of BOTH
      // CTAs); the partner contributes transaction bytes only - a remote arrive per k-block would put a
      // cluster-scope fence into its producer loop.  Its bytes for phase P cannot land before phase P-1
      // completed because its own "empty" wait is released by the leader's MMA of phase P-1.
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);     // one multicast commit from the leader's MMA thread
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 8);    // 4 epilogue warps x 2 CTAs (only the leader's is used)
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc_pair(tmem_slot, kTmemCols);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();                  // partner's barriers are initialised, TMEM is allocated in both
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cluster_id; t < num_tiles; t += num_clusters) {
        int m_blk, n_blk;
        tile_coords2(p, t, m_blk, n_blk);
        const int m0 = m_blk * (2 * BM) + int(rank) * BM;      // this CTA's rows of A
        const int n0 = n_blk * BN + int(rank) * BNH;           // this CTA's half of B
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb = sa + kABytes;
          const uint32_t leader_full = map_to_cta(smem_u32(&full_bar[stage]), 0);
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * kStageBytes);
          const int k0 = kb * BK;
          if constexpr (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_2d_pair(sa + j * (BK * 128), &tmA, leader_full, m0 + j * 64, k0);
          } else {
            tma_load_2d_pair(sa, &tmA, leader_full, k0, m0);
          }
          if constexpr (B_MN) {
#pragma unroll
            for (int j = 0; j < BNH / 64; ++j) tma_load_2d_pair(sb + j * (BK * 128), &tmB, leader_full, n0 + j * 64, k0);
          } else {
            tma_load_2d_pair(sb, &tmB, leader_full, k0, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      // c_format F32, a/b BF16, majors, N = 256 (>>3), M = 256 (>>4)
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (uint32_t(A_MN) << 15) |
                                 (uint32_t(B_MN) << 16) | (uint32_t(BN >> 3) << 17) | (uint32_t((2 * BM) >> 4) << 24);
      constexpr uint32_t a_step = A_MN ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      constexpr uint32_t b_step = B_MN ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = cluster_id; t < num_tiles; t += num_clusters, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * kStageBytes);
          const uint64_t adesc = make_smem_desc2(sa, A_MN);
          const uint64_t bdesc = make_smem_desc2(sa + kABytes, B_MN);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            umma_bf16_pair(d_tmem, adesc + uint64_t(k * a_step), bdesc + uint64_t(k * b_step), idesc,
                           (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit_pair(&tmem_full[as]);
      }
    }
  } else {
    // ===================== epilogue (warps 2..5, both CTAs) =====================
    const int quad = warp & 3;
    int it = 0;
    for (int t = cluster_id; t < num_tiles; t += num_clusters, ++it) {
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      int m_blk, n_blk;
      tile_coords2(p, t, m_blk, n_blk);
      const int m0 = m_blk * (2 * BM) + int(rank) * BM;
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
      if (lane == 0) {
        if (leader) mbar_arrive(&tmem_empty[as]);
        else mbar_arrive_cluster(map_to_cta(smem_u32(&tmem_empty[as]), 0));
      }
    }
  }

  __syncwarp();                        // re-converge: only lane 0 of the producer / MMA warps ran a role loop
  tc_fence_before();
  cluster_sync_all();                  // nobody leaves while the partner may still use this CTA's smem / TMEM
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, kTmemCols);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_tensor_map2(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld, int box_rows) {
  static EncodeTiledFn enc = nullptr;
  if (!enc) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return fail("cuTensorMapEncodeTiled is not available from the driver");
    enc = reinterpret_cast<EncodeTiledFn>(sym);
  }
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

template <bool A_MN, bool B_MN>
int launch_pair(const b200dp_gemm_args* a, cudaStream_t stream) {
  CUtensorMap tmA, tmB;
  int rc;
  if (A_MN) rc = make_tensor_map2(&tmA, a->A, a->K, a->M, a->lda, BK);
  else rc = make_tensor_map2(&tmA, a->A, a->M, a->K, a->lda, BM);
  if (rc) return rc;
  if (B_MN) rc = make_tensor_map2(&tmB, a->B, a->K, a->N, a->ldb, BK);
  else rc = make_tensor_map2(&tmB, a->B, a->N, a->K, a->ldb, BNH);
  if (rc) return rc;
  Gemm2Params p;
  p.M = a->M; p.N = a->N; p.K = a->K;
  p.D = a->D; p.ldd = a->ldd;
  p.bias = a->bias;
  p.mask = reinterpret_cast<const __nv_bfloat16*>(a->mask_src);
  p.ldmask = a->ldmask;
  p.relu = a->relu; p.d_fp32 = a->d_fp32;
  p.num_m_blocks = (a->M + 2 * BM - 1) / (2 * BM);
  p.num_n_blocks = (a->N + BN - 1) / BN;
  p.group_m = 8;
  const int tiles = p.num_m_blocks * p.num_n_blocks;
  int clusters = (a->max_ctas > 0 ? a->max_ctas : sm_count()) / 2;
  if (clusters < 1) clusters = 1;
  if (clusters > tiles) clusters = tiles;
  auto kern = gemm_bf16_pair_kernel<A_MN, B_MN>;
  static bool attr_set_dev[64] = {false};     // per device: the attribute belongs to the device's context
  int dev_ = 0;
  cudaGetDevice(&dev_);
  bool& attr_set = attr_set_dev[dev_ & 63];
  if (!attr_set) {
    B200DP_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
    attr_set = true;
  }
  kern<<<2 * clusters, kThreads, kSmemTotal, stream>>>(tmA, tmB, p);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

}  // namespace

// called from b200dp_gemm_bf16 (gemm_tcgen05.cu) after argument validation
int gemm_pair_dispatch(const b200dp_gemm_args* a, cudaStream_t s) {
  if (!a->a_mn_major && !a->b_mn_major) return launch_pair<false, false>(a, s);
  if (!a->a_mn_major && a->b_mn_major) return launch_pair<false, true>(a, s);
  if (a->a_mn_major && a->b_mn_major) return launch_pair<true, true>(a, s);
  return launch_pair<true, false>(a, s);
}

}  // namespace b200dp
