// CTA-pair variant of the Linear GEMM (tcgen05.mma.cta_group::2): two SMs of one TPC work on one
// 256x256 output tile.  Each CTA stages its own 128 rows of A and HALF of the B tile (128 of the 256
// columns); the pair's tensor cores read both halves, so per-SM shared-memory and L2->SM traffic
// drop by a third (32 KB instead of 48 KB per 64-deep k-block) and the freed shared memory buys a
// 6-stage TMA ring instead of 4.  Same math, epilogue (gemm_common.cuh) and operand layouts as
// gemm_tcgen05.cu (forward / dgrad / wgrad of ref.py:42-44,170,176 without transposed copies).
//
// Protocol differences to the single-CTA kernel:
//   * cluster of 2 CTAs; CTA rank 0 (the leader) issues every MMA for the pair;
//   * both producers credit their TMA bytes to the LEADER's "full" barrier (TMA .cta_group::2);
//   * the leader's tcgen05.commit multicasts the "slot free" and "accumulator ready" arrivals to both
//     CTAs; the peer's epilogue warps arrive remotely on the leader's "accumulator drained" barrier;
//   * TMEM is allocated/freed with cta_group::2 by the same warp of both CTAs, bracketed by cluster
//     barriers so neither CTA exits while its partner can still touch its shared memory or TMEM.
//
// Tail wave: T tiles over P pairs leave a last wave of T mod P tiles.  When that remainder would
// keep at most half of the pairs busy (a 4096x4096 wgrad: 256 tiles = 3 x 74 + 34), its tiles are
// computed as two independent 256x128 half tiles each (UMMA N = 128), so the last wave costs half
// a tile time instead of a whole one.  Half tiles need no workspace, no inter-CTA dependency and
// sum in the same order as whole tiles: results are bit-identical with and without the split.
#include "gemm_common.cuh"

namespace b200dp {
namespace {

using T = __nv_bfloat16;
constexpr int BM = kGemmBM;    // rows per CTA (UMMA M = 256 per pair)
constexpr int BN = 256;        // columns per pair
constexpr int BNH = 128;       // B columns staged by each CTA
constexpr int BK = Elem<T>::BK;
constexpr int UMMA_K = Elem<T>::UMMA_K;
constexpr int kABytes = BM * BK * 2;
constexpr int kBBytes = BNH * BK * 2;
constexpr int kStageBytes = kABytes + kBBytes;

// Depth of the TMA ring.  Forward GEMMs (both operands K-major) run alone and take 6 stages (192 KB).
// Backward's GEMMs (dgrad, wgrad: an MN-major operand) run BESIDE the gradient-sync / update kernels
// of the layers above; those kernels live on ordinary loads whose in-flight data sits in L1, and L1 is
// what the SM has left after the shared-memory carve-out.  With 6 stages + the column-sum staging the
// GEMM needed the 228 KB carve-out (28 KB of L1): the co-running sync kernel crawled and dragged the
// GEMM with it (2-GPU step 3.59 ms).  5 stages fit the 196 KB carve-out (60 KB of L1): 3.44 ms, for
// +0.5 % on a GEMM that runs alone.
template <bool A_MN, bool B_MN>
struct PairSmem {
  static constexpr int kStages = (A_MN || B_MN) ? 5 : 6;
  static constexpr int kBarrierOffset = kStages * kStageBytes;
  static constexpr int kColsumOffset = kBarrierOffset + 256;
  // the column sums are fused into dgrad only (A K-major, B MN-major)
  static constexpr int kColsumBytes = (!A_MN && B_MN) ? 2 * 4 * BN * 4 : 0;
  static constexpr int kTotal = kColsumOffset + kColsumBytes + 1024;
};

// work unit -> tile (and which half of it, -1 = the whole tile)
struct Unit { int m_blk, n_blk, half; };
__device__ __forceinline__ Unit unit_of(const GemmParams& p, int u) {
  Unit w;
  int t = u;
  w.half = -1;
  if (u >= p.full_tiles) {
    const int idx = u - p.full_tiles;
    t = p.full_tiles + (idx >> 1);
    w.half = idx & 1;
  }
  tile_coords(p, t, w.m_blk, w.n_blk);
  return w;
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

template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, kGemmMinCtas)
gemm_bf16_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmBh, const GemmParams p) {
  using S = PairSmem<A_MN, B_MN>;
  constexpr int kStages = S::kStages;
  constexpr int kBarrierOffset = S::kBarrierOffset;
  constexpr int kColsumOffset = S::kColsumOffset;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kBarrierOffset);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* cs_smem = reinterpret_cast<float*>(smem + kColsumOffset);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int num_tiles = p.num_m_blocks * p.num_n_blocks;
  const int num_units = p.full_tiles + 2 * (num_tiles - p.full_tiles);
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
      for (int u = cluster_id; u < num_units; u += num_clusters) {
        const Unit w = unit_of(p, u);
        const bool whole = w.half < 0;
        const int b_rows = whole ? BNH : BNH / 2;              // B columns this CTA stages
        const int m0 = w.m_blk * (2 * BM) + int(rank) * BM;    // this CTA's rows of A
        const int n0 = w.n_blk * BN + (whole ? 0 : w.half * BNH) + int(rank) * b_rows;
        const uint32_t tx = 2u * uint32_t(kABytes + b_rows * BK * 2);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb = sa + kABytes;
          const uint32_t leader_full = map_to_cta(smem_u32(&full_bar[stage]), 0);
          if (leader) mbar_expect_tx(&full_bar[stage], tx);
          const int k0 = kb * BK;
          if constexpr (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_2d_pair(sa + j * (BK * 128), &tmA, leader_full, m0 + j * 64, k0);
          } else {
            tma_load_2d_pair(sa, &tmA, leader_full, k0, m0);
          }
          if constexpr (B_MN) {
            tma_load_2d_pair(sb, &tmB, leader_full, n0, k0);
            if (whole) tma_load_2d_pair(sb + BK * 128, &tmB, leader_full, n0 + 64, k0);
          } else {
            tma_load_2d_pair(sb, whole ? &tmB : &tmBh, leader_full, k0, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc_whole = make_idesc<T>(2 * BM, BN, A_MN, B_MN);
      constexpr uint32_t idesc_half = make_idesc<T>(2 * BM, BN / 2, A_MN, B_MN);
      constexpr uint32_t a_step = A_MN ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      constexpr uint32_t b_step = B_MN ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int u = cluster_id; u < num_units; u += num_clusters, ++it) {
        const uint32_t idesc = u < p.full_tiles ? idesc_whole : idesc_half;
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * kStageBytes);
          const uint64_t adesc = make_smem_desc<BK>(sa, A_MN);
          const uint64_t bdesc = make_smem_desc<BK>(sa + kABytes, B_MN);
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
    for (int u = cluster_id; u < num_units; u += num_clusters, ++it) {
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const Unit w = unit_of(p, u);
      const int m0 = w.m_blk * (2 * BM) + int(rank) * BM;
      const int strip = w.m_blk * 2 + int(rank);
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t(quad * 32) << 16) + uint32_t(as * BN);
      uint64_t* release_bar = &tmem_empty[as];
      auto release = [&]() {
        if (lane == 0) {
          if (leader) mbar_arrive(release_bar);
          else mbar_arrive_cluster(map_to_cta(smem_u32(release_bar), 0));
        }
      };
      if (w.half < 0)
        epilogue_tile<T, BN>(p, taddr, m0, w.n_blk * BN, w.n_blk, strip, quad, lane, cs_smem, it & 1, release);
      else      // half tiles are never combined with the fused column sums (see launch_pair)
        epilogue_tile<T, BN / 2>(p, taddr, m0, w.n_blk * BN + w.half * BNH, 0, strip, quad, lane, cs_smem, it & 1,
                                 release);
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

template <bool A_MN, bool B_MN>
int launch_pair(const b200dp_gemm_args* a, cudaStream_t stream) {
  CUtensorMap tmA, tmB, tmBh;
  int rc;
  if (A_MN) rc = get_tensor_map(&tmA, Elem<T>::kTma, 2, a->A, a->K, a->M, a->lda, BK);
  else rc = get_tensor_map(&tmA, Elem<T>::kTma, 2, a->A, a->M, a->K, a->lda, BM);
  if (rc) return rc;
  if (B_MN) rc = get_tensor_map(&tmB, Elem<T>::kTma, 2, a->B, a->K, a->N, a->ldb, BK);
  else rc = get_tensor_map(&tmB, Elem<T>::kTma, 2, a->B, a->N, a->K, a->ldb, BNH);
  if (rc) return rc;
  GemmParams p;
  fill_gemm_params(a, &p);
  p.num_m_blocks = (a->M + 2 * BM - 1) / (2 * BM);
  p.num_n_blocks = (a->N + BN - 1) / BN;
  p.group_m = 8;
  p.cs_ld = p.num_n_blocks * BN;
  const int tiles = p.num_m_blocks * p.num_n_blocks;
  int clusters = (a->max_ctas > 0 ? a->max_ctas : sm_count()) / 2;
  if (clusters < 1) clusters = 1;
  // tail wave as half tiles when that halves its cost (never with the fused column sums, whose
  // tickets count whole 256-column blocks)
  p.full_tiles = tiles;
  const int rem = tiles % clusters;
  B200DP_REQUIRE(a->colsum_out == nullptr || (!A_MN && B_MN),
                 "gemm: the CTA-pair kernel fuses column sums into dgrad (A K-major, B MN-major) only");
  if (a->split_tail >= 0 && tiles > clusters && rem > 0 && 2 * rem <= clusters && a->colsum_out == nullptr)
    p.full_tiles = tiles - rem;
  tmBh = tmB;
  if (!B_MN && p.full_tiles < tiles) {
    rc = get_tensor_map(&tmBh, Elem<T>::kTma, 2, a->B, a->N, a->K, a->ldb, BNH / 2);
    if (rc) return rc;
  }
  const int units = p.full_tiles + 2 * (tiles - p.full_tiles);
  if (clusters > units) clusters = units;
  auto kern = gemm_bf16_pair_kernel<A_MN, B_MN>;
  static bool attr_done[64] = {false};
  constexpr int kSmemTotal = PairSmem<A_MN, B_MN>::kTotal;
  if (int e = ensure_smem_attr(kern, kSmemTotal, attr_done)) return e;
  kern<<<2 * clusters, kGemmThreads, kSmemTotal, stream>>>(tmA, tmB, tmBh, p);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

}  // namespace

int gemm_pair_regs_per_cta() {
  int best = -1;
  const void* ks[] = {(const void*)gemm_bf16_pair_kernel<false, false>, (const void*)gemm_bf16_pair_kernel<false, true>,
                      (const void*)gemm_bf16_pair_kernel<true, true>, (const void*)gemm_bf16_pair_kernel<true, false>};
  for (const void* k : ks) {
    cudaFuncAttributes at;
    if (cudaFuncGetAttributes(&at, k) != cudaSuccess) return -1;
    const int v = ((at.numRegs * 32 + 255) / 256) * 256 * (kGemmThreads / 32);
    if (v > best) best = v;
  }
  return best;
}

// called from b200dp_gemm_bf16 (gemm_tcgen05.cu) after argument validation
int gemm_pair_dispatch(const b200dp_gemm_args* a, cudaStream_t s) {
  if (!a->a_mn_major && !a->b_mn_major) return launch_pair<false, false>(a, s);
  if (!a->a_mn_major && a->b_mn_major) return launch_pair<false, true>(a, s);
  if (a->a_mn_major && a->b_mn_major) return launch_pair<true, true>(a, s);
  return launch_pair<true, false>(a, s);
}

}  // namespace b200dp
