// Linear-layer GEMMs on the 5th-gen tensor cores: tcgen05.mma with TMEM accumulators, operands
// staged by TMA (128B swizzle) through a 4-stage mbarrier ring, persistent over output tiles.
//
//   D[M,N] = epilogue( A[M,K] * B[N,K]^T )        bf16 in (kind::f16) or fp32 in (kind::tf32), fp32 accumulate
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
// box shape.  The fp32 instantiation reads the reference's own precision (ref.py:159 `.float()`)
// straight from fp32 buffers; the tensor core rounds the operands to TF32 and accumulates in fp32.
// It exists for K-major operands only: measured on B200, a kind::tf32 MMA with a transposed operand
// leaves the accumulator untouched, so the fp32 backward runs on transposed copies instead.
//
// Warp roles (192 threads): warp 0 = TMA producer (one lane), warp 1 = TMEM allocator + MMA issuer
// (one lane), warps 2..5 = epilogue (gemm_common.cuh).  Two TMEM accumulator stages let the
// epilogue of tile i overlap the main loop of tile i+1.
#include <mutex>

#include "gemm_common.cuh"

namespace b200dp {
namespace {

constexpr int BM = kGemmBM;    // UMMA M (cta_group::1)
constexpr int kStages = 4;

template <typename T, int BN, bool A_MN>
struct Smem {
  static constexpr int BK = Elem<T>::BK;
  static constexpr int kABytes = BM * 128;          // BM rows x one 128-byte swizzle row of K
  static constexpr int kBBytes = BN * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarrierOffset = kStages * kStageBytes;
  static constexpr int kColsumOffset = kBarrierOffset + 256;
  // column sums are fused into GEMMs whose A operand is K-major (dgrad; wgrad never asks for them):
  // keeping the staging out of the wgrad instantiation keeps it within the 196 KB shared-memory
  // carve-out, i.e. leaves 60 KB of L1 to the kernels that run beside it (see gemm_tcgen05_2cta.cu)
  static constexpr int kColsumBytes = A_MN ? 0 : 2 * 4 * BN * 4;
  static constexpr int kTotal = kColsumOffset + kColsumBytes + 1024;   // + alignment slack
};

template <typename T, int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kGemmThreads, kGemmMinCtas)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
  using S = Smem<T, BN, A_MN>;
  constexpr int BK = Elem<T>::BK;
  constexpr int UMMA_K = Elem<T>::UMMA_K;
  constexpr int E = 128 / Elem<T>::kBytes;          // elements per swizzle row (= BK)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::kBarrierOffset);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* cs_smem = reinterpret_cast<float*>(smem + S::kColsumOffset);

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
            for (int j = 0; j < BM / E; ++j) tma_load_2d(sa + j * (BK * 128), &tmA, &full_bar[stage], m0 + j * E, k0);
          } else {
            tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);
          }
          if constexpr (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / E; ++j) tma_load_2d(sb + j * (BK * 128), &tmB, &full_bar[stage], n0 + j * E, k0);
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
      constexpr uint32_t idesc = make_idesc<T>(BM, BN, A_MN, B_MN);
      // descriptor start-address step per UMMA (32 bytes of K per row), in 16-byte units
      constexpr uint32_t a_step = A_MN ? (UMMA_K * 128) >> 4 : 32 >> 4;
      constexpr uint32_t b_step = B_MN ? (UMMA_K * 128) >> 4 : 32 >> 4;
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
          const uint64_t adesc = make_smem_desc<BK>(sa, A_MN);
          const uint64_t bdesc = make_smem_desc<BK>(sa + S::kABytes, B_MN);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint32_t acc = (kb > 0 || k > 0) ? 1u : 0u;
            if constexpr (Elem<T>::kTf32)
              umma_tf32(d_tmem, adesc + uint64_t(k * a_step), bdesc + uint64_t(k * b_step), idesc, acc);
            else
              umma_bf16(d_tmem, adesc + uint64_t(k * a_step), bdesc + uint64_t(k * b_step), idesc, acc);
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
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t(quad * 32) << 16) + uint32_t(as * BN);
      uint64_t* release_bar = &tmem_empty[as];
      epilogue_tile<T, BN>(p, taddr, m_blk * BM, n_blk * BN, n_blk, m_blk, quad, lane, cs_smem, it & 1,
                           [&]() { if (lane == 0) mbar_arrive(release_bar); });
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

template <typename T, int BN, bool A_MN, bool B_MN>
int launch(const b200dp_gemm_args* a, cudaStream_t stream) {
  using S = Smem<T, BN, A_MN>;
  constexpr int BK = Elem<T>::BK;
  constexpr int EB = Elem<T>::kBytes;
  CUtensorMap tmA, tmB;
  // K-major operand [MN, K] (ld along K): box = [BM or BN rows] x 128 bytes of k.
  // MN-major operand stored [K, MN] (ld along MN): box = BK k-rows x 128 bytes of mn.
  int rc;
  if (A_MN) rc = get_tensor_map(&tmA, Elem<T>::kTma, EB, a->A, a->K, a->M, a->lda, BK);
  else rc = get_tensor_map(&tmA, Elem<T>::kTma, EB, a->A, a->M, a->K, a->lda, BM);
  if (rc) return rc;
  if (B_MN) rc = get_tensor_map(&tmB, Elem<T>::kTma, EB, a->B, a->K, a->N, a->ldb, BK);
  else rc = get_tensor_map(&tmB, Elem<T>::kTma, EB, a->B, a->N, a->K, a->ldb, BN);
  if (rc) return rc;

  B200DP_REQUIRE(a->colsum_out == nullptr || !A_MN, "gemm: column sums need a K-major A operand");
  GemmParams p;
  fill_gemm_params(a, &p);
  p.num_m_blocks = (a->M + BM - 1) / BM;
  p.num_n_blocks = (a->N + BN - 1) / BN;
  p.group_m = 16;
  p.cs_ld = p.num_n_blocks * BN;
  const int tiles = p.num_m_blocks * p.num_n_blocks;
  p.full_tiles = tiles;
  int grid = a->max_ctas > 0 ? a->max_ctas : sm_count();
  if (grid > tiles) grid = tiles;

  auto kern = gemm_kernel<T, BN, A_MN, B_MN>;
  static bool attr_done[64] = {false};
  if (int e = ensure_smem_attr(kern, S::kTotal, attr_done)) return e;
  kern<<<grid, kGemmThreads, S::kTotal, stream>>>(tmA, tmB, p);
  B200DP_CUDA_OK(cudaGetLastError());
  return B200DP_OK;
}

template <typename T, int BN>
int dispatch_major(const b200dp_gemm_args* a, cudaStream_t s) {
  if (!a->a_mn_major && !a->b_mn_major) return launch<T, BN, false, false>(a, s);
  if constexpr (!Elem<T>::kTf32) {
    if (!a->a_mn_major && a->b_mn_major) return launch<T, BN, false, true>(a, s);
    if (a->a_mn_major && a->b_mn_major) return launch<T, BN, true, true>(a, s);
    return launch<T, BN, true, false>(a, s);
  }
  return fail("gemm_tf32: operands must be K-major");
}

int validate(const b200dp_gemm_args* a, int elem_bytes) {
  B200DP_REQUIRE(a != nullptr, "gemm: null args");
  B200DP_REQUIRE(a->M > 0 && a->N > 0 && a->K > 0, "gemm: M, N, K must be positive (got %d %d %d)", a->M, a->N, a->K);
  B200DP_REQUIRE(a->A && a->B && a->D, "gemm: null operand pointer");
  const int vec = 16 / elem_bytes;
  B200DP_REQUIRE(a->lda % vec == 0 && a->ldb % vec == 0, "gemm: lda/ldb must be multiples of %d elements (16 bytes)", vec);
  B200DP_REQUIRE((reinterpret_cast<uintptr_t>(a->A) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->B) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(a->D) & 15) == 0,
                 "gemm: A, B, D must be 16-byte aligned");
  B200DP_REQUIRE(a->N % 8 == 0, "gemm: N must be a multiple of 8 (vectorised epilogue); got %d", a->N);
  B200DP_REQUIRE(a->ldd % (a->d_fp32 ? 4 : 8) == 0, "gemm: ldd must keep rows 16-byte aligned");
  if (a->mask_src) {
    B200DP_REQUIRE(a->ldmask % vec == 0 && (reinterpret_cast<uintptr_t>(a->mask_src) & 15) == 0,
                   "gemm: mask rows must be 16-byte aligned");
  }
  if (a->bias) B200DP_REQUIRE((reinterpret_cast<uintptr_t>(a->bias) & 15) == 0, "gemm: bias must be 16-byte aligned");
  B200DP_REQUIRE(a->cta_group >= 0 && a->cta_group <= 2, "gemm: cta_group must be 0, 1 or 2");
  if (a->colsum_out) {
    B200DP_REQUIRE(a->colsum_scratch != nullptr, "gemm: colsum_out needs colsum_scratch");
    B200DP_REQUIRE(a->colsum_scratch_bytes >= b200dp_gemm_colsum_scratch_bytes(a->M, a->N),
                   "gemm: colsum scratch too small (%lld < %lld bytes)", a->colsum_scratch_bytes,
                   b200dp_gemm_colsum_scratch_bytes(a->M, a->N));
    B200DP_REQUIRE((reinterpret_cast<uintptr_t>(a->colsum_scratch) & 15) == 0, "gemm: colsum scratch must be 16-byte aligned");
  }
  return B200DP_OK;
}

// 128x256 tiles unless the problem is too narrow to fill the machine with them
bool narrow_problem(const b200dp_gemm_args* a) {
  return a->N <= 128 ||
         ((long long)((a->M + 127) / 128) * ((a->N + 255) / 256) < sm_count() && a->N % 256 != 0);
}

// ---- tensor-map cache ----------------------------------------------------------------------------
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

struct MapKey {
  const void* base;
  long long rows, cols, ld;
  int box_rows, dt;
  bool operator==(const MapKey& o) const {
    return base == o.base && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows && dt == o.dt;
  }
};
constexpr int kMapCacheSize = 256;      // direct-mapped; a deep MLP step uses ~6 maps per layer
struct MapEntry { MapKey key; CUtensorMap map; bool valid; };
MapEntry g_map_cache[kMapCacheSize];
std::mutex g_map_mutex;

}  // namespace

// A tensor map encodes only (address, shape, strides, box): it stays valid for as long as those
// describe the caller's buffer, whatever the buffer holds, so caching by value is safe.
int get_tensor_map(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, const void* base, long long rows,
                   long long cols, long long ld, int box_rows) {
  const MapKey key{base, rows, cols, ld, box_rows, int(dt)};
  uint64_t h = reinterpret_cast<uintptr_t>(base) >> 4;
  h = (h ^ uint64_t(rows) * 0x9E3779B97F4A7C15ull) + uint64_t(cols) * 31 + uint64_t(ld) * 131 + uint64_t(box_rows) * 7 + uint64_t(dt);
  h ^= h >> 29;
  MapEntry& e = g_map_cache[h % kMapCacheSize];
  {
    std::lock_guard<std::mutex> g(g_map_mutex);
    if (e.valid && e.key == key) {
      *out = e.map;
      return B200DP_OK;
    }
  }
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return fail("cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {cuuint64_t(cols), cuuint64_t(rows)};
  cuuint64_t strides[1] = {cuuint64_t(ld) * cuuint64_t(elem_bytes)};
  cuuint32_t box[2] = {cuuint32_t(128 / elem_bytes), cuuint32_t(box_rows)};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with CUresult %d", int(r));
  std::lock_guard<std::mutex> g(g_map_mutex);
  e.key = key;
  e.map = *out;
  e.valid = true;
  return B200DP_OK;
}

void fill_gemm_params(const b200dp_gemm_args* a, GemmParams* p) {
  memset(p, 0, sizeof(*p));
  p->M = a->M; p->N = a->N; p->K = a->K;
  p->D = a->D; p->ldd = a->ldd;
  p->bias = a->bias;
  p->mask = a->mask_src;
  p->ldmask = a->ldmask;
  p->relu = a->relu; p->d_fp32 = a->d_fp32;
  if (a->colsum_out) {
    // scratch: [tickets: 4 KB][partials: strips x cs_ld floats]
    p->cs_tickets = reinterpret_cast<unsigned int*>(a->colsum_scratch);
    p->cs_partials = reinterpret_cast<float*>(reinterpret_cast<char*>(a->colsum_scratch) + 4096);
    p->cs_out = a->colsum_out;
    p->cs_out_bf16 = a->colsum_out_bf16;
    p->cs_strips = (a->M + kGemmBM - 1) / kGemmBM;
  }
}

int gemm_pair_dispatch(const b200dp_gemm_args* a, cudaStream_t s);   // gemm_tcgen05_2cta.cu
int gemm_pair_regs_per_cta();                                           // gemm_tcgen05_2cta.cu
int sync_regs_per_cta();                                                // allreduce_sgd.cu

namespace {
template <typename Kern>
int regs_per_cta(Kern k, int threads) {
  cudaFuncAttributes at;
  if (cudaFuncGetAttributes(&at, k) != cudaSuccess) return -1;
  const int per_warp = ((at.numRegs * 32 + 255) / 256) * 256;     // allocation granularity: 256 registers per warp
  return per_warp * (threads / 32);
}
}  // namespace

}  // namespace b200dp

using namespace b200dp;

extern "C" long long b200dp_gemm_colsum_scratch_bytes(int M, int N) {
  const long long strips = (M + kGemmBM - 1) / kGemmBM;
  const long long ld = (long long)((N + 255) / 256) * 256;
  return 4096 + strips * ld * 4;
}

extern "C" int b200dp_gemm_bf16(const b200dp_gemm_args* a, void* stream) {
  if (int rc = validate(a, 2)) return rc;
  B200DP_REQUIRE((a->N + 127) / 128 <= 1024, "gemm: N too large for the fused column sums");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  // auto: CTA pairs (256x256 tiles, measured 1.50-1.61 PFLOP/s vs 1.32-1.44 for single CTAs at
  // 8192x4096x4096) whenever there are enough 256x256 tiles to occupy most of the 74 SM pairs
  const long long tiles2 = (long long)((a->M + 255) / 256) * ((a->N + 255) / 256);
  const bool pair = a->cta_group == 2 || (a->cta_group == 0 && a->tile_n == 0 && tiles2 >= (3 * (sm_count() / 2)) / 4);
  if (pair) return gemm_pair_dispatch(a, s);
  if (a->tile_n == 128 || (a->tile_n == 0 && narrow_problem(a))) return dispatch_major<__nv_bfloat16, 128>(a, s);
  return dispatch_major<__nv_bfloat16, 256>(a, s);
}

// Same contract with fp32 operands (A, B, mask_src fp32; D fp32), K-major only: the tensor cores read
// them as TF32 (10-bit mantissa) and accumulate in fp32.  This is the path of `--dtype fp32` models that are too
// large for the fused SIMT kernel - the reference trains in fp32 only (ref.py:159).
extern "C" int b200dp_gemm_tf32(const b200dp_gemm_args* a, void* stream) {
  if (int rc = validate(a, 4)) return rc;
  B200DP_REQUIRE(a->d_fp32 != 0, "gemm_tf32: D must be fp32");
  // Measured on B200: tcgen05.mma.kind::tf32 with a transposed (MN-major) operand leaves the accumulator
  // untouched - TF32 operands are read K-major only (bf16 has no such limit).  Backward therefore works on
  // transposed copies (b200dp_transpose_f32).
  B200DP_REQUIRE(!a->a_mn_major && !a->b_mn_major,
                 "gemm_tf32: kind::tf32 reads K-major operands only - pass transposed copies (b200dp_transpose_f32)");
  B200DP_REQUIRE(a->cta_group != 2, "gemm_tf32: the CTA-pair kernel is bf16 only");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (a->tile_n == 128 || (a->tile_n == 0 && narrow_problem(a))) return dispatch_major<float, 128>(a, s);
  return dispatch_major<float, 256>(a, s);
}

// Registers one GEMM CTA and one fused-sync CTA occupy on an SM.  Their sum must stay within the SM's
// 65536 registers or the sync kernels cannot run beside backward's GEMMs (see kGemmMinCtas).
extern "C" int b200dp_overlap_register_budget(int* gemm_regs_per_cta, int* sync_regs_per_cta_out) {
  B200DP_REQUIRE(gemm_regs_per_cta && sync_regs_per_cta_out, "register_budget: null pointer");
  int g = gemm_pair_regs_per_cta();
  const int cands[] = {
      regs_per_cta(gemm_kernel<__nv_bfloat16, 256, false, false>, kGemmThreads),
      regs_per_cta(gemm_kernel<__nv_bfloat16, 256, false, true>, kGemmThreads),
      regs_per_cta(gemm_kernel<__nv_bfloat16, 256, true, true>, kGemmThreads),
      regs_per_cta(gemm_kernel<__nv_bfloat16, 128, false, false>, kGemmThreads),
      regs_per_cta(gemm_kernel<__nv_bfloat16, 128, false, true>, kGemmThreads),
      regs_per_cta(gemm_kernel<__nv_bfloat16, 128, true, true>, kGemmThreads),
      regs_per_cta(gemm_kernel<float, 256, false, false>, kGemmThreads),
      regs_per_cta(gemm_kernel<float, 128, false, false>, kGemmThreads)};
  for (int c : cands) {
    B200DP_REQUIRE(c > 0, "register_budget: cudaFuncGetAttributes failed");
    if (c > g) g = c;
  }
  *gemm_regs_per_cta = g;
  *sync_regs_per_cta_out = sync_regs_per_cta();
  B200DP_REQUIRE(g > 0 && *sync_regs_per_cta_out > 0, "register_budget: cudaFuncGetAttributes failed");
  return B200DP_OK;
}
