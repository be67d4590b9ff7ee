// Pieces shared by the single-CTA (gemm_tcgen05.cu) and CTA-pair (gemm_tcgen05_2cta.cu) Linear GEMMs:
// element traits (bf16 -> kind::f16, fp32 -> kind::tf32), the persistent tile walk, UMMA shared-memory
// and instruction descriptors, the epilogue (TMEM -> registers -> bias / ReLU / ReLU-mask -> global,
// and the fused column sums that produce the bias gradient of the layer below), and the
// host-side tensor-map cache.
#pragma once
#include <string.h>

#include "common.cuh"

namespace b200dp {

constexpr int kGemmBM = 128;          // rows of D per CTA (= TMEM lanes)
constexpr int kGemmThreads = 192;     // warp 0 TMA producer, warp 1 MMA issuer, warps 2..5 epilogue
constexpr int kEpiThreads = 128;
// Register budget of the GEMM kernels.  A persistent GEMM CTA (192 threads) must leave room on its SM
// for one CTA of the fused gradient sync (256 threads x 128 registers = 32 K of the SM's 64 K
// registers): the sync kernels run on a second stream WHILE backward's GEMMs run, and their cross-GPU
// barrier needs all of their CTAs resident at once.  The kernels compile to ~117 registers (23 K per
// CTA); b200dp_overlap_register_budget + a test keep the sum within the register file.
constexpr int kGemmMinCtas = 1;
constexpr int kEpiBarrierId = 1;      // named barrier of the four epilogue warps

// ---- element traits ---------------------------------------------------------------------------
// One k-block is one 128-byte swizzle atom along K: 64 bf16 or 32 fp32 (read by the tensor core as
// TF32).  One tcgen05.mma consumes 32 bytes of K per row.
template <typename T> struct Elem;
template <> struct Elem<__nv_bfloat16> {
  static constexpr int kBytes = 2, BK = 64, UMMA_K = 16;
  static constexpr uint32_t kIdescFormat = 1;   // BF16 (kind::f16)
  static constexpr bool kTf32 = false;
  static constexpr CUtensorMapDataType kTma = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
};
template <> struct Elem<float> {
  static constexpr int kBytes = 4, BK = 32, UMMA_K = 8;
  static constexpr uint32_t kIdescFormat = 2;   // TF32 (kind::tf32)
  static constexpr bool kTf32 = true;
  static constexpr CUtensorMapDataType kTma = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
};

struct GemmParams {
  int M, N, K;
  void* D;
  long long ldd;
  const float* bias;
  const void* mask;            // same element type as A/B
  long long ldmask;
  int relu;
  int d_fp32;
  int num_m_blocks, num_n_blocks;
  int group_m;                 // rasterisation: tiles are visited in bands of group_m row-blocks
  // fused column sums of the stored result (db of the layer below = sum over rows of its dY):
  // every 128-row strip writes its partial sums to cs_partials[strip][cs_ld]; the last epilogue warp
  // to finish a column block (ticket) adds the strips in strip order - a fixed summation order, so
  // the result is bit-reproducible - and writes cs_out.  Tickets start at zero and are left at zero.
  float* cs_partials;
  unsigned int* cs_tickets;
  void* cs_out;
  int cs_out_bf16;
  int cs_strips;               // ceil(M / 128)
  int cs_ld;                   // columns per partial row (N rounded up to the tile width)
  // tail wave of the CTA-pair kernel: tiles [0, full_tiles) are computed whole, the rest as two
  // independent half-width tiles each (gemm_tcgen05_2cta.cu)
  int full_tiles;
};

// Tile index -> (row block, column block).  Consecutive tiles (= the CTAs of one wave) walk down a
// band of `group_m` row blocks before moving to the next column block, so a wave touches a compact
// patch of the output and both operand panels stay L2-resident.
__device__ __forceinline__ void tile_coords(const GemmParams& p, int t, int& m_blk, int& n_blk) {
  const int per_group = p.group_m * p.num_n_blocks;
  const int g = t / per_group;
  const int r = t - g * per_group;
  const int m_first = g * p.group_m;
  const int gsize = min(p.group_m, p.num_m_blocks - m_first);
  n_blk = r / gsize;
  m_blk = m_first + (r - n_blk * gsize);
}

// UMMA shared-memory descriptor, SWIZZLE_128B (layout type 2), descriptor version 1 (sm_100).
// K-major : rows of 128 B along K; 8-row groups 1024 B apart (SBO).
// MN-major: k-rows of 128 B along M/N; 8-k groups 1024 B apart (SBO); the next 128 B along M/N live
//           in the next TMA box, BK*128 B further (LBO).
template <int BK>
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, bool mn_major) {
  const uint64_t lbo = mn_major ? uint64_t((BK * 128) >> 4) : 1ull;
  const uint64_t sbo = 1024 >> 4;
  return uint64_t((smem_addr & 0x3FFFF) >> 4) | (lbo << 16) | (sbo << 32) | (1ull << 46) | (2ull << 61);
}

// c_format F32 (bits 4-5 = 1), a/b format (bits 7-9 / 10-12), majors (15/16), N>>3 (17-22), M>>4 (24-28)
template <typename T>
__host__ __device__ constexpr uint32_t make_idesc(int umma_m, int umma_n, bool a_mn, bool b_mn) {
  return (1u << 4) | (Elem<T>::kIdescFormat << 7) | (Elem<T>::kIdescFormat << 10) | (uint32_t(a_mn) << 15) |
         (uint32_t(b_mn) << 16) | (uint32_t(umma_n >> 3) << 17) | (uint32_t(umma_m >> 4) << 24);
}

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ void epi_bar_sync() {
  asm volatile("bar.sync %0, %1;" ::"n"(kEpiBarrierId), "n"(kEpiThreads) : "memory");
}

// ---- epilogue ----------------------------------------------------------------------------------
// Lane l of a warp holds row l of the warp's 32-row quadrant, 32 consecutive columns in f[].
// Returns in lane l the sum over the warp's 32 rows of column l: a butterfly that halves the number
// of live columns per step (16+8+4+2+1 = 31 shuffles instead of 32 five-step reductions).
__device__ __forceinline__ float warp_transpose_colsum(float (&f)[32], int lane) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const bool upper = (lane & o) != 0;
#pragma unroll
    for (int i = 0; i < o; ++i) {
      const float send = upper ? f[i] : f[i + o];
      const float keep = upper ? f[i + o] : f[i];
      f[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
    }
  }
  return f[0];
}

template <typename T>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, const uint32_t (&v)[32], int row, bool row_ok,
                                               int col0, int lane, float* cs_row /* this warp's [BN] or null */,
                                               int cs_col) {
  if (col0 >= p.N) {                       // warp-uniform: the whole chunk is outside the matrix
    if (cs_row != nullptr) cs_row[cs_col + lane] = 0.0f;
    return;
  }
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
  if (p.mask != nullptr && row_ok) {
    if constexpr (Elem<T>::kTf32) {
      const float* mrow = reinterpret_cast<const float*>(p.mask) + (long long)row * p.ldmask + col0;
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        if (col0 + j < p.N) {
          const float4 m4 = __ldg(reinterpret_cast<const float4*>(mrow + j));
          if (!(m4.x > 0.0f)) f[j] = 0.0f;
          if (!(m4.y > 0.0f)) f[j + 1] = 0.0f;
          if (!(m4.z > 0.0f)) f[j + 2] = 0.0f;
          if (!(m4.w > 0.0f)) f[j + 3] = 0.0f;
        }
      }
    } else {
      const __nv_bfloat16* mrow = reinterpret_cast<const __nv_bfloat16*>(p.mask) + (long long)row * p.ldmask + col0;
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
  }
  if (row_ok) {
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
  if (cs_row != nullptr) {
    if (!row_ok) {
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = 0.0f;     // rows beyond M hold bias/garbage, not data
    }
    cs_row[cs_col + lane] = warp_transpose_colsum(f, lane);
  }
}

// The part of the column-sum fusion that runs once per tile, after the TMEM stage was released.  One
// CTA barrier (all four quadrants are in shared memory), then every epilogue warp works alone: it adds
// the four quadrants of ITS quarter of the columns (fixed order), publishes that piece of the strip's
// partial row, and takes a ticket (release: fence + atomic by one lane).  The warp that draws the last
// ticket of a column block (4 per 128-row strip) adds the strips in strip order - a fixed summation
// order, so the result is bit-reproducible - and writes the bias gradient.  No warp waits for another
// warp's fence or atomic; the shared buffer is double-buffered by tile parity, so the barrier of the
// next tile is the only thing that orders its reuse.
template <int BN>
__device__ __forceinline__ void colsum_finish(const GemmParams& p, const float* cs_buf /* [4][BN] */, int strip,
                                              bool strip_ok, int n_blk, int n0, int quad, int lane) {
  epi_bar_sync();
  constexpr int CPW = BN / 4;                      // columns per warp
  if (strip_ok) {
#pragma unroll
    for (int j = lane; j < CPW; j += 32) {
      const int c = quad * CPW + j;
      const float s = ((cs_buf[c] + cs_buf[BN + c]) + cs_buf[2 * BN + c]) + cs_buf[3 * BN + c];
      __stcg(p.cs_partials + (long long)strip * p.cs_ld + n0 + c, s);
    }
  }
  __syncwarp();
  int last = 0;
  if (lane == 0 && strip_ok) {
    __threadfence();                               // cumulative: covers the warp's stores ordered by __syncwarp
    last = atomicAdd(p.cs_tickets + n_blk, 1u) == unsigned(4 * p.cs_strips - 1);
  }
  last = __shfl_sync(0xffffffffu, last, 0);
  if (!last) return;
  __threadfence();
  constexpr int V = BN / 128;                      // float4 vectors per lane: lane owns 4*V consecutive columns
  float4 acc[V];
#pragma unroll
  for (int v = 0; v < V; ++v) acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
  const float* base = p.cs_partials + n0 + lane * 4 * V;
  constexpr int U = 8;                             // strips in flight per round
  for (int st = 0; st < p.cs_strips; st += U) {
    float4 t[U][V];
#pragma unroll
    for (int u = 0; u < U; ++u) {
#pragma unroll
      for (int v = 0; v < V; ++v) {
        t[u][v] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (st + u < p.cs_strips)
          t[u][v] = __ldcg(reinterpret_cast<const float4*>(base + (long long)(st + u) * p.cs_ld) + v);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
#pragma unroll
      for (int v = 0; v < V; ++v) {
        acc[v].x += t[u][v].x; acc[v].y += t[u][v].y; acc[v].z += t[u][v].z; acc[v].w += t[u][v].w;
      }
    }
  }
#pragma unroll
  for (int v = 0; v < V; ++v) {
    const int col = n0 + lane * 4 * V + v * 4;
    const float o[4] = {acc[v].x, acc[v].y, acc[v].z, acc[v].w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (col + e < p.N) {
        if (p.cs_out_bf16) reinterpret_cast<__nv_bfloat16*>(p.cs_out)[col + e] = __float2bfloat16_rn(o[e]);
        else reinterpret_cast<float*>(p.cs_out)[col + e] = o[e];
      }
    }
  }
  if (lane == 0) p.cs_tickets[n_blk] = 0;          // reusable by the next launch on the stream
}

// One (128-row strip x BN columns) tile of one CTA.  `taddr`: TMEM address of this warp's lane
// quadrant in the accumulator stage.  `release()` is called once every tcgen05.ld of the tile has
// completed (the MMA warp may then overwrite the stage).
template <typename T, int BN, typename Release>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, uint32_t taddr, int m0, int n0, int n_blk,
                                              int strip, int quad, int lane, float* cs_smem, int parity,
                                              Release release) {
  const int row = m0 + quad * 32 + lane;
  const bool row_ok = row < p.M;
  const bool cs = p.cs_out != nullptr;
  float* cs_buf = cs_smem + parity * (4 * BN);     // [4][BN] of this tile (double-buffered by tile parity)
  float* cs_row = cs ? cs_buf + quad * BN : nullptr;
  constexpr int NC = BN / 32;
  // One register set, load -> wait -> process.  (A two-set ping-pong that kept the next tcgen05.ld in
  // flight was measured and dropped: it cost 50 more registers per thread, did nothing for a GEMM that
  // runs alone - the epilogue hides under the next tile's main loop either way - and made the step with
  // a co-running gradient sync 6 % slower.)
  uint32_t va[32];
#pragma unroll 1
  for (int c = 0; c < NC; ++c) {
    tmem_ld_32x32b_x32(taddr + c * 32, va);
    tmem_ld_wait();
    epilogue_chunk<T>(p, va, row, row_ok, n0 + c * 32, lane, cs_row, c * 32);
  }
  tc_fence_before();
  __syncwarp();
  release();
  if (cs) colsum_finish<BN>(p, cs_buf, strip, m0 < p.M, n_blk, n0, quad, lane);
}

// ---- host side ---------------------------------------------------------------------------------
// 2-D tensor map over a row-major [rows, cols] matrix with leading dimension ld (elements),
// box = [box_rows, 128 bytes], 128B swizzle, zero fill out of bounds.  Maps are cached per
// (pointer, shape, box): a training step re-launches the same GEMMs on the same buffers.
int get_tensor_map(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, const void* base, long long rows,
                   long long cols, long long ld, int box_rows);

// per-device one-time cudaFuncSetAttribute(MaxDynamicSharedMemorySize)
template <typename Kern>
int ensure_smem_attr(Kern kern, int bytes, bool (&done)[64]) {
  int dev = 0;
  cudaGetDevice(&dev);
  bool& d = done[dev & 63];
  if (!d) {
    B200DP_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    d = true;
  }
  return B200DP_OK;
}

void fill_gemm_params(const b200dp_gemm_args* a, GemmParams* p);

}  // namespace b200dp
