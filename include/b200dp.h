/* libb200dp - C ABI of the hand-written sm_100a kernels and the native runtime of b200dp.
 *
 * Conventions
 *   - every function returns 0 (B200DP_OK) on success, non-zero on failure; the text of the last
 *     failure on the calling thread is returned by b200dp_last_error().
 *   - pointers are device pointers unless the name says "host"; `stream` is a cudaStream_t passed
 *     as void* (NULL = legacy default stream).  Nothing here synchronises the host unless stated.
 *   - matrices are row-major; `ld*` are leading dimensions in ELEMENTS.
 *   - bf16 = __nv_bfloat16 (2 bytes), fp32 = float.
 *
 * Each declaration cites the part of the reference (ref.py = dataParallelTraining_NN_MPI.py) whose
 * work it performs.
 */
#ifndef B200DP_H_
#define B200DP_H_

#ifdef __cplusplus
extern "C" {
#endif

#define B200DP_OK 0
#define B200DP_ERR 1
#define B200DP_MAX_RANKS 8
#define B200DP_MAX_LAYERS 64

const char* b200dp_last_error(void);
int b200dp_version(void);

/* ------------------------------------------------------------------------------------------------
 * Linear-layer GEMM on tcgen05 / TMEM / TMA:  D[M,N] = epilogue(A[M,K] * B[N,K]^T), bf16 inputs,
 * fp32 accumulation.  Covers aten::addmm of the forward (ref.py:42-44,170) and both mm's of
 * AddmmBackward0 (ref.py:176):
 *   forward  : A = X  (K-major, lda=K)    B = W  (K-major,  [N,K], ldb=K)   bias, relu
 *   dgrad    : A = dY (K-major)           B = W  (MN-major: stored [K,N], ldb=N)   mask_src = X
 *   wgrad    : A = dY (MN-major: stored [K,M], lda=M)   B = X (MN-major: stored [K,N], ldb=N)
 * Epilogue order: + bias[n] -> relu -> zero where mask_src[m,n] <= 0 -> store (bf16 or fp32).
 * Requirements: pointers 16-byte aligned, lda/ldb/ldd/ldmask multiples of 8 elements (4 for an
 * fp32 D), N % 8 == 0.  M/N/K tails are handled (TMA zero fill, guarded stores).
 */
typedef struct b200dp_gemm_args {
  int M, N, K;
  const void* A;        /* bf16 */
  long long lda;
  int a_mn_major;       /* 0: A stored [M,K];  1: A stored [K,M] */
  const void* B;        /* bf16 */
  long long ldb;
  int b_mn_major;       /* 0: B stored [N,K];  1: B stored [K,N] */
  void* D;              /* bf16 or fp32 [M,N] */
  long long ldd;
  int d_fp32;
  const float* bias;    /* fp32 [N] or NULL */
  int relu;
  const void* mask_src; /* bf16 [M,N] or NULL */
  long long ldmask;
  int tile_n;           /* 0 = auto, 128 or 256 */
  int max_ctas;         /* 0 = one persistent CTA per SM; >0 caps the grid (SM budget when overlapping) */
  int cta_group;        /* 0 = auto; 1 = one CTA per 128x256 tile; 2 = CTA pair (cta_group::2) per 256x256 tile */
  int split_tail;       /* CTA-pair kernel: 0 = auto (a last wave that would keep <= half of the SM pairs busy runs
                           as half-width tiles), -1 = never */
  /* Fused bias gradient (the `sum` of AddmmBackward0, ref.py:176): when colsum_out is set the
   * epilogue also writes colsum_out[n] = sum_m D[m,n] (the values it stores, before rounding), i.e.
   * the db of the layer whose dY this GEMM produces.  Fixed summation order (bit-reproducible).
   * Needs a K-major A (a_mn_major = 0); the CTA-pair kernel fuses it into the dgrad layout only
   * (a_mn_major = 0, b_mn_major = 1) - other layouts take the one-CTA kernel (cta_group = 1).
   * colsum_scratch: b200dp_gemm_colsum_scratch_bytes(M, N) bytes whose first 4096 bytes are zero
   * before the first use (ticket counters; every launch leaves them zero). */
  void* colsum_out;     /* bf16 or fp32 [N], or NULL */
  int colsum_out_bf16;
  void* colsum_scratch;
  long long colsum_scratch_bytes;
} b200dp_gemm_args;

int b200dp_gemm_bf16(const b200dp_gemm_args* args, void* stream);
long long b200dp_gemm_colsum_scratch_bytes(int M, int N);
/* Registers that one persistent GEMM CTA and one fused-sync CTA occupy on an SM.  The step schedule runs
 * the sync kernels beside backward's GEMMs, so gemm + sync must fit the SM's 65536 registers. */
int b200dp_overlap_register_budget(int* gemm_regs_per_cta, int* sync_regs_per_cta);
/* Same contract with fp32 A, B and mask_src and an fp32 D: the tensor cores read the operands as TF32
 * (kind::tf32) and accumulate in fp32.  The reference computes in fp32 only (ref.py:159); this is the
 * path of fp32 models beyond the fused SIMT kernel.  lda/ldb/ldmask multiples of 4 elements.
 * K-major operands only (a_mn_major = b_mn_major = 0): on B200 a kind::tf32 MMA with a transposed
 * operand is a no-op, so backward passes transposed copies made with b200dp_transpose_f32. */
int b200dp_gemm_tf32(const b200dp_gemm_args* args, void* stream);
/* dst[c, r] = src[r, c] for an fp32 [rows, cols] matrix (leading dimensions in elements). */
int b200dp_transpose_f32(const float* src, long long lds, int rows, int cols, float* dst, long long ldd,
                         void* stream);

/* ------------------------------------------------------------------------------------------------
 * Whole-network SIMT kernel for MLPs too small for tensor cores (the reference's own 2-3-1 net,
 * ref.py:41-45): forward + MSE + backward + batch reduction of the gradients in one launch
 * (everything ref.py:170-176 does).  fp32.  `flat_numel` floats of `params` / `grads` use the
 * offsets below (weights row-major [out,in]).  scratch: b200dp_tiny_mlp_scratch_bytes(); its first
 * 4 bytes (a ticket counter) must be zero before the first use and are left zero by every launch.
 */
typedef struct b200dp_mlp_desc {
  int num_layers;
  int sizes[B200DP_MAX_LAYERS + 1];
  long long w_off[B200DP_MAX_LAYERS];
  long long b_off[B200DP_MAX_LAYERS];
  long long flat_numel;
} b200dp_mlp_desc;

long long b200dp_tiny_mlp_scratch_bytes(const b200dp_mlp_desc* m, int rows);
int b200dp_tiny_mlp_fwd_bwd(const b200dp_mlp_desc* m, const float* params, const float* x, const float* y,
                            int rows, float* grads, float* loss, void* scratch, long long scratch_bytes,
                            void* stream);

/* Allocation-free variant of b200dp_tiny_mlp_fwd_bwd (no counterpart in the reference: the reference
 * never manages device memory at all).  For callers that cannot call cudaMalloc where they train -
 * a latency-critical loop, a plugin inside a framework whose pool owns the whole GPU, a process on
 * a fixed memory budget.  The caller passes inputs and outputs only; the cross-block scratch is a
 * built-in device array of B200DP_TINY_NOALLOC_SCRATCH_BYTES that every GPU using the library holds
 * until exit.  Problems that need more scratch (256 + blocks*(flat_numel+1)*4 bytes, blocks <= SM
 * count, 0 when one block suffices) are refused with an error.  The array is shared by all calls:
 * calls that may overlap on one GPU race - serialise them.  The first call on each device looks up
 * the array's device address (cudaGetSymbolAddress); make that first call before any stream
 * capture.  Calls no cudaMalloc/cudaFree. */
#define B200DP_TINY_NOALLOC_SCRATCH_BYTES (1 << 20)
int b200dp_tiny_mlp_fwd_bwd_noalloc(const b200dp_mlp_desc* m, const float* params, const float* x,
                                    const float* y, int rows, float* grads, float* loss, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused MSE loss + gradient (nn.MSELoss() + MseLossBackward, ref.py:94,173,176):
 *   *loss = mean((pred-target)^2),  dy = grad_scale * 2 (pred-target) / numel.
 * denom > 0 replaces numel as the divisor (outputs padded with zero predictions against zero targets
 * add nothing to the sum but must not enter the mean).
 * grad_scale is 1 for plain training; a rank whose gradient enters the all-reduce with weight w_r
 * (sample weighting of uneven shards) passes w_r here, so the gradients arrive pre-weighted and any
 * reduction - including the NVSwitch's unweighted sum - yields the weighted mean.
 * Deterministic two-stage reduction through `scratch` (b200dp_mse_scratch_bytes(numel) bytes; the
 * first 4 bytes are a ticket counter that must start at zero and is left at zero).
 */
long long b200dp_mse_scratch_bytes(long long numel);
int b200dp_mse_loss_grad(const void* pred, int pred_bf16, const float* target, long long numel, long long denom,
                         void* dy, int dy_bf16, float* loss, float grad_scale, void* scratch,
                         long long scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Regression head (last Linear with 1..4 outputs, ref.py:44): GEMV + bias + MSE + dY in one kernel,
 * and its backward in one pass over H:  dH = (dY W) * [H > 0] (pass NULL for the first layer; H, W, dH are
 * bf16, or fp32 when elems_fp32 is set - fp32 training),
 * dW = dY^T H, db = sum dY (fp32 or bf16, fixed-order reduction through scratch).
 * scratch: b200dp_head_scratch_bytes(rows, K, nout); first 4 bytes zero before first use.
 */
long long b200dp_head_scratch_bytes(int rows, int K, int nout);
int b200dp_head_fwd_mse(const void* H, long long ldh, const void* W, int elems_fp32, const float* bias, const float* target,
                        int rows, int K, int nout, float* yhat, float* dy, float* loss, float grad_scale,
                        void* scratch, long long scratch_bytes, void* stream);
/* db_prev (NULL = skip; needs dH): also writes the column sums of dH, i.e. the bias gradient of the
 * layer below (fused, fixed order) - K values in the gradient dtype. */
int b200dp_head_bwd(const float* dy, const void* W, const void* H, long long ldh, int elems_fp32, int rows, int K, int nout,
                    void* dH, long long lddh, void* dW, void* db, void* db_prev, int grads_bf16, void* scratch,
                    long long scratch_bytes, void* stream);

/* Bias gradient of a GEMM layer: out[n] = sum_rows A[r,n]  (the `sum` of AddmmBackward0, ref.py:176). */
long long b200dp_colsum_scratch_bytes(int rows, int n);
int b200dp_colsum(const void* A, int a_fp32, long long lda, int rows, int n, void* out, int out_bf16, void* scratch,
                  long long scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Local optimizer step, torch.optim.SGD(momentum) semantics (ref.py:91,211):
 *   g = grads*grad_scale;  buf = first_step ? g : mu*buf + g;  p -= lr*buf;  shadow = bf16(p).
 */
int b200dp_sgd_momentum(float* params, const void* grads, int grads_bf16, float* momentum, void* shadow_bf16,
                        long long n, float lr, float mu, float grad_scale, int first_step, void* stream);
/* Same with a cap on the grid (0 = none): keeps an update that overlaps GEMMs below the full HBM rate. */
int b200dp_sgd_momentum_ex(float* params, const void* grads, int grads_bf16, float* momentum, void* shadow_bf16,
                           long long n, float lr, float mu, float grad_scale, int first_step, int max_blocks,
                           void* stream);
int b200dp_cast_f32_bf16(const float* src, void* dst, long long n, void* stream);
/* Mini-batch assembly on the device: dst[r,:] = src[idx[r],:], rows of `row_bytes` bytes (multiple of
 * 16; idx is int64 on the device).  Stands in for the DataLoader's shuffle + collate when the batch
 * is smaller than the shard (ref.py:146,155). */
int b200dp_gather_rows(const void* src, long long src_row_bytes, const long long* idx, int rows,
                       long long row_bytes, void* dst, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Pinned host staging memory for input batches (the reference's DataLoader tensors never leave the
 * CPU, ref.py:146,155; here every step's inputs cross PCIe).  2 MB-aligned anonymous memory advised to
 * huge pages, touched, and registered with cudaHostRegister: measured 55.2-55.5 GB/s H2D on every
 * allocation, where cudaHostAlloc / pin_memory() buffers vary between 45 and 55 GB/s.
 */
int b200dp_host_alloc(long long bytes, void** ptr, long long* mapped_bytes);
int b200dp_host_free(void* ptr, long long mapped_bytes);

/* ------------------------------------------------------------------------------------------------
 * Fused all-reduce + weighting + SGD-momentum over NVLink peer memory - the reference's
 * gather / root average / send / optimizer.step sequence (ref.py:185-211) as ONE kernel per bucket.
 *   g = (sum_r weights[r]*grads_r) / div  (rank order);  buf = first ? g : mu*buf+g;  p -= lr*buf
 * mode 0 one-shot (every rank reduces everything), 1 two-shot (rank r owns slice r and pushes the
 * updated parameters to all peers), 2 NVLS (multimem.ld_reduce / multimem.st; needs mc_* addresses;
 * the switch cannot weight its inputs, so all weights[r] must be equal - pre-scale the gradients).
 * A peer that does not arrive within timeout_ms sets error_flag and the launch changes nothing.
 * Pointer tables hold every rank's buffer as mapped into THIS process (index `rank` = own buffer).
 * `flags[r]`: rank r's signal pad, B200DP_SYNC_MAX_BLOCKS*B200DP_MAX_RANKS u32, zero-initialised
 * once.  `epochs`: local, B200DP_SYNC_MAX_BLOCKS u32, zero-initialised once.  `error_flag`: local
 * u32, set to 1+peer when a peer does not arrive within timeout_ms.
 * All ranks must issue the same sequence of calls with the same n / mode / max_blocks.
 */
#define B200DP_SYNC_MAX_BLOCKS 160

typedef struct b200dp_sync_args {
  int rank, world;
  const void* grads[B200DP_MAX_RANKS];   /* fp32 or bf16 */
  float* params[B200DP_MAX_RANKS];       /* fp32 master */
  void* shadow[B200DP_MAX_RANKS];        /* bf16 shadow weights, or all NULL */
  unsigned int* flags[B200DP_MAX_RANKS];
  float* momentum;                       /* local */
  unsigned int* epochs;                  /* local */
  unsigned int* error_flag;              /* local */
  float weights[B200DP_MAX_RANKS];
  float div;
  float lr, mu;
  int first_step;
  int grads_bf16;
  long long offset;                      /* first element of the bucket inside the flat buffers */
  long long n;                           /* bucket length in elements (multiple of 8) */
  int mode;
  int push_master;                       /* modes 1/2: also write fp32 params into the peers */
  long long replicate_tail;              /* modes 1/2 without push_master: the last `replicate_tail` elements of the
                                            bucket (a layer's biases, which kernels read in fp32) still get their
                                            fp32 value written into every peer; 0 = none */
  int max_blocks;                        /* 0 = one CTA per SM; SM budget of the comm kernel */
  int timeout_ms;                        /* 0 = wait forever */
  const void* mc_grads;                  /* NVLS multicast addresses (mode 2) */
  void* mc_shadow;
  float* mc_params;
} b200dp_sync_args;

int b200dp_allreduce_sgd(const b200dp_sync_args* args, void* stream);
/* One-GPU emulation of a whole world in a single cooperative launch (tests; single-GPU boxes). */
long long b200dp_sync_args_device_bytes(void);
int b200dp_allreduce_sgd_emulated(const b200dp_sync_args* per_rank, int world, void* device_args, void* stream);
/* Cross-GPU barrier on the stream (blocking semantics of the MPI calls at ref.py:185,203). */
int b200dp_peer_barrier(const b200dp_sync_args* args, void* stream);
/* Every rank copies `bytes` from args->params[root] into args->params[rank] (ref.py:87-88). */
int b200dp_broadcast(const b200dp_sync_args* args, int root, long long bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Symmetric (peer-mapped) memory.  Replaces MPI.COMM_WORLD as the transport (ref.py:61-63): each
 * rank allocates, exports a handle, peers import it and get a pointer valid in their own address
 * space.  Two flavours:
 *   ipc : cudaMalloc + cudaIpcMemHandle (64 opaque bytes, exchanged by the host-side store).
 *   vmm : cuMemCreate + POSIX fd; peers duplicate the fd with pidfd_getfd().  Needed for NVLS:
 *         b200dp_mc_* bind the same allocations into a multicast object.
 */
int b200dp_ipc_alloc(long long bytes, void** ptr, unsigned char* handle64);
int b200dp_ipc_open(const unsigned char* handle64, void** ptr);
int b200dp_ipc_close(void* ptr);
int b200dp_ipc_free(void* ptr);

int b200dp_vmm_supported(int* fd_export, int* multicast);
int b200dp_vmm_alloc(long long bytes, void** ptr, int* fd, long long* padded_bytes);
int b200dp_vmm_import(int owner_pid, int owner_fd, long long padded_bytes, void** ptr);
int b200dp_vmm_free(void* ptr, long long padded_bytes);
int b200dp_mc_create(int world, long long padded_bytes, int* fd, unsigned long long* mc_handle);
int b200dp_mc_import(int owner_pid, int owner_fd, unsigned long long* mc_handle);
int b200dp_mc_add_device(unsigned long long mc_handle);
int b200dp_mc_bind(unsigned long long mc_handle, void* local_ptr, long long padded_bytes, void** mc_ptr);
int b200dp_mc_release(unsigned long long mc_handle, void* mc_ptr, long long padded_bytes);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H_ */
