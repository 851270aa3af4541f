/* C ABI of libpdjob_b200.so - the sm_100a data plane of the B200 PaddleJob
 * runtime.  Every entry point returns 0 on success and a non-zero code on
 * failure; pdj_last_error() then returns a thread-local description.
 *
 * All kernels are launched on the cudaStream_t passed in and never
 * synchronise the device, so they can be captured in CUDA graphs.
 *
 * Reference parity notes cite files of lfeng-nl/paddle-operator, which has no
 * device code: it only publishes the rank/endpoint environment
 * (controllers/paddlejob_helper.go:215-377) that Paddle's own collectives and
 * parameter servers bootstrap from.  The entry points here are the hot paths
 * of the workloads that environment launches.
 */
#ifndef PDJOB_B200_H_
#define PDJOB_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

struct PdjComm;      /* see csrc/common.cuh; built by the Python side */
struct PdjArena;
struct PdjMulticast;

#define PDJ_AR_TWOSHOT 0
#define PDJ_AR_ONESHOT 1
#define PDJ_AR_NVLS 2

#define PDJ_BN_MAX_ROW_BLOCKS 512

const char* pdj_last_error(void);

/* ---- symmetric-memory runtime (csrc/comm_runtime.cu) ----------------------
 * Stands in for the ConfigMap rendezvous of the reference
 * (controllers/paddlejob_controller.go:289-306): besides PADDLE_* strings the
 * ranks exchange one file descriptor each, naming their arena. */
int pdj_device_count(void);
int pdj_can_access_peer(int dev, int peer);
int pdj_multicast_supported(int device);
int pdj_arena_round_size(int device, size_t bytes, int nranks, size_t* out);
int pdj_arena_create(int device, size_t bytes, struct PdjArena** out);
int pdj_arena_export_fd(struct PdjArena* a, int* fd);
int pdj_arena_import_fd(int device, int fd, size_t bytes, struct PdjArena** out);
void* pdj_arena_ptr(struct PdjArena* a);
size_t pdj_arena_bytes(struct PdjArena* a);
int pdj_arena_destroy(struct PdjArena* a);
int pdj_multicast_create(int device, int nranks, size_t bytes, struct PdjMulticast** out, int* fd);
int pdj_multicast_import(int device, int fd, size_t bytes, struct PdjMulticast** out);
int pdj_multicast_add_device(struct PdjMulticast* m);
int pdj_multicast_bind(struct PdjMulticast* m, struct PdjArena* local, void** mc_ptr);
int pdj_multicast_destroy(struct PdjMulticast* m);

/* ---- collectives fused with scale / cast (csrc/collectives.cu) -------------
 * Collective-mode wiring in the reference: controllers/paddlejob_helper.go:
 * 261-265,319-322 (PADDLE_TRAINER_ENDPOINTS, PADDLE_TRAINERS_NUM,
 * PADDLE_TRAINER_ID).  Offsets are bytes from the arena base and identical on
 * every rank.  emulate != 0 runs all ranks as one cooperative launch over
 * arenas that live on a single device (tests). */
int pdj_barrier(const struct PdjComm* c, int emulate, cudaStream_t st);
int pdj_allreduce_bf16(const struct PdjComm* c, size_t in_off, size_t out_off, size_t n_elems, float scale,
                       int algo, int out_f32, int nblocks, int emulate, cudaStream_t st);
/* Gradient reduce-scatter + SGD-momentum + parameter broadcast in one kernel (K1b).  The four flat buffers - bf16
 * gradients and parameters, fp32 masters and momentum - live in the arena at the given byte offsets; the call covers
 * the elements [first_elem, first_elem + n_elems).  Each rank reduces its 1/N slice of every peer's gradients (P2P
 * loads, or the switch's multimem.ld_reduce with PDJ_AR_NVLS), updates its master / momentum slice
 * (hyper = {lr, momentum, weight_decay, grad_scale} in device memory; weight decay on elements < wd_split) and
 * stores the new master and bf16 parameter into every rank's copy.  After it every rank holds identical
 * parameters and masters; only the owner's momentum slice is current. */
int pdj_allreduce_sgd(const struct PdjComm* c, size_t grad_off, size_t param_off, size_t master_off, size_t mom_off,
                      size_t first_elem, size_t n_elems, size_t wd_split, const float* hyper, float scale, int algo,
                      int nblocks, int emulate, cudaStream_t st);
int pdj_alltoall(const struct PdjComm* c, size_t in_off, size_t out_off, size_t chunk_bytes, int nblocks,
                 int emulate, cudaStream_t st);
int pdj_broadcast(const struct PdjComm* c, int root, size_t off, size_t bytes, int nblocks, int emulate,
                  cudaStream_t st);

/* ---- PS mode: GPU-sharded embedding tables (csrc/embedding.cu) -------------
 * The PSERVER / TRAINER roles of the reference (api/v1/paddlejob_types.go:44-48,
 * PADDLE_PSERVERS_IP_PORT_LIST at controllers/paddlejob_helper.go:258-260;
 * pull / push semantics in docs/design-arch.md:49-74).  The table is cut into P shards (list_cap, the job's
 * spec.ps.replicas; 0 = one per rank): row `id` belongs to shard id % P, which lives on rank (id % P) % nranks in
 * slot (id % P) / nranks, at row id / P of the slot.  All offsets are arena offsets.
 *   lookup: out_deep[b, s, :] = bf16(table[id[b,s]]), out_wide[b] = sum_s wide[id[b,s]]
 *   push  : sync  - staging[id] += grad (peer red.add), row registered with its owner
 *           async - table[id]  -= lr * grad immediately (asynchronous SGD)
 *   apply : owner applies SGD (0) or Adagrad (1) to its touched rows, clears staging */
struct PdjEmbTable {
  size_t deep_off;      /* fp32 [rows_per_shard, dim] */
  size_t wide_off;      /* fp32 [rows_per_shard] (has_wide) */
  size_t gdeep_off;     /* staging, same shapes */
  size_t gwide_off;
  size_t acc_deep_off;  /* Adagrad accumulators, same shapes */
  size_t acc_wide_off;
  size_t flag_off;      /* u8 [rows_per_shard rounded up to 128]: row touched this step */
  size_t list_off;      /* 0: separate arrays; S > 0: one S-byte record per row holds all six parts (see csrc/embedding.cu) */
  size_t count_off;     /* u32 [2]: touched rows this step, last step */
  long long rows_per_shard;
  unsigned int list_cap; /* number of table shards P (the job's pserver count); 0 = one per rank */
  int dim;              /* power of two, 4..128 */
  int has_wide;
  size_t shadow_off;    /* record layout only, 0 = none: bf16 copy of the deep weights at the head of the record, directly
                           followed by the wide weight (wide_off = shadow_off + 2 * dim) - the 64 bytes a lookup reads.
                           pdj_emb_apply keeps it equal to bf16(weights); asynchronous SGD cannot (use 0). */
};
int pdj_emb_lookup(const struct PdjComm* c, const struct PdjEmbTable* t, size_t ids_off, size_t out_deep_off,
                   size_t out_wide_off, int B, int S, int barrier, int nblocks, int emulate, cudaStream_t st);
/* Same lookup, writing rows of `out_pitch` bf16 values (a multiple of 8, >= S * dim): the S * dim embedding values,
 * then the sample's `dense_n` dense features (fp32 [B, dense_n] at arena offset dense_off, converted), then zeros -
 * i.e. the padded input of the MLP tower in place, with no concatenation / conversion / fill kernels. */
int pdj_emb_lookup2(const struct PdjComm* c, const struct PdjEmbTable* t, size_t ids_off, size_t out_deep_off,
                    size_t out_wide_off, int B, int S, int barrier, int out_pitch, size_t dense_off, int dense_n,
                    int nblocks, int emulate, cudaStream_t st);
/* Same push, reading the deep gradient from rows of g_pitch bf16 values (the gradient of that padded input). */
int pdj_emb_push2(const struct PdjComm* c, const struct PdjEmbTable* t, size_t ids_off, size_t gdeep_off,
                  size_t gwide_off, int B, int S, int async_sgd, float lr, float grad_scale, int g_pitch, int nblocks,
                  int emulate, cudaStream_t st);
int pdj_emb_push(const struct PdjComm* c, const struct PdjEmbTable* t, size_t ids_off, size_t gdeep_off,
                 size_t gwide_off, int B, int S, int async_sgd, float lr, float grad_scale, int nblocks, int emulate,
                 cudaStream_t st);
int pdj_emb_apply(const struct PdjComm* c, const struct PdjEmbTable* t, int optimizer, float lr, float eps,
                  int nblocks, int emulate, cudaStream_t st);

/* ---- tcgen05 / TMA fused linear layer (csrc/gemm_tcgen05.cu) -----------------
 * out[M,N] = epilogue(A[M,K] . B[N,K]^T): bf16 operands (row pitches lda/ldb in
 * elements, multiples of 8), fp32 accumulation in tensor memory, bf16 output with
 * row pitch ldo.  Epilogue: + bias[N] (fp32, optional), ReLU (optional), zero where
 * mask[M,N] <= 0 (bf16, optional; the ReLU derivative for the backward GEMM).
 * The MLP of the Wide&Deep workload (deploy/examples/wide_and_deep.yaml). */
int pdj_linear_block_n(int N);
int pdj_linear_bf16(const void* a, const void* b, void* out, const float* bias, const void* mask, int M, int N, int K,
                    int lda, int ldb, int ldo, int relu, cudaStream_t st);
/* pdj_linear_bf16 with B given as b_kn[K][N] (row pitch ldb >= N): out = epilogue(a . b_kn), e.g. dX = dZ . W with the
 * weight W[out][in] read as stored. */
int pdj_linear_bf16_kn(const void* a, const void* b_kn, void* out, const float* bias, const void* mask, int M, int N, int K,
                       int lda, int ldb, int ldo, int relu, cudaStream_t st);
/* 1x1 convolution over NHWC activations as a GEMM on the same tcgen05 pipeline, with the
 * per-channel sum / sum of squares of the (bf16-rounded) output accumulated in the epilogue:
 *   out[M,N] = A[M,K] . B[N,K]^T,   partial[p][0][n] = sum_m out, partial[p][1][n] = sum_m out^2
 * for p < pdj_conv1x1_partial_rows(M, N) (fp32; pass partial = NULL to skip the statistics).
 * N must be a multiple of 64.  Feed `partial` to pdj_bn_fwd_pre: the BatchNorm that follows then
 * needs no statistics pass over the activation.  ResNet-50's bottleneck 1x1 convolutions
 * (deploy/examples/resnet.yaml workload). */
int pdj_conv1x1_partial_rows(int M, int N);
int pdj_conv1x1_bf16(const void* a, const void* b, void* out, float* partial, int M, int N, int K, int lda, int ldb,
                     int ldo, cudaStream_t st);
/* out[M,N] = addend[M,N] + A[M,K] . B[N,K]^T on the same kernel (the addend box is fetched by TMA into
 * the epilogue's staging buffer; out may alias addend).  The input-gradient GEMM of a bottleneck's
 * first 1x1 convolution with the skip connection's gradient folded in. */
int pdj_conv1x1_add_bf16(const void* a, const void* b, const void* addend, void* out, int M, int N, int K, int lda,
                         int ldb, int ldc, int ldo, cudaStream_t st);
/* Both of the above with B given as b_kn[K][N] (row pitch ldb >= N elements, a multiple of 8): the weight of a 1x1
 * convolution used transposed for its input gradient, read as stored (MN-major tcgen05 operand).  addend may be NULL. */
int pdj_conv1x1_kn_bf16(const void* a, const void* b_kn, const void* addend, void* out, int M, int N, int K, int lda, int ldb,
                        int ldc, int ldo, cudaStream_t st);
/* General NHWC convolution forward (square filter k <= 7, symmetric padding, stride) as an implicit GEMM on the same
 * pipeline: out[N][P][Q][Cout] from x[N][H][W][Cin] and the channels-last weight w[Cout][k][k][Cin].  The A operand is
 * read with im2col-mode TMA loads (128 consecutive output pixels x 64 channels per filter tap; the TMA unit walks
 * rows and images and zero-fills the padding), so nothing is materialised.  `partial` as above with
 * pdj_conv_im2col_partial_rows(...) rows.  The input gradient of a stride-1 convolution is the same call on dy with
 * the flipped, transposed weight.  Cin % 64 == 0, Cout % 64 == 0.  ResNet-50's 3x3 and strided convolutions. */
int pdj_conv_im2col_partial_rows(int N, int H, int W, int Cout, int ksize, int pad, int stride);
int pdj_conv_im2col_fwd(const void* x, const void* w, void* out, float* partial, int N, int H, int W, int Cin, int Cout,
                        int ksize, int pad, int stride, cudaStream_t st);
/* ResNet stem on the same pipeline: the 7x7 / stride-2 convolution expressed as a 4x4 / stride-1
 * convolution (16 -> 64 channels) over the space-to-depth input x[N][Ho+3][Wo+3][16] (see
 * pdj_preprocess_u8_s2d); w[64][4][4][16] bf16, out[N][Ho][Wo][64] bf16, Wo <= 128.  The A operand is read
 * through a TMA tensor map with overlapping pixel windows: no im2col buffer.  `partial` as above, with
 * pdj_stem_s2d_partial_rows(N, Ho) rows (NULL to skip). */
int pdj_stem_s2d_partial_rows(int N, int Ho);
int pdj_stem_s2d_fwd(const void* x, const void* w, void* out, float* partial, int N, int Ho, int Wo, cudaStream_t st);
/* Weight gradient of the stem: dw[64][4][4][16] bf16 from x and dy[N][Ho][Wo][64]; a persistent tcgen05 kernel
 * (MN-major operands straight from the NHWC tensors, one output row per k-step, X rows reused from a
 * shared-memory ring) plus a fixed-order reduction of the per-CTA partials in `ws`
 * (pdj_stem_s2d_wgrad_ws_bytes() bytes). */
int pdj_stem_s2d_wgrad_ws_bytes(void);
int pdj_stem_s2d_wgrad(const void* x, const void* dy, void* dw, void* ws, int N, int Ho, int Wo, cudaStream_t st);

/* Weight gradient of a 3x3 / pad 1 / stride 1 NHWC convolution on tcgen05 (csrc/wgrad_tcgen05.cu):
 * dw[Cout][3][3][Cin] bf16 (the memory of a channels-last [Cout, Cin, 3, 3] weight) from x[N][H][W][Cin] and
 * dy[N][H][W][Cout].  One shared-memory X tile with its halo serves all nine taps through row-shifted matrix
 * descriptors; accumulators stay in tensor memory for the whole kernel (split-K over pixel blocks), a second
 * kernel adds the per-CTA partials held in `ws` (pdj_conv3x3_wgrad_ws_bytes bytes) in a fixed order.
 * Cin % 32 == 0, Cout % 64 == 0, W <= 254.  ResNet-50's bottleneck 3x3 convolutions (deploy/examples/resnet.yaml). */
size_t pdj_conv3x3_wgrad_ws_bytes(int N, int H, int W, int Cin, int Cout);
int pdj_conv3x3_wgrad(const void* x, const void* dy, void* dw, void* ws, int N, int H, int W, int Cin, int Cout,
                      cudaStream_t st);

/* Weight gradient of a linear layer / 1x1 convolution on the same MN-major tcgen05 pipeline:
 * dw[No][Nc] (bf16, row-major) = dy[M][No]^T . x[M][Nc], row pitches ldy / ldx in elements (multiples of 8), No and Nc
 * multiples of 8; split-K over 128-row blocks with a fixed-order reduce of the per-CTA partials in `ws`
 * (pdj_wgrad_mn_ws_bytes bytes).  The MLP of deploy/examples/wide_and_deep.yaml and ResNet's 1x1 convolutions. */
size_t pdj_wgrad_mn_ws_bytes(int M, int No, int Nc);
int pdj_wgrad_mn(const void* x, const void* dy, void* dw, void* ws, int M, int No, int Nc, int ldx, int ldy,
                 cudaStream_t st);
/* pdj_wgrad_mn that also returns db[No] (bf16) = column sums of dy (the bias gradient), computed from the dY tiles the
 * kernel streams anyway.  `ws`: pdj_wgrad_mn_db_ws_bytes(M, No, Nc) bytes. */
size_t pdj_wgrad_mn_db_ws_bytes(int M, int No, int Nc);
int pdj_wgrad_mn_db(const void* x, const void* dy, void* dw, void* db, void* ws, int M, int No, int Nc, int ldx, int ldy,
                    cudaStream_t st);

/* ---- ResNet-50 step kernels (csrc/bn_act.cu, csrc/optim_misc.cu) -----------
 * The workload of deploy/examples/resnet.yaml:13-19. NHWC bf16 activations
 * viewed as [M, C]; fp32 statistics. */
size_t pdj_bn_workspace_bytes(void);
int pdj_bn_fwd(const void* x, const void* res, void* y, int M, int C, const float* gamma, const float* beta,
               float eps, float momentum, float* running_mean, float* running_var, float* mean_rstd,
               float* scale_shift, int relu, void* workspace, cudaStream_t st);
int pdj_bn_bwd(const void* dy, const void* y, const void* x, void* dx, void* dres, int M, int C,
               const float* gamma, const float* mean_rstd, float* dgamma_dbeta, void* dgb_bf16, float* coef,
               int relu, void* workspace, cudaStream_t st);
/* Variants with a 1-bit ReLU gate: forward also writes one byte per 8 elements (bit j = y_j > 0), backward
 * reads that instead of y - 1/16 of the bytes on the two backward passes. */
int pdj_bn_fwd_gate(const void* x, const void* res, void* y, void* gate, int M, int C, const float* gamma,
                    const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                    float* mean_rstd, float* scale_shift, int relu, void* workspace, cudaStream_t st);
int pdj_bn_bwd_gate(const void* dy, const void* y, const void* gate, const void* x, void* dx, void* dres, int M, int C,
                    const float* gamma, const float* mean_rstd, float* dgamma_dbeta, void* dgb_bf16, float* coef,
                    int relu, void* workspace, cudaStream_t st);
/* pdj_bn_bwd_gate with the two bf16 gradients at separate addresses (dgamma_bf16 / dbeta_bf16, both or neither). */
int pdj_bn_bwd_gate2(const void* dy, const void* y, const void* gate, const void* x, void* dx, void* dres, int M, int C,
                     const float* gamma, const float* mean_rstd, float* dgamma_dbeta, void* dgamma_bf16,
                     void* dbeta_bf16, float* coef, int relu, void* workspace, cudaStream_t st);
/* Forward with the statistics pre-reduced to `nb` partial rows ([nb][2][C] fp32) by the kernel that
 * produced x (pdj_conv1x1_bf16): finalisation + apply only. */
int pdj_bn_fwd_pre(const void* x, const void* res, void* y, void* gate, int M, int C, const float* gamma,
                   const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                   float* mean_rstd, float* scale_shift, int relu, const float* partial, int nb, cudaStream_t st);
/* No counterpart in the reference (it has no device code and no allocator
 * story).  Same result as pdj_bn_fwd, but takes no workspace and allocates
 * nothing: its block partials and tickets live in a file-scope device array of
 * PDJ_BN_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until
 * exit.  For FFI callers without an allocator and for rank loops whose HBM
 * budget is frozen (a late cudaMalloc/cudaFree would fail or synchronise the
 * device under spinning peers).  Rejects C > PDJ_BN_NOALLOC_MAX_C or
 * M > PDJ_BN_NOALLOC_MAX_ROWS.  The first call on a device resolves the array's
 * address: make it before any stream capture.  All calls share the array, so
 * calls that may overlap on the GPU race; serialise them. */
#define PDJ_BN_NOALLOC_SCRATCH_BYTES ((2u << 20) + 256u)
#define PDJ_BN_NOALLOC_MAX_C 2048
#define PDJ_BN_NOALLOC_MAX_ROWS (1 << 24)
int pdj_bn_fwd_noalloc(const void* x, const void* res, void* y, int M, int C, const float* gamma,
                       const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                       float* mean_rstd, float* scale_shift, int relu, cudaStream_t st);
/* NHWC bf16 max pooling with a saved arg-max (csrc/pool.cu).  idx holds one byte per output element. */
int pdj_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int s, int p,
                    cudaStream_t st);
int pdj_maxpool_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int k, int s, int p,
                    cudaStream_t st);
int pdj_sgd_momentum(float* w, float* mom, const void* grad, void* param, size_t n, size_t wd_split,
                     const float* hyper, cudaStream_t st);
int pdj_preprocess_u8(const void* in, void* out, size_t n_px, const float* mean3, const float* std3,
                      cudaStream_t st);
/* uint8 NHWC image -> normalised bf16 in the stem's space-to-depth layout [N, H/2+3, W/2+3, 16] (border untouched). */
int pdj_preprocess_u8_s2d(const void* in, void* out, int N, int H, int W, const float* mean3, const float* std3,
                          cudaStream_t st);
/* out[c] = sum_r x[r, c] for row-major bf16 x with row pitch ld (the bias gradient of a linear layer). */
int pdj_colsum_bf16(const void* x, float* out, int M, int N, int ld, cudaStream_t st);
/* Deterministic weighted column sums in one launch: res[c] = sum_r w[r] x[r,c] (x bf16 [M,N], N % 8 == 0),
 * res[N+j] = sum_r w[r] xe[r,j] (xe fp32 [M,Ne], optional), res[N+Ne] = sum_r w[r] when `ones`;
 * w[r] = wgt[r] * gscale[0], either may be NULL (= 1).  Result columns [dst_start[k], +dst_n[k]) go to dst[k] as
 * bf16 (dst_bf16[k] != 0) or fp32, k < ndst <= 4.  `ws`: pdj_colsum_w_ws_bytes() bytes, first 4 zero before the
 * first use (left zero by every call).  Bias gradients; every parameter gradient of the Wide&Deep head. */
int pdj_colsum_w_ws_bytes(int M, int N, int Ne, int ones);
int pdj_colsum_w(const void* x, int ld, int M, int N, const float* wgt, const float* gscale, const float* xe, int lde,
                 int Ne, int ones, void* ws, void* const* dst, const int* dst_start, const int* dst_n,
                 const int* dst_bf16, int ndst, cudaStream_t st);
/* Wide&Deep head in one pass (one warp per sample): z = h . w_out + b_out + wide_sum + dense . w_d + b_d,
 * loss[b] = BCE-with-logits(z, labels[b]), dlogit[b] = (sigmoid(z) - labels[b]) * grad_scale.  h bf16 [B,N]
 * (row pitch ld), w_out / w_d / biases bf16 (biases may be NULL), dense fp32 [B,Dd], wide_sum fp32 or NULL.
 * pdj_wd_head_dz then forms the gradient entering the top ReLU layer, mask applied:
 * dz[b,j] = h[b,j] > 0 ? dlogit[b] * w_out[j] : 0. */
int pdj_wd_head(const void* h, int ld, const void* w_out, const void* b_out, const float* wide_sum, const float* dense,
                const void* w_d, const void* b_d, const float* labels, float* loss, float* dlogit, int B, int N, int Dd,
                float grad_scale, cudaStream_t st);
int pdj_wd_head_dz(const float* dlogit, const void* w_out, const void* h, int ldh, void* dz, int ldz, int B, int N,
                   cudaStream_t st);
int pdj_softmax_xent(const void* logits, const int64_t* labels, void* dlogits, float* loss_sum, int rows,
                     int ncls, float grad_scale, cudaStream_t st);

#ifdef __cplusplus
}
#endif
#endif /* PDJOB_B200_H_ */
