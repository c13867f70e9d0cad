/*
 * kubeml_b200 native runtime — C ABI.
 *
 * One shared library (kubeml_b200/_native/libkubeml_b200.so), built for sm_100a only, called from
 * Python through ctypes with raw device pointers and a cudaStream_t.  Every function returns 0 on
 * success or a non-zero code; kml_last_error() returns the message of the last failure on the
 * calling thread.  Nothing here allocates device memory unless its comment says so.
 *
 * Reference parity is cited per entry point as path:line into zzengcs/kubeML.
 */
#ifndef KUBEML_B200_H
#define KUBEML_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KML_MAX_RANKS 8

const char* kml_last_error(void);
int kml_version(void);
/* Number of kernels this library has launched in the calling process (bench "gpu_launches"). */
unsigned long long kml_launch_count(void);
void kml_reset_launch_count(void);
/* Streaming multiprocessors of the current device (148 on B200): the number every split / tile plan of the
 * Python side is sized against. */
int kml_num_sms(void);

/* ------------------------------------------------------------------------------------------------
 * Weight arena: device memory allocated through the CUDA VMM API so that it can be exported to the
 * other per-GPU worker processes as a POSIX fd and mapped there as peer memory, and bound to an
 * NVSwitch multicast object.  Replaces the RedisAI tensor store of the reference
 * (python/kubeml/kubeml/network.py:416-461, ml/pkg/model/model.go:135-161, 249-269).
 * ---------------------------------------------------------------------------------------------- */
typedef struct kml_arena kml_arena_t;

/* Allocates `bytes` (rounded up to the allocation granularity) on `device`.  want_fabric_fd != 0
 * makes the allocation exportable.  *out receives the handle. */
int kml_arena_create(int device, size_t bytes, int exportable, kml_arena_t** out);
int kml_arena_destroy(kml_arena_t* a);
void* kml_arena_ptr(const kml_arena_t* a);
size_t kml_arena_size(const kml_arena_t* a);
/* Returns a new fd owned by the caller (close it after sending it to the peers). */
int kml_arena_export_fd(kml_arena_t* a, int* fd_out);
/* Maps a peer's exported allocation (fd received over a Unix socket) into this process with
 * read/write access for `device`.  The fd may be closed afterwards. */
int kml_arena_import_fd(int device, int fd, size_t bytes, kml_arena_t** out);
/* Grants `device` access to an arena owned by this process (single-process multi-GPU). */
int kml_arena_grant_access(kml_arena_t* a, int device);

/* Multicast (NVLS) object spanning `num_devices` GPUs.  Rank 0 creates and exports it, the others
 * import; everyone then adds its device, binds its arena and maps the multicast VA. */
typedef struct kml_mcast kml_mcast_t;
int kml_mcast_supported(int device, int* supported);
int kml_mcast_create(int num_devices, size_t bytes, kml_mcast_t** out);
int kml_mcast_export_fd(kml_mcast_t* m, int* fd_out);
int kml_mcast_import_fd(int fd, int num_devices, size_t bytes, kml_mcast_t** out);
int kml_mcast_add_device(kml_mcast_t* m, int device);
int kml_mcast_bind(kml_mcast_t* m, int device, kml_arena_t* arena, void** mc_ptr_out);
int kml_mcast_destroy(kml_mcast_t* m);

/* ------------------------------------------------------------------------------------------------
 * K-AVG merge (the north-star kernel).  Replaces Model.Update + ParallelSGD.Average + Model.Save +
 * N x __load_model (ml/pkg/model/model.go:249-302, parallelSGD.go:26-54, network.py:416-442) and the
 * HTTP /next barrier (ml/pkg/train/api.go:100-126) with one kernel per rank.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  /* replica[r]: address, valid on the calling GPU, of rank r's fp32 weight region (peer mapped).
   * Only ranks in `participants` are read; every rank in `write_mask` is written. */
  float* replica[KML_MAX_RANKS];
  long long* ireplica[KML_MAX_RANKS]; /* int64 region (BatchNorm num_batches_tracked), may be NULL */
  uint32_t* flags[KML_MAX_RANKS];     /* flags[r]: rank r's signal pad, >= 2*KML_MAX_RANKS words  */
  float* mcast;                       /* multicast address of the fp32 region, or NULL           */
  size_t n_f32;                       /* floats per replica (multiple of 4, 16-byte aligned)     */
  size_t n_i64;
  size_t n_sgd;                       /* floats covered by grad/momentum_buf (parameter prefix)  */
  int my_rank;
  uint32_t participants;              /* bit r: rank r contributes a replica this round          */
  uint32_t write_mask;                /* bit r: rank r receives the merged model                 */
  uint32_t generation;                /* strictly increasing per merge; 0 disables the barrier   */
  /* Fused last local step of the round + optimizer reset (network.py:121-128): applied to this
   * rank's own replica before the exchange when fuse_sgd != 0. */
  int fuse_sgd;
  const float* grad;
  float* momentum_buf;                /* zeroed after the merge when non-NULL (optimizer reset)   */
  float lr, weight_decay, momentum;
  void* w_bf16;                       /* if non-NULL: bf16 copy of the first n_sgd merged floats,
                                         refreshed by this rank after the final barrier           */
  unsigned long long timeout_cycles;  /* spin limit for the cross-rank barrier; 0 = default       */
  int* status;                        /* device int; on a barrier timeout bit r is set for every
                                         rank r that did not arrive, and the kernel returns without
                                         touching any replica or optimizer state                  */
  /* fuse_sgd == 2: the fused last local step is Adam (function_vgg11.py's optimizer) instead of SGD:
   * momentum_buf = exp_avg, state2 = exp_avg_sq (both zeroed after the merge: the optimizer reset), step count
   * on the device (step_dev[0], reset to 0 after the merge), lr / weight_decay as above. */
  float* state2;
  float beta1, beta2, eps;
  int* step_dev;
  /* fuse_sgd applies the step to the first n_fuse floats only (0 = all n_sgd): the caller already updated the
   * rest - the layers whose gradients were final early in the backward pass - overlapped with that pass. */
  size_t n_fuse;
} kml_merge_params;

/* Bits of *status beyond the per-rank "did not arrive" bits 0..7. */
#define KML_MERGE_EVICTED 0x40000000   /* the other ranks declared THIS rank dead: it did not merge   */
#define KML_MERGE_LATE    0x20000000   /* a peer vanished after the merge had started (bits 0..7: who) */
/* Failure handling of the device merge (the reference divides by the notifications it received and lets
 * an epoch stand if one function answered, ml/pkg/train/job.go:383-398, train/util.go:144-166):
 * marks rank `who` dead from generation `from_gen` on in every pad listed in p->flags and raises its
 * flags to `tombstone` so that nobody waits for it; from_gen = 0 clears the mark (the rank re-joins). */
int kml_merge_mark_dead(const kml_merge_params* p, int who, uint32_t from_gen, uint32_t tombstone, void* stream);

/* Launches the merge on `stream` of the current device.  With generation == 0 no cross-rank
 * barrier is executed and the calling rank reduces the whole buffer (single-GPU emulation). */
int kml_kavg_merge(const kml_merge_params* p, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Optimizers over the flat arena (torch.optim.SGD / Adam semantics; function_lenet.py:79-81,
 * function_resnet34.py:62, function_vgg11.py).  One launch covers every parameter.
 * w_bf16 (optional) receives the updated weights rounded to bf16 for the tensor-core operands.
 * ---------------------------------------------------------------------------------------------- */
int kml_sgd_step(float* w, const float* g, float* momentum_buf, size_t n, float lr,
                 float weight_decay, float momentum, int nesterov, void* w_bf16, void* stream);
int kml_adam_step(float* w, const float* g, float* exp_avg, float* exp_avg_sq, size_t n, float lr,
                  float beta1, float beta2, float eps, float weight_decay, int step, void* w_bf16,
                  void* stream);
/* Same update with the step count on the device: step_dev = int[2] {steps done, internal ticket},
 * both zero at (re)set; the kernel advances it, so a CUDA-graph replay keeps the bias correction right. */
int kml_adam_step_dev(float* w, const float* g, float* exp_avg, float* exp_avg_sq, size_t n, float lr,
                      float beta1, float beta2, float eps, float weight_decay, int* step_dev, void* w_bf16,
                      void* stream);
int kml_cast_f32_to_bf16(const float* src, void* dst_bf16, size_t n, void* stream);
int kml_fill_f32(float* dst, float value, size_t n, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused softmax-cross-entropy (F.cross_entropy, mean reduction; function_resnet34.py:69) forward +
 * backward + accuracy count in one launch.  logits [rows, classes] fp32 or bf16 (logits_bf16 != 0),
 * labels int64.  Outputs: *loss (mean), dlogits (same dtype/shape as logits, already divided by
 * rows, may be NULL), *correct (int, argmax == label, may be NULL).  `scratch` must hold
 * kml_softmax_ce_scratch_bytes(rows) bytes of device memory; the kernel leaves it re-usable.
 * ---------------------------------------------------------------------------------------------- */
size_t kml_softmax_ce_scratch_bytes(int rows);
int kml_softmax_ce(const void* logits, int logits_bf16, const long long* labels, int rows,
                   int classes, float* loss, void* dlogits, int* correct, void* scratch,
                   void* stream);

/* ------------------------------------------------------------------------------------------------
 * tcgen05 / TMEM / TMA GEMM family (csrc/gemm_tcgen05.cu).  The reference reaches these shapes
 * through cuDNN / cuBLAS inside torch 1.7.1 (ml/environment/Dockerfile:3,20); SURVEY 2.4 lists them.
 * All operands bf16, fp32 accumulation in TMEM.
 * ---------------------------------------------------------------------------------------------- */
/* BatchNorm fused behind a cluster-mode GEMM (kml_epilogue.bn_fuse).  The GEMM's CTAs accumulate the
 * per-channel sums, synchronise the whole grid in software (the grid must be co-resident:
 * kml_gemm_max_coresident), and finish the layer while the tile is still on chip - convolution +
 * BatchNorm (+ residual + ReLU) in ONE launch forward, data gradient + BatchNorm backward in one launch
 * backward.  The reference gets these as separate cuDNN/ATen calls.  Output must be bf16.
 *   mode 1, forward : tile = conv output x.  Writes raw_out = x (bf16, backward needs it) and
 *                     out = relu?(gamma*(x-mean)*invstd + beta + res?); save_mean/save_invstd, running
 *                     statistics (momentum, unbiased variance) and num_batches_tracked like kml_bn_fwd_train.
 *   mode 2, backward: tile (+ addend_bf16) masked by mask_y is the gradient g entering the BatchNorm whose
 *                     forward saved bn_x / save_mean / save_invstd.  Writes raw_out = g when non-NULL (the
 *                     residual branch needs it), out = gamma*invstd*(g - sum_g/M - xhat*sum_gxhat/M),
 *                     dgamma, dbeta.
 * scratch: the kml_bn_scratch_bytes() launch state shared with the kml_bn_* kernels of the same stream. */
typedef struct {
  int mode;
  void* scratch;
  const float* gamma;
  const float* beta;               /* forward */
  float* running_mean;             /* forward, may be NULL */
  float* running_var;
  long long* num_batches_tracked;  /* forward, may be NULL */
  float* save_mean;                /* forward: written; backward: read */
  float* save_invstd;
  const void* res;                 /* forward: bf16 residual, may be NULL */
  void* raw_out;
  int relu;                        /* forward: ReLU after the residual add */
  float eps, momentum;
  const void* mask_y;              /* backward: bf16 output of the layer (ReLU mask), may be NULL */
  const void* bn_x;                /* backward: bf16 BatchNorm input saved by forward */
  float* dgamma;
  float* dbeta;
} kml_bn_fuse;

/* Largest grid (in CTAs) of the cluster-mode GEMM variant (block_n, operand majors) with clusters of
 * `cluster_size` that is guaranteed to be resident at once on the current device: the bound a fused
 * BatchNorm launch must respect.  Returns <= 0 on error. */
int kml_gemm_max_coresident(int block_n, int a_mn, int b_mn, int cluster_size);
/* Diagnostics: resident CTAs per SM of the 64-wide instantiation (bn_mode 0/1/2) per the occupancy calculator;
 * kml_last_error() then holds the function attributes it was computed from. */
int kml_gemm_blocks_per_sm(int block_n, int b_mn, int bn_mode);

typedef struct {
  const float* bias;        /* per output column, added by K-split 0; may be NULL                   */
  const void* addend_bf16;  /* bf16 tensor shaped like the output, added before ReLU; may be NULL   */
  int relu;
  int out_f32;              /* 0: bf16 output, 1: fp32 output                                       */
  int splits;               /* K-splits; 0/1 = off.  Combined with red.add into a zeroed fp32 output, or */
  int slabs;                /* (slabs != 0) each split stores its partial tile to out + split*M*ldo (fp32
                               [splits][M][ldo]) for a consumer that folds them (kml_bn_*)             */
  int block_n;              /* 64, 128, 256, or 0 = choose                                          */
  int cluster;              /* != 0: the `splits` CTAs of a tile form a thread-block cluster (<= 8) and add
                               their partial tiles through distributed shared memory: the output is complete
                               (bf16 or fp32, bias/addend/ReLU applied), no slabs, no atomics            */
  float* stats;             /* cluster mode only, may be NULL: fp32 [2][N], += per-column sum and sum of
                               squares of the (rounded) output over all rows - the BatchNorm batch
                               statistics, produced where the data already is (zero it before the GEMM;
                               kml_bn_apply_train re-zeroes it after use)                                */
  /* cluster mode, data-gradient GEMMs feeding a BatchNorm backward (all optional): the output becomes the
     masked gradient g = out * (mask_y > 0), and when bn_x is given `stats` receives the two sums that
     layer's backward needs, sum g and sum g * (bn_x - bn_mean) * bn_invstd, instead of sum / sum of squares */
  const void* mask_y;       /* bf16, shaped like the output                                          */
  const void* bn_x;         /* bf16, shaped like the output: the BatchNorm input saved by forward     */
  const float* bn_mean;     /* fp32 [N]                                                               */
  const float* bn_invstd;   /* fp32 [N]                                                               */
  const kml_bn_fuse* bn_fuse; /* cluster mode only, may be NULL: finish a BatchNorm in this launch         */
} kml_epilogue;

/* out[M,N] = A * B^T.  a_mn = 0: A is [M][K] (lda = row stride); a_mn = 1: A is [K][M].  Same for B
 * with N.  Leading dimensions in elements, multiples of 8.  (A MN-major with B K-major is not built.) */
int kml_gemm_bf16(const void* A, int a_mn, long long lda, const void* B, int b_mn, long long ldb,
                  void* out, long long ldo, int M, int N, int K, const kml_epilogue* epi, void* stream);
/* Stride-1 NHWC convolution as implicit GEMM: x[N,H,W,Cin], w[Cout,KH,KW,Cin] -> y[N,Ho,Wo,Cout].
 * Needs Cin % 64 == 0, Cout % 8 == 0 and a power-of-two output grid (else use im2col + gemm). */
int kml_conv2d_fprop(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout,
                     int KH, int KW, int pad, const kml_epilogue* epi, void* stream);
/* The same with a stride (ResNet's 3x3/2 and 1x1/2 convolutions): still no im2col buffer, the tensor map
 * walks the input with an element stride.  Output grid (H+2p-K)/s+1 must be a power of two. */
int kml_conv2d_fprop_strided(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout,
                             int KH, int KW, int pad, int stride, const kml_epilogue* epi, void* stream);
int kml_conv2d_wgrad_strided(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout,
                             int KH, int KW, int pad, int stride, int splits, int block_n, void* stream);
/* Data gradient of the same convolution from the forward filter (no transposed weight copy). */
int kml_conv2d_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                     int KH, int KW, int pad, const kml_epilogue* epi, void* stream);
/* Weight gradient dw[Cout,KH,KW,Cin] (fp32; zero it first when splits > 1). */
int kml_conv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout,
                     int KH, int KW, int pad, int splits, int block_n, void* stream);

/* ------------------------------------------------------------------------------------------------
 * NHWC bf16 layers around the GEMMs (csrc/norm.cu, csrc/nn_misc.cu); ATen/cuDNN in the reference
 * (BatchNorm2d, MaxPool2d, AdaptiveAvgPool2d, accuracy, data augmentation).
 * ---------------------------------------------------------------------------------------------- */
size_t kml_bn_scratch_bytes(int C);
/* y = relu?(bn(x) + res?) with batch statistics; updates running stats and num_batches_tracked.
 * Input: either x (bf16 [M][C]) or x_slabs (fp32 [nslabs][M][C], the split-K partial outputs of the
 * producing GEMM, slab_stride elements apart) whose sum is the input; in the slab case x_out receives
 * the folded input as bf16 for the backward pass.  scratch: kml_bn_scratch_bytes() bytes, zeroed
 * ONCE at allocation and then owned by these kernels (it carries a launch counter). */
/* Same forward when the producing GEMM already accumulated the batch statistics (kml_epilogue.stats):
 * an element-wise pass, no reduction.  stats: kml_bn_stats_bytes(C) bytes = fp32 [2][C] + a ticket word,
 * zeroed once at allocation; this kernel clears it again after use. */
size_t kml_bn_stats_bytes(int C);
int kml_bn_apply_train(const void* x, float* stats, const void* res, void* y, const float* gamma, const float* beta,
                       float* running_mean, float* running_var, long long* num_batches_tracked, float* save_mean,
                       float* save_invstd, int M, int C, int relu, float eps, float momentum, void* stream);
/* Backward counterpart: g = masked gradient and stats = {sum g, sum g*xhat} come from the producing
 * data-gradient GEMM (kml_epilogue.mask_y / bn_x); dx = gamma*invstd*(g - s1/M - xhat*s2/M), dgamma = s2,
 * dbeta = s1.  Element-wise; clears stats after use. */
int kml_bn_bwd_apply(const void* g, const void* x, const float* gamma, const float* save_mean,
                     const float* save_invstd, float* stats, void* dx, float* dgamma, float* dbeta, int M, int C,
                     void* stream);
int kml_bn_fwd_train(const void* x, const float* x_slabs, int nslabs, long long slab_stride, void* x_out,
                     const void* res, void* y, const float* gamma, const float* beta, float* running_mean,
                     float* running_var, long long* num_batches_tracked, float* save_mean,
                     float* save_invstd, void* scratch, int M, int C, int relu, float eps, float momentum,
                     void* stream);
int kml_bn_fwd_eval(const void* x, const float* x_slabs, int nslabs, long long slab_stride, const void* res,
                    void* y, const float* gamma, const float* beta, const float* running_mean,
                    const float* running_var, int M, int C, int relu, float eps, void* stream);
/* dy (bf16) or dy_slabs (fp32 split-K slabs) [+ dy_add]: grad of the block output; y: block output
 * (ReLU mask) or NULL; x: saved conv output.  Writes dx, optionally dres (masked grad for the
 * residual branch), dgamma, dbeta. */
int kml_bn_bwd(const void* dy, const float* dy_slabs, int nslabs, long long slab_stride, const void* dy_add,
               const void* y, const void* x, const float* gamma, const float* save_mean,
               const float* save_invstd, void* dx, void* dres, float* dgamma, float* dbeta, void* scratch,
               int M, int C, void* stream);
int kml_maxpool_fwd(const void* x, void* y, void* idx_u8, int N, int H, int W, int C, int k, int stride,
                    int pad, void* stream);
int kml_maxpool_bwd(const void* dy, const void* idx_u8, void* dx, int N, int H, int W, int C, int k,
                    int stride, int pad, void* stream);
int kml_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int kml_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);
/* cols[m][(kh*KW+kw)*C + c], row stride ldk (zero padded): for convolutions the shifted-TMA path
 * cannot express (stride 2, Cin < 64, non power-of-two grids). */
int kml_im2col(const void* x, void* cols, int N, int H, int W, int C, int KH, int KW, int stride, int pad,
               int ldk, void* stream);
int kml_col2im(const void* dcols, const void* addend, void* dx, int N, int H, int W, int C, int KH, int KW,
               int stride, int pad, int ldk, void* stream);
int kml_colsum_bf16(const void* x, float* out, int rows, int cols, long long ld, void* stream);
/* kml_image_prep with the flip / crop offsets drawn in the kernel from a hash of (seed, call number, sample);
 * state = uint64[2] on the device, zero at start, advanced by the kernel (CUDA-graph replay safe).  Replaces the
 * per-sample PIL transforms of ml/experiments/kubeml/function_resnet34.py:17-30. */
int kml_image_prep_rand(const void* x_u8, void* y_bf16, unsigned long long* state, unsigned seed, int N, int H, int W,
                        int C, int pad, const float* mean3, const float* std3, const long long* labels_in,
                        long long* labels_out /* both or neither: labels_out[0..N) = labels_in[0..N) */, void* stream);
/* slots[idx] = %globaltimer (ns), stream-ordered: per-phase timeline of a captured step (no reference
 * counterpart; the reference logs one wall-clock line around its merge, ml/pkg/train/job.go:397,412). */
int kml_stamp(unsigned long long* slots, int idx, void* stream);
/* dst_host[0..n) = src[0..n): device scalars stored straight into page-locked (device-visible) host memory by a
 * kernel, stream-ordered; the host reads them after synchronising with the stream.  Stands where the reference
 * calls loss.item() on every batch (python/kubeml/kubeml/network.py:288-300). */
int kml_post_f32(const float* src, float* dst_host, int n, void* stream);
int kml_pack_rows_bf16(const float* src, void* dst_bf16, int rows, int cols, int ld, void* stream);
/* VGG-style layers (ReLU fused in the producing GEMM, Dropout p=0.5 in the classifier, the
 * AdaptiveAvgPool2d((7,7)) of torchvision vgg on a 1x1 map; ml/experiments/kubeml/function_vgg11.py). */
int kml_relu_bwd(const void* dy, const void* y, void* g, float scale, size_t n, void* stream);
/* g = scale * v * (y > 0) with v = dy (bf16) or the sum of `nslabs` fp32 split-K slabs; dbias (optional)
 * receives the column sums of g (the bias gradient) in the same pass. */
int kml_relu_bwd_cols(const void* dy, const float* dy_slabs, int nslabs, size_t slab_stride, const void* y, void* g,
                      float scale, float* dbias, int M, int C, void* stream);
/* y(bf16) = relu?(sum of fp32 slabs + bias): the epilogue of a split-K GEMM. */
int kml_bias_act(const float* x_slabs, int nslabs, size_t slab_stride, const float* bias, void* y, int relu, int M,
                 int C, long long ld_x, long long ld_y, void* stream);
int kml_dropout_fwd(const void* x, void* y, void* mask_u8, unsigned long long* counter, unsigned int layer,
                    float p, size_t n, void* stream);
int kml_dropout_bwd(const void* dy, const void* mask_u8, void* dx, float p, size_t n, void* stream);
int kml_bump_counter(unsigned long long* counter, void* stream);
int kml_replicate_fwd(const void* x, void* y, int N, int C, int R2, void* stream);
int kml_replicate_bwd(const void* dy, void* dx, int N, int C, int R2, void* stream);
/* uint8 NHWC images -> normalised bf16 NHWC; aug[n] = flip | dy<<8 | dx<<16 (crop offsets in
 * [0, 2*pad]) or NULL (function_resnet34.py:17-30 on the GPU). mean3/std3 are host arrays. */
int kml_image_prep(const void* x_u8, void* y_bf16, const int* aug, int N, int H, int W, int C, int pad,
                   const float* mean3, const float* std3, void* stream);

/* ------------------------------------------------------------------------------------------------
 * kml_softmax_ce_noalloc — the same fused softmax-cross-entropy, for callers that may not allocate.
 * (No counterpart in the reference; it exists for integrators, see below.)
 *
 * Who this is for: code that has to run the loss without ever asking the CUDA allocator for
 * memory — a host that has handed all of its HBM to pre-planned arenas and must not fail or
 * fragment on a late cudaMalloc, a latency-critical loop that cannot afford the device-wide
 * synchronisation a cudaMalloc/cudaFree can imply, or an embedding where another component owns
 * the allocator.  kml_softmax_ce needs a caller-provided scratch buffer; this variant keeps that
 * scratch inside the library, in a fixed `static __device__` array, so the caller passes only
 * inputs and outputs.  It calls no cudaMalloc, cudaFree or other allocating routine.
 *
 * Limits and rules:
 *  - rows <= KML_SOFTMAX_CE_NOALLOC_MAX_ROWS (kml_softmax_ce_noalloc_max_rows()); larger batches
 *    are rejected with an error message, use kml_softmax_ce for those.  The array costs
 *    16 + 4*MAX_ROWS bytes on every GPU that loads the library, until the process exits.
 *  - The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *    that device current) and caches it; make that first call before any stream capture begins.
 *  - All calls on a device share the one array.  Calls are safe back to back on one stream (or
 *    otherwise ordered); two calls that can overlap on the GPU (different streams, no dependency)
 *    race on it and give wrong results.
 *  - Arguments and results are exactly those of kml_softmax_ce, minus `scratch`.
 * ---------------------------------------------------------------------------------------------- */
#define KML_SOFTMAX_CE_NOALLOC_MAX_ROWS 16384
int kml_softmax_ce_noalloc_max_rows(void);
int kml_softmax_ce_noalloc(const void* logits, int logits_bf16, const long long* labels, int rows,
                           int classes, float* loss, void* dlogits, int* correct, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KUBEML_B200_H */
