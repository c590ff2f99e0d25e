/* mtb200.h — C ABI of libmtb200.so: the sm_100a kernels and the device runtime.
 *
 * Every function returns 0 on success or a non-zero code; mtb_last_error() gives the text.
 * All pointers are device pointers unless the name says host.  `stream` is a cudaStream_t.
 * The reference ships no native code; each entry point cites the line
 * of mpipy.py whose TensorFlow / MPI library call it replaces.
 */
#ifndef MTB200_H_
#define MTB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __DRIVER_TYPES_H__
typedef struct CUstream_st* cudaStream_t;
#endif

enum {
  MTB_E_OK = 0,
  MTB_E_CUDA = 1,     /* a CUDA runtime call failed */
  MTB_E_ARG = 2,      /* bad argument / unsupported shape */
  MTB_E_TIMEOUT = 3,  /* a device-side wait on a peer ran out of time */
  MTB_E_LIMIT = 4     /* request exceeds a fixed capacity */
};

/* ---- runtime ------------------------------------------------------------------------ */
const char* mtb_last_error(void);
int mtb_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem);
/* static arena (replaces TF's BFC allocator with growth, mpipy.py:69-71): one cudaMalloc,
 * shapes are static so nothing is ever freed mid-run */
int mtb_arena_alloc(size_t bytes, void** out_ptr);
int mtb_arena_free(void* ptr);
/* CUDA IPC: export / import an arena so peers can map it (replaces host-staged MPI buffers) */
int mtb_ipc_export(void* ptr, void* out_handle64);
int mtb_ipc_import(const void* handle64, void** out_ptr);
int mtb_ipc_close(void* ptr);
int mtb_enable_peer_access(int peer_device);

/* ---- tcgen05 GEMM engine (cuBLAS matmul behind tf.matmul, mpipy.py:165,167, and its grads) */
size_t mtb_gemm_workspace_bytes(int M, int N, int bn, int splits);
size_t mtb_gemm_counter_bytes(int M, int N, int bn);
int mtb_gemm_bf16(const void* A, const void* B, void* out, int M, int N, int K, int a_mn_major, int b_mn_major,
                  long long lda, long long ldb, long long ldc, int out_bf16, const float* bias, int relu,
                  float keep_prob, unsigned long long seed, const long long* step_ptr, int bn, int splits,
                  void* workspace, void* counters, cudaStream_t stream);
/* GEMM -> reduce-scatter in ONE kernel: the product is added (red.v4.f32 over NVLink) into the flat gradient buffer of
 * whichever rank owns each 16-byte vector, so the cross-GPU reduction of this gradient overlaps the multiplication tile
 * by tile and mtb_fused_allreduce_update (prered_* ranges) skips it.  Slice r = elements [r*per, (r+1)*per). */
typedef struct {
  int world, rank;
  long long base;       /* flat element offset of the product's element (0,0) */
  long long per;        /* elements per owner slice, multiple of 4 (same partition as the two-shot / nvls update) */
  void* out[16];        /* every rank's flat fp32 gradient buffer, peer-mapped */
  int flags;            /* bit 0: add straight from the TMEM row layout (one 16-byte atomic per lane) instead of staging
                         * each chunk through shared memory so that a warp adds whole 128-byte row segments */
} mtb_gemm_rs;
int mtb_gemm_bf16_rs(const void* A, const void* B, int M, int N, int K, int a_mn_major, int b_mn_major, long long lda,
                     long long ldb, long long ldc, int bn, const mtb_gemm_rs* rs, cudaStream_t stream);
/* fp32 operands and output, TF32 tensor-core math (kind::tf32), fp32 accumulation: the fp32 training path. */
int mtb_gemm_tf32(const float* A, const float* B, float* out, int M, int N, int K, int a_mn_major, int b_mn_major,
                  long long lda, long long ldb, long long ldc, const float* bias, int relu, float keep_prob,
                  unsigned long long seed, const long long* step_ptr, int bn, int splits, void* workspace,
                  void* counters, cudaStream_t stream);

/* 5x5 SAME convolution on 14x14 maps as an implicit GEMM on the same engine (tf.nn.conv2d mpipy.py:159 and its two
 * gradients): the producer addresses the NHWC tensor through a 4-D TMA map, no im2col / col2im matrix is ever built.
 * Any channel counts that are multiples of one 128-byte row (32 fp32 / 64 bf16); tf32 != 0: fp32 tensors, TF32 math.
 * mode 0 fprop (x, w[5,5,Cin,Cout]) -> out[B,14,14,Cout] (+ bias, ReLU); mode 1 dgrad (x = dy, w) -> out[B,14,14,Cin];
 * mode 2 wgrad (x = layer input, w = dy) -> out[5,5,Cin,Cout] fp32, split-K over the batch (workspace / counters sized
 * with mtb_gemm_workspace_bytes(25*Cin, Cout, bn, splits) / mtb_gemm_counter_bytes, bn = 256 / 128 / 64 by Cout). */
int mtb_conv5x5_gemm(int mode, const void* x, const void* w, void* out, int B, int Cin, int Cout, int tf32,
                     const float* bias, int relu, int splits, void* workspace, void* counters, cudaStream_t stream);

/* Split-K GEMM with built-in scratch, for callers that cannot allocate or pass device scratch (a plug-in inside a
 * process whose allocator owns the whole GPU, an allocation-free control loop, a bare C caller without a tensor
 * library).  No counterpart in the reference.  Same product/epilogue as mtb_gemm_bf16 (bias, ReLU, bf16/fp32
 * output; no dropout).  Scratch is a fixed 4 MiB + 4 KiB device array inside the library, one per GPU, kept until
 * exit; shapes whose split-K partials need more return MTB_E_LIMIT.  Allocates nothing.  First call per device
 * resolves the array's address: do it at start-up, before any stream capture.  Calls share the array: never let two
 * of them overlap on one GPU.  bn: 0 (= 64), 64 or 128; anything else is MTB_E_ARG. */
size_t mtb_gemm_noalloc_max_workspace_bytes(void);
int mtb_gemm_bf16_splitk_noalloc(const void* A, const void* B, void* out, int M, int N, int K, int a_mn_major,
                                 int b_mn_major, long long lda, long long ldb, long long ldc, int out_bf16,
                                 const float* bias, int relu, int bn, int splits, cudaStream_t stream);

/* ---- init / data / casts (tf.truncated_normal mpipy.py:38-52; extract_data mpipy.py:215) ---- */
int mtb_init_trunc_normal(float* dst, long long n, float stddev, unsigned long long seed, unsigned stream_id,
                          cudaStream_t stream);
int mtb_fill_f32(float* dst, long long n, float value, cudaStream_t stream);
int mtb_zero_multi(void* const* ptrs, const long long* n_words, int count, cudaStream_t stream); /* <= 8 ranges of 32-bit words */
int mtb_normalize_u8_dev(const uint8_t* src, float* dst, long long n, cudaStream_t stream);
int mtb_synthetic_batch(float* images, int* labels, long long n_img, long long row_offset,
                        unsigned long long seed, int learnable, cudaStream_t stream);
int mtb_batch_window(const float* data, const int* labels, const long long* step_ptr, long long rows, int B,
                     float* x_out, int* y_out, cudaStream_t stream);                       /* mpipy.py:80-82 */
int mtb_cast_f32_to_bf16(const float* src, void* dst, long long n, cudaStream_t stream);
int mtb_cast_bf16_to_f32(const void* src, float* dst, long long n, cudaStream_t stream);
int mtb_l2_half_sumsq(const float* x, long long n, float* out_accum, cudaStream_t stream);   /* mpipy.py:57-58 */

/* ---- conv1 fused: conv5x5 + bias + ReLU + 2x2 max-pool, and its wgrad (mpipy.py:156-158) ---- */
int mtb_conv1_fwd(const float* x, const float* w, const float* bias, void* out_bf16, uint8_t* idx, int B, int C1,
                  cudaStream_t stream);
int mtb_conv1_wgrad(const float* x, const void* dP_bf16, const void* P_bf16, const uint8_t* idx, float* dW, float* db,
                    int B, int C1, cudaStream_t stream);

/* tensor-core versions for C1 == 32 (csrc/conv1_sm100.cu): same arguments minus C1, same output formats */
int mtb_conv1_fwd_tc(const float* x, const float* w, const float* bias, void* out_bf16, uint8_t* idx, int B,
                     cudaStream_t stream);
int mtb_conv1_wgrad_tc(const float* x, const void* dP_bf16, const void* P_bf16, const uint8_t* idx, float* dW, float* db,
                       int B, cudaStream_t stream);
/* "_win" forms: x / labels are the rank's resident shard (win_rows rows) and the batch is the reference's sequential
 * window offset = (step*B) % (rows-B) (mpipy.py:80-82) for the step stored on the device at *win_step, so a captured
 * step needs neither a host feed nor a separate gather kernel.  win_step == NULL: the pointer is the batch itself. */
int mtb_conv1_fwd_tc_win(const float* x, const float* w, const float* bias, void* out_bf16, uint8_t* idx, int B,
                         const long long* win_step, long long win_rows, cudaStream_t stream);
/* ... and, being the first kernel of a step, it can also clear the step's accumulators (up to 8 ranges of 32-bit
 * words, as mtb_zero_multi) so that no separate launch is needed for that */
int mtb_conv1_fwd_tc_win_zero(const float* x, const float* w, const float* bias, void* out_bf16, uint8_t* idx, int B,
                              const long long* win_step, long long win_rows, void* const* zero_ptrs,
                              const long long* zero_words, int zero_count, cudaStream_t stream);
int mtb_conv1_wgrad_tc_win(const float* x, const void* dP_bf16, const void* P_bf16, const uint8_t* idx, float* dW,
                           float* db, int B, const long long* win_step, long long win_rows, cudaStream_t stream);
int mtb_fc2_softmax_ce_fast_win(const void* h_bf16, const float* W, const float* bias, const int* labels,
                                float* logits, float* probs, float* dlogits, float* loss_accum, int* err_accum, int B,
                                int F, int NC, const long long* win_step, long long win_rows, cudaStream_t stream);

/* ---- pooling + generic-conv helpers (mpipy.py:158,161; tf.nn.conv2d mpipy.py:159 via im2col) ---- */
int mtb_maxpool2x2_fwd(const void* in, void* out, uint8_t* idx, int B, int H, int W, int C, cudaStream_t stream);
int mtb_unpool_relu_bwd(const void* dP, const void* P, const uint8_t* idx, void* dA, int B, int H, int W, int C,
                        cudaStream_t stream);
int mtb_im2col_5x5(const void* x, void* col, int B, int H, int W, int C, cudaStream_t stream);
int mtb_col2im_5x5(const void* dcol, void* dx, int B, int H, int W, int C, cudaStream_t stream);

/* ---- conv2 as implicit GEMM on tcgen05, no im2col (tf.nn.conv2d + relu + max_pool, mpipy.py:159-161, and the
 *      Conv2DBackpropInput / Conv2DBackpropFilter ops TF autodiff adds for mpipy.py:65).  14x14 maps, 32 -> 64 ch. */
int mtb_conv2_supported(int H, int cin, int cout);
int mtb_conv2_fprop_pool(const void* x_bf16, const void* w_bf16, const float* bias, void* out_bf16, uint8_t* idx, int B,
                         cudaStream_t stream);
int mtb_conv2_dgrad(const void* dy_bf16, const void* w_bf16, void* dx_bf16, int B, cudaStream_t stream);
size_t mtb_conv2_wgrad_workspace_bytes(int B);
int mtb_conv2_wgrad(const void* x_bf16, const void* dy_bf16, float* dW, float* db, void* workspace, int B,
                    cudaStream_t stream);
/* variants with MaxPoolGrad + ReluGrad fused in front: take the pooled gradient + pooled activation + argmax */
int mtb_conv2_dgrad_unpool(const void* dflat_bf16, const void* pooled_bf16, const uint8_t* idx, const void* w_bf16,
                           void* dx_bf16, int B, cudaStream_t stream);
int mtb_conv2_wgrad_unpool(const void* x_bf16, const void* dflat_bf16, const void* pooled_bf16, const uint8_t* idx,
                           float* dW, float* db, void* workspace, int B, cudaStream_t stream);
/* occupancy probe (scripts only): `grid` blocks of `block` threads that hold their SM slots for `cycles` clocks without
 * memory traffic; fat = about 56 live registers per thread instead of a handful */
int mtb_debug_occupy(int grid, int block, int fat, long long cycles, float* sink, cudaStream_t stream);
/* probe of the shared-memory descriptor semantics the implicit-GEMM convolutions rely on (test only) */
int mtb_debug_shift_mma(const void* A, const void* B, float* out, int mode, int row_bytes, int rows_total, int shift,
                        int base_off, cudaStream_t stream);

/* ---- FC2 + softmax + cross-entropy, forward and backward (mpipy.py:167, :55-56, :67-68) ---- */
int mtb_fc2_softmax_ce(const void* h_bf16, const float* W, const float* bias, const int* labels, float* logits,
                       float* probs, float* dlogits, float* loss_accum, int* err_accum, int B, int F, int NC,
                       cudaStream_t stream);
int mtb_fc2_dgrad(const float* dlogits, const float* W, const void* h_bf16, void* dh_bf16, int B, int F, int NC,
                  float scale, cudaStream_t stream);
int mtb_fc2_wgrad(const void* h_bf16, const float* dlogits, float* dW, float* db, int B, int F, int NC,
                  cudaStream_t stream);
/* engine variants: weights staged in shared memory; one backward pass giving dW2, db2, dh and colsum(dh) = db of FC1 */
int mtb_fc2_softmax_ce_fast(const void* h_bf16, const float* W, const float* bias, const int* labels, float* logits,
                            float* probs, float* dlogits, float* loss_accum, int* err_accum, int B, int F, int NC,
                            cudaStream_t stream);
int mtb_fc2_bwd_fused(const void* h_bf16, const float* dlogits, const float* W, void* dh_bf16, float* dW, float* db2,
                      float* db1, int B, int F, int NC, float scale, cudaStream_t stream);
int mtb_colsum_bf16(const void* x_bf16, float* out, long long R, int C, cudaStream_t stream);

/* ---- fp32-activation twins of the generic layer kernels (the fp32 / TF32 training path: dtype="fp32").
 * Same semantics and layouts as the functions of the same name without the suffix; every tensor is fp32. */
int mtb_conv1_fwd_f32(const float* x, const float* w, const float* bias, float* out, uint8_t* idx, int B, int C1,
                      cudaStream_t stream);
int mtb_conv1_wgrad_f32(const float* x, const float* dP, const float* P, const uint8_t* idx, float* dW, float* db,
                        int B, int C1, cudaStream_t stream);
int mtb_maxpool2x2_fwd_f32(const float* in, float* out, uint8_t* idx, int B, int H, int W, int C, cudaStream_t stream);
int mtb_unpool_relu_bwd_f32(const float* dP, const float* P, const uint8_t* idx, float* dA, int B, int H, int W, int C,
                            cudaStream_t stream);
int mtb_im2col_5x5_f32(const float* x, float* col, int B, int H, int W, int C, cudaStream_t stream);
int mtb_col2im_5x5_f32(const float* dcol, float* dx, int B, int H, int W, int C, cudaStream_t stream);
int mtb_fc2_softmax_ce_f32(const float* h, const float* W, const float* bias, const int* labels, float* logits,
                           float* probs, float* dlogits, float* loss_accum, int* err_accum, int B, int F, int NC,
                           cudaStream_t stream);
int mtb_fc2_dgrad_f32(const float* dlogits, const float* W, const float* h, float* dh, int B, int F, int NC,
                      float scale, cudaStream_t stream);
int mtb_fc2_wgrad_f32(const float* h, const float* dlogits, float* dW, float* db, int B, int F, int NC,
                      cudaStream_t stream);
int mtb_colsum_f32(const float* x, float* out, long long R, int C, cudaStream_t stream);
int mtb_fc2_softmax_ce_fast_win_f32(const float* h, const float* W, const float* bias, const int* labels,
                                    float* logits, float* probs, float* dlogits, float* loss_accum, int* err_accum,
                                    int B, int F, int NC, const long long* win_step, long long win_rows,
                                    cudaStream_t stream);
int mtb_fc2_bwd_fused_f32(const float* h, const float* dlogits, const float* W, float* dh, float* dW, float* db2,
                          float* db1, int B, int F, int NC, float scale, cudaStream_t stream);

/* ---- K1: fused cross-GPU allreduce + average + L2 + optimizer + LR schedule + bf16 cast
 *      (Gather/mean/assign mpipy.py:121-152; l2 mpipy.py:57-58; lr mpipy.py:60-64; momentum mpipy.py:65) */
typedef struct mtb_fused_args {
  int world, rank, algo /*0 one-shot, 1 two-shot, 2 nvls*/, optimizer /*-1 none, 0 sgd, 1 momentum, 2 adam*/;
  const void* grads[16];
  void* params[16];
  void* shadow[16];
  const void* mc_grads;
  void* mc_params;
  void* mc_shadow;
  void* slot_m;
  void* slot_v;
  void* flags[16];
  void* seq;
  void* status;
  long long offset, count;
  float scale, lr0, lr_decay, lr_override;
  int batch;
  long long decay_rows;
  const void* step_ptr;
  float momentum, beta1, beta2, eps, wd;
  int n_ranges;
  long long range_begin[8], range_end[8];
  int grid, block;
  /* Sharded fp32 master (two-shot / nvls only): inside these element ranges the owner of a slice keeps the only
   * up-to-date fp32 copy and broadcasts just the bf16 shadow, 2 instead of 6 bytes per parameter over NVLink.
   * Use it for tensors whose consumers read the shadow (the big GEMM / convolution weights). */
  int n_sharded;
  long long shard_begin[4], shard_end[4];
  void* advance_step;   /* if set: *advance_step += 1 once every block of this launch has finished (global_step=iter_, mpipy.py:66) */
  void* done_counter;   /* 32-bit scratch for that, zero between launches */
  /* Loss mailbox (needs advance_step): the block that bumps the step also stores {u32 new step : f32 *loss_src} as one
   * 8-byte word into loss_mail[(new step - 1) % loss_ring], a ring in pinned (device-visible) host memory, so the host
   * reads every step's loss (the fetch of sess.run([optimizer, loss]), mpipy.py:83-85) without a copy or a sync. */
  const void* loss_src;
  void* loss_mail;
  int loss_ring;        /* power of two */
  /* Pre-reduced ranges (two-shot / nvls, world > 1): element ranges of the gradient buffer whose owner slices already
   * hold the sum over ranks because the producing GEMM added its tiles into the owners directly (mtb_gemm_bf16_rs).
   * The owner reads them locally instead of reducing over peers and zeroes them for the next step. */
  int n_prered;
  long long prered_begin[2], prered_end[2];
  /* Launches that may be in flight at the same time (a bucket on a communication stream under the rest of backward, the
   * closing bucket on the main stream) must use different channels: each has its own flag and sequence blocks. */
  int chan;             /* 0 (up to 448 blocks) or 1 (up to 63 blocks) */
  int no_flags;         /* 1: no flag rounds inside this launch; the caller runs mtb_p2p_barrier before it (peers' gradients
                         * complete) and after it (peer loads and stores done) on the same stream.  Lets the launch use any
                         * number of short-lived blocks (grid < 0: one pass per block). */
  /* algo 3, "push": one-shot for a small bucket with NO flag round.  Every rank stores its gradients into its row of
   * each peer's receive area as flag-in-data words {f32, tag, f32, tag} (tag = *step_ptr + 1; rows double buffered by the
   * step's parity); the receiver polls the words it needs until their tags match and reduces in rank order.  The area
   * must be cleared (by its owner, between barriers) whenever *step_ptr is moved backwards. */
  void* push_recv[16];  /* every rank's receive area, peer-mapped: [2][world][2 * push_stride] floats, zeroed */
  long long push_stride;/* payload floats per row, >= count, multiple of 4 */
  /* bf16 gradient wire format: inside [lp_begin, lp_end) gradients are read as bf16 from grads_lp (peer loads, or
   * multimem.ld_reduce with fp32 accumulation in the switch) and everything after the load stays fp32. */
  const void* grads_lp[16];
  const void* mc_grads_lp;
  long long lp_begin, lp_end;
  unsigned dst_mask;    /* two-shot only: bit r set = rank r receives the result; 0 = every rank.  Bit 0 alone is the
                         * reference's root-only assign (mpipy.py:129) when the kernel averages parameters. */
  /* L2 term of the loss (mpipy.py:57-58): with l2_blocks set the launch sums 0.5*w^2 (weights before the update) over the
   * range_* it walks and stores {step+1 : value} into row l2_row of the slots (its own entry on every rank when it only
   * walked an owner slice, entry 0 locally otherwise).  The launch that mails the loss adds wd * (all entries listed in
   * l2_rows_n) to *loss_src.  Needs done_counter and step_ptr. */
  void* l2_blocks;      /* float[grid] scratch, one per channel */
  void* l2_slots[16];   /* every rank's uint64[4][16], peer-mapped */
  int l2_row;
  int l2_rows_n[4];     /* entries to add per row: 0 unused, 1 local, world = one per slice owner */
  void* trace;          /* optional uint64[2]: %globaltimer (ns) when the launch's first block starts and when it closes */
  const void* err_src;  /* int32 error count of the batch, mailed as {step : errors} into aux_mail (same ring size) */
  void* aux_mail;
} mtb_fused_args;
int mtb_fused_allreduce_update(const mtb_fused_args* a, cudaStream_t stream);
int mtb_p2p_barrier(const mtb_fused_args* a, cudaStream_t stream);          /* comm.Barrier(), mpipy.py:93 */
int mtb_advance_step(long long* step_ptr, cudaStream_t stream);             /* global_step=iter_, mpipy.py:66 */
int mtb_fused_max_blocks(void);
int mtb_fused_args_size(void);          /* sizeof(mtb_fused_args), for bindings to check their layout against */
int mtb_fused_max_l2_blocks(void);      /* capacity needed for l2_blocks, in floats */
int mtb_fused_flag_bytes(void);

#ifdef __cplusplus
}
#endif
#endif /* MTB200_H_ */
