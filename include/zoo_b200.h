/* zoo_b200.h -- C ABI of libzoo_cuda.so, the hand-written sm_100a kernels of zoo-b200.
 *
 * Conventions
 *   - every launcher returns 0 on success, ZOO_ERR_* otherwise; zoo_last_error() gives the text;
 *   - pointers are device pointers unless the comment says "host";
 *   - `stream` is a cudaStream_t passed as void*; launchers enqueue work and return (no sync);
 *   - bf16 matrices are row-major with a leading dimension in ELEMENTS; rows must be 16-byte
 *     aligned where noted (TMA / vector accesses).
 * Python binds these symbols with ctypes in zoo/native/lib.py.
 */
#ifndef ZOO_B200_H_
#define ZOO_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

#define ZOO_OK 0
#define ZOO_ERR_ARG 1
#define ZOO_ERR_CUDA 2
#define ZOO_ERR_TOO_LARGE 3

const char* zoo_last_error(void);
int zoo_abi_version(void);
int zoo_device_sm_count(void);

/* ---- K3: Dense layer GEMMs on tcgen05 tensor cores (csrc/gemm/dense_tcgen05.cu) -------------
 * Replaces the Dense stacks over MKL sgemm of the reference
 * (zoo/.../models/recommendation/NeuralCF.scala:73-96, WideAndDeep.scala:146-157).
 * act: 0 none, 1 relu, 2 gelu(erf), 3 sigmoid, 4 tanh. */
int zoo_dense_fwd_bf16(const void* x, long long ldx, const void* w, long long ldw, const float* bias,
                       void* y, long long ldy, int M, int N, int K, int act, void* stream);
int zoo_dense_dgrad_bf16(const void* dy, long long lddy, const void* w, long long ldw, const void* mask,
                         long long ldmask, int mask_cols, void* dx, long long lddx, int M, int N, int K,
                         void* stream);
/* Forward that also writes the sign bits of Y (bit n%32 of relu_bits[m*bits_wpr + n/32] = Y[m,n] > 0), and the dgrad
 * that takes them as its ReLU mask (mask_cols: multiple of 32, or <= 0 for all K columns). */
int zoo_dense_fwd_bf16_bits(const void* x, long long ldx, const void* w, long long ldw, const float* bias,
                            void* y, long long ldy, int M, int N, int K, int act, unsigned int* relu_bits,
                            int bits_wpr, void* stream);
int zoo_dense_dgrad_bf16_bits(const void* dy, long long lddy, const void* w, long long ldw, const unsigned int* mask_bits,
                              int bits_wpr, int mask_cols, void* dx, long long lddx, int M, int N, int K, void* stream);
int zoo_dense_wgrad_bf16(const void* dy, long long lddy, const void* x, long long ldx, float* dw,
                         long long lddw, int M, int N, int K, void* stream);
/* wgrad + bias gradient in one kernel: db[N] += column sums of dY (an extra 16-column MMA against a tile of ones). */
int zoo_dense_wgrad_bias_bf16(const void* dy, long long lddy, const void* x, long long ldx, float* dw,
                              long long lddw, float* db, int M, int N, int K, void* stream);
int zoo_colsum_bf16(const void* x, long long ldx, int rows, int cols, float* out, void* stream);

/* ---- K8: blockwise attention, forward and backward (csrc/ops/attention.cu, csrc/ops/attention_bwd.cu) ------------
 * Replaces the dense [batch, heads, L, L] attention of TransformerLayer.scala:163-181 (mm -> mask -> softmax -> mm) and its
 * autograd.  out[B,H,L,64] = softmax(scale * q k^T + mask[b,key] (+causal)) v; q, k, v, out contiguous bf16 [B,H,L,64]; mask
 * additive fp32 [B,L] or null.  Head dimension 64 only.  The _lse form also writes one fp32 log-sum-exp per row (log2 domain),
 * which the backward needs to recompute the probabilities block by block. */
int zoo_attention_fwd_bf16(const void* q, const void* k, const void* v, void* out, const float* mask, int B, int H, int L, int D,
                           float scale, int causal, void* stream);
int zoo_attention_fwd_lse_bf16(const void* q, const void* k, const void* v, void* out, float* lse, const float* mask, int B, int H,
                               int L, int D, float scale, int causal, void* stream);
/* delta: fp32 [B,H,L] scratch (written); dq: fp32 [B,H,L,64], ZERO on entry (key blocks accumulate); dk, dv: bf16 [B,H,L,64]. */
int zoo_attention_bwd_bf16(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                           float* delta, const float* mask, float* dq, void* dk, void* dv, int B, int H, int L, int D,
                           float scale, int causal, void* stream);

/* ---- K2/K4: embedding gather and sparse optimizer (csrc/ops/embedding.cu) --------------------
 * Replaces LookupTable / LookupTableSparse inside zoo Embedding / SparseEmbedding
 * (zoo/.../pipeline/api/keras/layers/Embedding.scala:82-101, SparseEmbedding.scala:76-89).
 * tables: HOST array of `world` device pointers (row r lives on shard r % world, local row r / world). */
int zoo_embedding_gather(const long long* ids, int id_stride, int id_col, int id_col2, int B, int bag,
                         int mean, const void* const* tables, long long rows_a,
                         const void* const* tables2, long long rows_b, int world, int D,
                         int table_dtype, void* out, long long ldo, void* save_a, void* save_b, void* stream);
int zoo_embedding_scratch_init(int* rep_of_row, long long rows_local, void* stream);
/* opt: 0 SGD, 1 Adam (touched rows only), 2 Adagrad; hyper: fp32[16] {lr,b1,b2,eps,wd,step,...}.
 * Scratch (caller-owned): rep_of_row int32[rows_local] all INT_MAX, accum fp32[E*D] zero,
 * dup_flag int32[E] zero on entry; all three are restored on exit. */
int zoo_embedding_sparse_update(const long long* ids, int id_stride, int id_col, int B, int bag,
                                const void* dout, long long ldg, const void* mul, long long ldmul,
                                float grad_scale, long long rows, int world, int rank, int D,
                                float* w, void* w16, float* m, float* v, int opt, const float* hyper,
                                int* rep_of_row, float* accum, int* dup_flag, void* stream);
/* No counterpart in the reference (it has no device memory to manage).  Allocation-free variant of
 * zoo_embedding_sparse_update for callers that may not allocate device memory where they run
 * (latency-critical serving threads, host callbacks inside another runtime's stream, processes
 * whose HBM budget is already committed): the scratch is a static device array inside the
 * library, so local rows <= 1,048,576, B*bag <= 8,192 and D <= 128, else ZOO_ERR_TOO_LARGE.
 * The array (~8 MiB) is held on every GPU that loads the library.  Calls share it: do not let two
 * calls overlap on the GPU.  The first call on a device resolves the array's device address;
 * make it outside stream capture. */
int zoo_embedding_sparse_update_noalloc(const long long* ids, int id_stride, int id_col, int B, int bag,
                                        const void* dout, long long ldg, const void* mul, long long ldmul,
                                        float grad_scale, long long rows, int world, int rank, int D,
                                        float* w, void* w16, float* m, float* v, int opt,
                                        const float* hyper, void* stream);
/* Same as zoo_embedding_sparse_update plus win_list (int32 [2*B*bag + 4], any contents): pass 2 compacts the
 * representative entries into it and pass 3 runs only over that list. */
int zoo_embedding_sparse_update_w(const long long* ids, int id_stride, int id_col, int B, int bag,
                                  const void* dout, long long ldg, const void* mul, long long ldmul,
                                  float grad_scale, long long rows, int world, int rank, int D,
                                  float* w, void* w16, float* m, float* v, int opt, const float* hyper,
                                  int* rep_of_row, float* accum, int* dup_flag, int* win_list, int bump_step, void* stream);
/* Backward of a row-sharded table: ids_src / dout_src / mul_src are HOST arrays of `world` device pointers into the
 * symmetric heap (every rank's ids, gradient-row window, optional multiplier rows); the owner pulls the rows it needs over
 * NVLink.  accum fp32 [world*B*bag, D], dup_flag int32 [world*B*bag], win_list int32 [2*world*B*bag + 4].
 * Bracket with zoo_p2p_barrier. */
int zoo_embedding_sparse_update_multi(const long long* const* ids_src, int id_stride, int id_col, int B, int bag,
                                      const void* const* dout_src, long long ldg, const void* const* mul_src,
                                      long long ldmul, float grad_scale, long long rows, int world, int rank,
                                      int D, float* w, void* w16, float* m, float* v, int opt, const float* hyper,
                                      int* rep_of_row, float* accum, int* dup_flag, int* win_list, int bump_step, void* stream);
/* Backward dispatch of a row-sharded table: every source GPU pushes (local row index, final gradient row) of each entry
 * into the owner's receive buffers (recv_ids: int64 [world*cap], -1 = empty; recv_grad: bf16 [world*cap, D]; both HOST
 * arrays of `world` device pointers into the symmetric heap).  counters: int32 [world] on this GPU.  After a barrier the
 * owner runs zoo_embedding_sparse_update_w over its receive buffer with world = 1. */
int zoo_embedding_dispatch(const long long* ids, int id_stride, int id_col, int B, int bag, const void* dout, long long ldg,
                           const void* mul, long long ldmul, float grad_scale, long long rows, int world, int rank, int D,
                           void* const* recv_ids, void* const* recv_grad, int* counters, int cap, void* stream);
int zoo_embedding_dense_grad(const long long* ids, int id_stride, int id_col, int B, int bag,
                             const void* dout, long long ldg, const void* mul, long long ldmul,
                             float grad_scale, long long rows, int D, float* dtable, void* stream);

/* ---- K1 (local part): fused optimizers over the flat parameter arena (csrc/ops/optim.cu) -----
 * Replaces the 8-pass MKL Adam of zoo/.../keras/optimizers/Adam.scala:59-106 and
 * AdamWeightDecay.scala:75-135; opt ids in csrc/ops/optim_math.cuh. */
int zoo_dense_opt_step(float* p, float* g, float* m, float* v, void* p16, long long n, int opt,
                       const float* hyper, int zero_grad, void* stream);
int zoo_hyper_step(float* hyper, const float* src, void* stream);
long long zoo_sumsq_scratch_bytes(void);
int zoo_sumsq(const float* x, long long n, void* scratch, float* out, int accumulate, void* stream);
int zoo_clip_l2(float* g, long long n, const float* sumsq, float max_norm, void* stream);
int zoo_clip_const(float* g, long long n, float lo, float hi, void* stream);
int zoo_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream);
int zoo_cast_bf16_to_f32(const void* in, float* out, long long n, void* stream);

/* K3, whole-tower form (csrc/gemm/mlp3_tcgen05.cu): three Dense + ReLU layers K0 -> N1 -> N2 -> 64 in one persistent kernel; the
 * intermediate activations stay in shared memory as the next layer's A operand and are written once for backward. */
int zoo_mlp3_supported(int K0, int N1, int N2, int N3);
int zoo_mlp3_fwd_bf16(const void* x, long long ldx, const void* w1, long long ldw1, const float* b1, const void* w2, long long ldw2,
                      const float* b2, const void* w3, long long ldw3, const float* b3, void* h1, long long ldh1, void* h2,
                      long long ldh2, void* h3, long long ldh3, int M, int K0, int N1, int N2, int N3, unsigned int* bits1,
                      unsigned int* bits2, unsigned int* bits3, void* stream);

/* ---- K5: fused losses (csrc/ops/loss.cu) ------------------------------------------------------
 * ZooClassNLLCriterion.scala:48-121 + SoftMax, BinaryCrossEntropy.scala:36-42. dtype 0 fp32, 1 bf16. */
int zoo_softmax_nll(const void* logits, long long ldl, const long long* labels, int B, int C, int dtype,
                    int label_offset, long long padding_value, int has_padding, const float* class_w,
                    float eps, float loss_scale, float grad_scale, float* loss_out, float* row_loss,
                    void* dlogits, long long ldd, float* probs, long long ldp, void* stream);
int zoo_sigmoid_bce(const void* logits, const float* targets, long long n, int dtype, int from_probs, float eps,
                    float loss_scale, float grad_scale, float* loss_out, void* dlogits, void* stream);
/* K5b (csrc/ops/head_loss.cu): last Dense layer + softmax + NLL + the head's dgrad / wgrad / bias gradient in one pass over the
 * head's input (NeuralCF.scala:89-96).  x [B, K] bf16, w [C, K] bf16 (C <= 8, K % 8 == 0, K <= 256); dw / db are accumulated. */
int zoo_head_softmax_nll(const void* x, long long ldx, const void* w, long long ldw, const float* bias,
                         const long long* labels, int B, int K, int C, int zero_based, long long padding_value,
                         int has_padding, int size_average, float eps, int relu_cols, float* loss_out, void* dx,
                         long long ldd, float* dw, long long lddw, float* db, float* logits_out, long long ldl,
                         float* valid_scratch, int impl, void* stream);

/* ---- K1: fused reduce-scatter + optimizer + all-gather over NVLink peer memory (csrc/comm/rs_opt_ag.cu)
 * Replaces BigDL AllReduceParameter + OptimMethod.optimize (Topology.scala:1128,1187-1206).  Pointer
 * arguments are HOST arrays of `world` device pointers (index = rank) into the symmetric heap. */
int zoo_symm_alloc(long long nbytes, void** ptr, void* ipc_handle64);
int zoo_symm_open(const void* ipc_handle64, void** ptr);
int zoo_symm_close(void* ptr);
int zoo_symm_free(void* ptr);
/* Large-page symmetric memory (cuMemCreate + POSIX fd export / cuMemMap on the peer), csrc/comm/symm_vmm.cu */
int zoo_vmm_alloc(long long bytes, void** ptr, int* fd, long long* mapped_bytes);
int zoo_vmm_import(int fd, long long mapped_bytes, void** ptr);
int zoo_vmm_free(void* ptr, long long mapped_bytes);
int zoo_close_fd(int fd);
/* zero_grad: bit 0 = zero the gradient range afterwards; bit 1 = replicated update (every rank reduces and updates the
 * whole range on its own replica: one NVLink round trip less, for ranges of a few MB). */
int zoo_rs_opt_ag(void* const* g, void* const* p, void* const* p16, void* const* m, void* const* v,
                  void* const* flags, void* const* epoch, void* const* hyper, long long start, long long end,
                  int world, int rank, int device_slot, int opt, int zero_grad, int grid, void* stream);
/* One-to-all push over peer memory: dst[p] + dst_off <- src[0..bytes) for every rank p (dst: HOST array of `world` device
 * pointers into the symmetric heap).  Offsets and source must be 16-byte aligned. */
int zoo_symm_bcast(void* const* dst, long long dst_off, const void* src, long long bytes, int world, void* stream);
int zoo_p2p_barrier(void* const* flags, void* const* epoch, int world, int rank, int slot, void* stream);

/* ---- K6: persistent LSTM / GRU, forward and backward-through-time (csrc/ops/rnn.cu) -----------
 * Replaces the cell loops of LSTM.scala / GRU.scala / InternalRecurrent.scala and of the PyTorch
 * AutoML models (VanillaLSTM_pytorch.py:29-56).  torch.nn.LSTM/GRU weight layout, fp32, H <= 128.
 * `gates` holds x W_ih^T + b for all steps on entry and the activated gates on exit (save != 0);
 * backward overwrites it with the pre-activation gate gradients. */
int zoo_rnn_supported(int H);
int zoo_lstm_forward(float* gates, const float* whh, const float* h0, const float* c0, float* hout, float* cout,
                     float* hT, float* cT, int B, int T, int H, int save, int hard_sigmoid, void* stream);
int zoo_lstm_backward(float* gates, const float* whh, const float* hout, const float* cout, const float* h0,
                      const float* c0, const float* dy, long long dy_sb, long long dy_st, const float* dhT,
                      const float* dcT, float* dh0, float* dc0, int B, int T, int H, int hard_sigmoid, void* stream);
int zoo_gru_forward(const float* gx, float* saved, const float* whh, const float* bhn, const float* h0,
                    float* hout, float* hT, int B, int T, int H, int hard_sigmoid, void* stream);
int zoo_gru_backward(float* saved, const float* whh, const float* hout, const float* h0, const float* dy,
                     long long dy_sb, long long dy_st, const float* dhT, float* dh0, int B, int T, int H,
                     int hard_sigmoid, void* stream);

/* ---- K7: dilated causal conv1d with fused bias/ReLU/dropout/residual epilogue (csrc/ops/tcn.cu) -
 * North-star TCN forecaster (no TCN in the reference); also MTNet's CNN front-end
 * (pyzoo/zoo/automl/model/MTNet_keras.py:398-406).  x [B,Cin,T], w [Cout,Cin,K], fp32. */
int zoo_tcn_conv_forward(const float* x, const float* w, const float* bias, const float* res, float* out, int B,
                         int Cin, int Cout, int T, int K, int dilation, int relu, float drop_p,
                         unsigned long long seed, void* stream);
int zoo_tcn_conv_dgrad(const float* g, const float* w, float* dx, int B, int Cin, int Cout, int T, int K,
                       int dilation, void* stream);
int zoo_tcn_conv_wgrad(const float* g, const float* x, float* dw, int B, int Cin, int Cout, int T, int K, int dilation,
                       void* stream);
int zoo_tcn_prep_grad(const float* dout, const float* out, const float* res, float* g, float* dres, long long n,
                      int relu, float drop_p, void* stream);

/* ---- K13/K14/K15: serving pre/post-processing and recommender top-N (csrc/ops/serving.cu) ------
 * PreProcessing.scala:24-43 + ClusterServing.scala:209-216; PostProcessing.scala:83-116;
 * Recommender.scala:46-104.  dtype 0 fp32, 1 bf16. */
int zoo_serving_preproc(const void* src, const long long* offsets, const int* hw, void* out, int B, int C, int OH,
                        int OW, int RH, int RW, const float* mean /*host*/, const float* stdv /*host*/, int swap_rb,
                        int chw, int out_bf16, void* stream);
int zoo_serving_topk(const void* x, long long ldx, int rows, int n, int k, int dtype, int softmax, int* out_idx,
                     float* out_val, void* stream);
int zoo_pair_score(const void* probs, long long ldp, long long P, int C, int dtype, int* cls, float* prob,
                   long long* rank_key, void* stream);
int zoo_segment_head(const long long* keys, long long P, int n, unsigned char* keep, void* stream);

/* ---- K2/K4 grouped form (csrc/ops/embedding_group.cu): parameter blocks and entry points in zoo_emb_group.h ---- */

/* ---- K1, small-range form (csrc/comm/ar_push.cu): one-handshake all-reduce with a bf16 wire --------------------
 * Replaces AllReduceParameter's put / aggregate of gradient slices (Topology.scala:1128-1206; FP16-compressed blocks,
 * Topology.scala:1043-1047) for dense parts of a few MB.  zoo_ar_push: bf16 copies of g[0..n) into every rank's mailbox
 * (mail: HOST array of `world` device pointers, each [2][world][n] bf16; flags: [2][world][1024] uint32, zero at start;
 * epoch: local uint32 [2][1024], zero at start) -- never waits.  zoo_ar_reduce: waits for every rank's push of this step and
 * leaves the fp32 sum (rank order) in g.  Same `grid` on every rank for both; n % 8 == 0. */
int zoo_ar_push(float* g, void* const* mail, void* const* flags, unsigned int* epoch, long long n, int world, int rank,
                int grid, void* stream);
int zoo_ar_reduce(float* g, void* const* mail, void* const* flags, unsigned int* epoch, long long n, int world, int rank,
                  int grid, void* stream);

#ifdef __cplusplus
}
#endif
#include "zoo_emb_group.h"
#endif /* ZOO_B200_H_ */
