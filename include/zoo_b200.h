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
int zoo_dense_wgrad_bf16(const void* dy, long long lddy, const void* x, long long ldx, float* dw,
                         long long lddw, int M, int N, int K, void* stream);
int zoo_colsum_bf16(const void* x, long long ldx, int rows, int cols, float* out, void* stream);

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

/* ---- K5: fused losses (csrc/ops/loss.cu) ------------------------------------------------------
 * ZooClassNLLCriterion.scala:48-121 + SoftMax, BinaryCrossEntropy.scala:36-42. dtype 0 fp32, 1 bf16. */
int zoo_softmax_nll(const void* logits, long long ldl, const long long* labels, int B, int C, int dtype,
                    int label_offset, long long padding_value, int has_padding, const float* class_w,
                    float eps, float loss_scale, float grad_scale, float* loss_out, float* row_loss,
                    void* dlogits, long long ldd, float* probs, long long ldp, void* stream);
int zoo_sigmoid_bce(const void* logits, const float* targets, long long n, int dtype, int from_probs, float eps,
                    float loss_scale, float grad_scale, float* loss_out, void* dlogits, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ZOO_B200_H_ */
