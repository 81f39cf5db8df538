// K5: fused classification losses (forward + gradient in one pass).
//
//   softmax_nll : logits -> softmax -> clip(p, eps, 1-eps) -> -w[t] * log p[t]   (+ d/dlogits)
//   sigmoid_bce : logit  -> sigmoid -> clip -> -(y log p + (1-y) log(1-p))       (+ d/dlogit)
//
// Restates ZooClassNLLCriterion (zoo/.../keras/objectives/ZooClassNLLCriterion.scala:48-121:
// probability input clipped to [1e-8, 1-1e-8], optional class weights, paddingValue rows
// contribute nothing, zero- or one-based labels, size-average) composed with the SoftMax the
// reference puts inside the model (NeuralCF.scala:93-96), SparseCategoricalCrossEntropy.scala:47-80
// and BinaryCrossEntropy.scala:36-42.  The reference evaluates these per sample on a thread
// pool; here one warp owns one row and the loss, the probabilities and the logits gradient
// leave the kernel together.
#include "../common/zoo_common.cuh"

namespace zoo {

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

// loss_out[0] += sum of per-row losses * loss_scale ; loss_out[1] += number of non-padding rows.
template <typename T>
__global__ void __launch_bounds__(256)
softmax_nll_kernel(const T* __restrict__ logits, long long ldl, const long long* __restrict__ labels, int B, int C,
                   int label_offset, long long padding_value, int has_padding, const float* __restrict__ class_w,
                   float eps, float loss_scale, float grad_scale, float* __restrict__ loss_out,
                   float* __restrict__ row_loss, T* __restrict__ dlogits, long long ldd, float* __restrict__ probs,
                   long long ldp) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  __shared__ float blk_loss[8];
  __shared__ float blk_cnt[8];
  float my_loss = 0.f, my_cnt = 0.f;
  if (row < B) {
    const T* x = logits + (long long)row * ldl;
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, ldf(x + c));
    mx = warp_max(mx);
    float sum = 0.f;
    for (int c = lane; c < C; c += 32) sum += __expf(ldf(x + c) - mx);
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    const long long raw = labels[row];
    const bool pad = has_padding && raw == padding_value;
    const long long t = raw - label_offset;
    const bool valid = !pad && t >= 0 && t < C;
    float w = 1.f, pt = 1.f;
    bool clipped = false;
    if (valid) {
      w = class_w ? class_w[t] : 1.f;
      pt = __expf(ldf(x + t) - mx) * inv;
      clipped = (pt < eps) || (pt > 1.f - eps);
      pt = fminf(fmaxf(pt, eps), 1.f - eps);
    }
    // d(-w log clip(p_t))/dlogit_c = w (p_c - [c == t]) when p_t is inside the clip range, else 0
    const float gs = (valid && !clipped) ? w * grad_scale : 0.f;
    for (int c = lane; c < C; c += 32) {
      const float p = __expf(ldf(x + c) - mx) * inv;
      if (probs) probs[(long long)row * ldp + c] = p;
      if (dlogits) stf(dlogits + (long long)row * ldd + c, gs * (p - (c == t ? 1.f : 0.f)));
    }
    if (lane == 0) {
      const float l = valid ? -w * __logf(pt) : 0.f;
      if (row_loss) row_loss[row] = l;
      my_loss = l;
      my_cnt = valid ? 1.f : 0.f;
    }
  }
  if (lane == 0) { blk_loss[threadIdx.x >> 5] = my_loss; blk_cnt[threadIdx.x >> 5] = my_cnt; }
  __syncthreads();
  if (threadIdx.x == 0 && loss_out) {
    float l = 0.f, n = 0.f;
    for (int i = 0; i < 8; ++i) { l += blk_loss[i]; n += blk_cnt[i]; }
    atomicAdd(loss_out, l * loss_scale);
    atomicAdd(loss_out + 1, n);
  }
}

// Few classes (C <= 16, e.g. 5 ratings or binary): one THREAD per row, the row lives in registers.
template <typename T>
__global__ void __launch_bounds__(256)
softmax_nll_small_kernel(const T* __restrict__ logits, long long ldl, const long long* __restrict__ labels, int B, int C,
                         int label_offset, long long padding_value, int has_padding, const float* __restrict__ class_w,
                         float eps, float loss_scale, float grad_scale, float* __restrict__ loss_out,
                         float* __restrict__ row_loss, T* __restrict__ dlogits, long long ldd, float* __restrict__ probs,
                         long long ldp) {
  pdl_wait();
  pdl_launch_dependents();
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  __shared__ float blk_loss[8];
  __shared__ float blk_cnt[8];
  float l = 0.f, cnt = 0.f;
  if (row < B) {
    const T* x = logits + (long long)row * ldl;
    float v[16];
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      v[c] = c < C ? ldf(x + c) : -INFINITY;
      mx = fmaxf(mx, v[c]);
    }
    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      v[c] = c < C ? __expf(v[c] - mx) : 0.f;
      sum += v[c];
    }
    const float inv = 1.f / sum;
    const long long raw = labels[row];
    const bool pad = has_padding && raw == padding_value;
    const long long t = raw - label_offset;
    const bool valid = !pad && t >= 0 && t < C;
    float w = 1.f, pt = 1.f;
    bool clipped = false;
    if (valid) {
      w = class_w ? class_w[t] : 1.f;
#pragma unroll
      for (int c = 0; c < 16; ++c) if (c == (int)t) pt = v[c] * inv;
      clipped = (pt < eps) || (pt > 1.f - eps);
      pt = fminf(fmaxf(pt, eps), 1.f - eps);
      l = -w * __logf(pt);
      cnt = 1.f;
    }
    const float gs = (valid && !clipped) ? w * grad_scale : 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      if (c < C) {
        const float p = v[c] * inv;
        if (probs) probs[(long long)row * ldp + c] = p;
        if (dlogits) stf(dlogits + (long long)row * ldd + c, gs * (p - (c == (int)t ? 1.f : 0.f)));
      }
    }
    if (row_loss) row_loss[row] = l;
  }
  l = warp_sum(l);
  cnt = warp_sum(cnt);
  if ((threadIdx.x & 31) == 0) { blk_loss[threadIdx.x >> 5] = l; blk_cnt[threadIdx.x >> 5] = cnt; }
  __syncthreads();
  if (threadIdx.x == 0 && loss_out) {
    float a = 0.f, n = 0.f;
    for (int i = 0; i < 8; ++i) { a += blk_loss[i]; n += blk_cnt[i]; }
    atomicAdd(loss_out, a * loss_scale);
    atomicAdd(loss_out + 1, n);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
sigmoid_bce_kernel(const T* __restrict__ logits, const float* __restrict__ targets, long long n, int from_probs,
                   float eps, float loss_scale, float grad_scale, float* __restrict__ loss_out,
                   T* __restrict__ dlogits) {
  __shared__ float wsum[8];
  float acc = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float x = ldf(logits + i);
    const float y = targets[i];
    const float p_raw = from_probs ? x : 1.f / (1.f + __expf(-x));
    // 1 - 1e-8 rounds to 1 in fp32, so p and 1-p are floored separately (no log(0), no 0/0)
    const float p = fmaxf(p_raw, eps), q = fmaxf(1.f - p_raw, eps);
    acc += -(y * __logf(p) + (1.f - y) * __logf(q));
    if (dlogits) {
      float g;
      if (from_probs) g = -(p_raw < eps ? 0.f : y / p) + (1.f - p_raw < eps ? 0.f : (1.f - y) / q);
      else g = p_raw - y;
      stf(dlogits + i, g * grad_scale);
    }
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0 && loss_out) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += wsum[i];
    atomicAdd(loss_out, s * loss_scale);
  }
}

}  // namespace zoo

using namespace zoo;

// dtype: 0 = fp32 logits/dlogits, 1 = bf16.  loss_out: fp32[2] = {sum(loss)*loss_scale, #valid rows},
// accumulated (zero it first).  dlogits / probs / row_loss / class_w may be null.
ZOO_API int zoo_softmax_nll(const void* logits, long long ldl, const long long* labels, int B, int C, int dtype,
                            int label_offset, long long padding_value, int has_padding, const float* class_w,
                            float eps, float loss_scale, float grad_scale, float* loss_out, float* row_loss,
                            void* dlogits, long long ldd, float* probs, long long ldp, void* stream) {
  if (!logits || !labels || B < 0 || C <= 0) { zoo_set_last_error("softmax_nll: bad arguments", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (B == 0) return ZOO_OK;
  const unsigned blocks = (unsigned)(((long long)B * 32 + 255) / 256);
  cudaStream_t s = (cudaStream_t)stream;
  if (C <= 16) {
    const unsigned sb = (unsigned)((B + 255) / 256);
    if (dtype == 0)
      softmax_nll_small_kernel<float><<<sb, 256, 0, s>>>((const float*)logits, ldl, labels, B, C, label_offset, padding_value,
                                                         has_padding, class_w, eps, loss_scale, grad_scale, loss_out, row_loss,
                                                         (float*)dlogits, ldd, probs, ldp);
    else
      softmax_nll_small_kernel<__nv_bfloat16><<<sb, 256, 0, s>>>((const __nv_bfloat16*)logits, ldl, labels, B, C, label_offset,
                                                                 padding_value, has_padding, class_w, eps, loss_scale, grad_scale,
                                                                 loss_out, row_loss, (__nv_bfloat16*)dlogits, ldd, probs, ldp);
    ZOO_CUDA_CHECK(cudaGetLastError());
    return ZOO_OK;
  }
  if (dtype == 0)
    softmax_nll_kernel<float><<<blocks, 256, 0, s>>>((const float*)logits, ldl, labels, B, C, label_offset, padding_value,
                                                     has_padding, class_w, eps, loss_scale, grad_scale, loss_out, row_loss,
                                                     (float*)dlogits, ldd, probs, ldp);
  else
    softmax_nll_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>((const __nv_bfloat16*)logits, ldl, labels, B, C, label_offset,
                                                             padding_value, has_padding, class_w, eps, loss_scale, grad_scale,
                                                             loss_out, row_loss, (__nv_bfloat16*)dlogits, ldd, probs, ldp);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_sigmoid_bce(const void* logits, const float* targets, long long n, int dtype, int from_probs, float eps,
                            float loss_scale, float grad_scale, float* loss_out, void* dlogits, void* stream) {
  if (!logits || !targets || n < 0) { zoo_set_last_error("sigmoid_bce: bad arguments", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (n == 0) return ZOO_OK;
  long long b = (n + 255) / 256;
  if (b > 148 * 8) b = 148 * 8;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == 0)
    sigmoid_bce_kernel<float><<<(unsigned)b, 256, 0, s>>>((const float*)logits, targets, n, from_probs, eps, loss_scale, grad_scale, loss_out, (float*)dlogits);
  else
    sigmoid_bce_kernel<__nv_bfloat16><<<(unsigned)b, 256, 0, s>>>((const __nv_bfloat16*)logits, targets, n, from_probs, eps, loss_scale, grad_scale, loss_out, (__nv_bfloat16*)dlogits);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
