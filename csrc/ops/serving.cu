// K13 / K14 / K15 -- Cluster Serving batch pre/post-processing and the recommender's top-N.
//
// K13 serving_preproc: decoded u8 HWC images (any size each) -> bilinear resize -> optional centre
//   crop -> per-channel normalise -> CHW (or HWC) fp32/bf16 written straight into the batch slot.
//   Replaces the per-image OpenCV chain + Scala copy loop of the reference
//   (zoo/.../serving/PreProcessing.scala:24-43, ClusterServing.scala:209-216) and the image
//   preprocessing chain ImageResize -> ImageCenterCrop -> ImageChannelNormalize -> ImageMatToTensor
//   (zoo/.../feature/image/*.scala).  Resize uses OpenCV's INTER_LINEAR geometry
//   (src = (dst + 0.5) * scale - 0.5, clamped to the border).
// K14 serving_topk: per-row (optional softmax) + top-k -> (index, probability) pairs
//   (zoo/.../serving/PostProcessing.scala:83-116, TensorUtils.scala:22).
// K15 recommend: per-pair argmax class + its probability, and the "first N of every key segment"
//   filter used after sorting (zoo/.../models/recommendation/Recommender.scala:46-104).
#include "../common/zoo_common.cuh"
#include <float.h>

namespace zoo {

struct PreprocArgs {
  const uint8_t* src;          // packed images
  const long long* offsets;    // [B] byte offset of image b in `src`
  const int* hw;               // [B,2] source height, width
  void* out;                   // [B,C,OH,OW] (chw) or [B,OH,OW,C]
  int B, C, OH, OW;            // output geometry (after crop)
  int RH, RW;                  // resize target (before crop); crop offset = ((RH-OH)/2, (RW-OW)/2)
  float mean[4], inv_std[4];
  int swap_rb;                 // BGR <-> RGB
  int chw;                     // 1: CHW planes, 0: HWC
  int out_bf16;
};

__global__ void __launch_bounds__(256) serving_preproc_kernel(PreprocArgs a) {
  const int b = blockIdx.y;
  const int npix = a.OH * a.OW;
  const int sh = a.hw[2 * b], sw = a.hw[2 * b + 1];
  const uint8_t* img = a.src + a.offsets[b];
  const float sy = (float)sh / (float)a.RH, sx = (float)sw / (float)a.RW;
  const int cy = (a.RH - a.OH) / 2, cx = (a.RW - a.OW) / 2;
  const int C = a.C;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += gridDim.x * blockDim.x) {
    const int oy = p / a.OW, ox = p - oy * a.OW;
    float fy = ((float)(oy + cy) + 0.5f) * sy - 0.5f;
    float fx = ((float)(ox + cx) + 0.5f) * sx - 0.5f;
    int y0 = (int)floorf(fy), x0 = (int)floorf(fx);
    float wy = fy - (float)y0, wx = fx - (float)x0;
    if (y0 < 0) { y0 = 0; wy = 0.f; }
    if (x0 < 0) { x0 = 0; wx = 0.f; }
    int y1 = y0 + 1, x1 = x0 + 1;
    if (y0 >= sh - 1) { y0 = sh - 1; y1 = y0; wy = 0.f; }
    if (x0 >= sw - 1) { x0 = sw - 1; x1 = x0; wx = 0.f; }
    const uint8_t* r0 = img + ((size_t)y0 * sw) * C;
    const uint8_t* r1 = img + ((size_t)y1 * sw) * C;
    for (int c = 0; c < C; ++c) {
      const int sc = (a.swap_rb && C >= 3 && c < 3) ? 2 - c : c;
      const float v00 = r0[x0 * C + sc], v01 = r0[x1 * C + sc], v10 = r1[x0 * C + sc], v11 = r1[x1 * C + sc];
      const float top = v00 + wx * (v01 - v00), bot = v10 + wx * (v11 - v10);
      const float v = (top + wy * (bot - top) - a.mean[c]) * a.inv_std[c];
      const size_t o = a.chw ? (((size_t)b * C + c) * a.OH + oy) * a.OW + ox : (((size_t)b * a.OH + oy) * a.OW + ox) * C + c;
      if (a.out_bf16) reinterpret_cast<__nv_bfloat16*>(a.out)[o] = __float2bfloat16(v);
      else reinterpret_cast<float*>(a.out)[o] = v;
    }
  }
}

// ---------------------------------------------------------------------- top-k --
constexpr int kMaxTopK = 8;

template <typename T> __device__ __forceinline__ float to_float(T v);
template <> __device__ __forceinline__ float to_float<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

// one warp per row
template <typename T>
__global__ void __launch_bounds__(128) serving_topk_kernel(const T* __restrict__ x, long long ldx, int rows, int n, int k,
                                                           int softmax, int* __restrict__ out_idx, float* __restrict__ out_val) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const T* row = x + (size_t)warp * ldx;
  float tv[kMaxTopK];
  int ti[kMaxTopK];
#pragma unroll
  for (int i = 0; i < kMaxTopK; ++i) { tv[i] = -FLT_MAX; ti[i] = 0x7fffffff; }
  float mx = -FLT_MAX, sum = 0.f;
  for (int j = lane; j < n; j += 32) {
    const float v = to_float<T>(row[j]);
    if (softmax) {                       // online max / sum-exp
      const float nm = fmaxf(mx, v);
      sum = sum * __expf(mx - nm) + __expf(v - nm);
      mx = nm;
    }
    if (v > tv[kMaxTopK - 1]) {          // insert into the lane's sorted list (ties keep the lower index)
      float cv = v; int ci = j;
#pragma unroll
      for (int i = 0; i < kMaxTopK; ++i) {
        if (cv > tv[i]) { const float t0 = tv[i]; const int t1 = ti[i]; tv[i] = cv; ti[i] = ci; cv = t0; ci = t1; }
      }
    }
  }
  if (softmax) {
    const float gm = warp_max(mx);
    sum = warp_sum(sum * __expf(mx - gm));
    mx = gm;
  }
  int head = 0;                          // k rounds of warp arg-max over the lanes' list heads
  for (int r = 0; r < k; ++r) {
    float hv = -FLT_MAX; int hi = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < kMaxTopK; ++i) if (i == head) { hv = tv[i]; hi = ti[i]; }
    float bv = hv; int bi = hi;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (bi == hi && hi != 0x7fffffff) ++head;
    if (lane == 0) {
      out_idx[(size_t)warp * k + r] = bi == 0x7fffffff ? -1 : bi;
      out_val[(size_t)warp * k + r] = softmax ? __expf(bv - mx) / sum : bv;
    }
  }
}

// ------------------------------------------------------------------ recommend --
// probs [P, C] -> cls (1-based argmax, as the reference's predictClass) and its probability, plus a
// 64-bit key that sorts ascending by (class, probability) -- negate/flip on the caller side.
template <typename T>
__global__ void __launch_bounds__(256) pair_score_kernel(const T* __restrict__ probs, long long ldp, long long P, int C,
                                                         int* __restrict__ cls, float* __restrict__ prob,
                                                         long long* __restrict__ rank_key) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const T* r = probs + i * ldp;
  float best = to_float<T>(r[0]); int bi = 0;
  for (int c = 1; c < C; ++c) { const float v = to_float<T>(r[c]); if (v > best) { best = v; bi = c; } }
  cls[i] = bi + 1;
  prob[i] = best;
  if (rank_key) rank_key[i] = ((long long)(bi + 1) << 32) | (long long)__float_as_uint(fmaxf(best, 0.f));
}

// keys sorted so that equal keys are adjacent and best-first: keep[i] = i is among the first n of its segment
__global__ void __launch_bounds__(256) segment_head_kernel(const long long* __restrict__ keys, long long P, int n,
                                                           unsigned char* __restrict__ keep) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  keep[i] = (i < n || keys[i - n] != keys[i]) ? 1 : 0;
}

}  // namespace zoo

using namespace zoo;

ZOO_API int zoo_serving_preproc(const void* src, const long long* offsets, const int* hw, void* out, int B, int C, int OH,
                                int OW, int RH, int RW, const float* mean, const float* stdv, int swap_rb, int chw,
                                int out_bf16, void* stream) {
  if (!src || !offsets || !hw || !out || B <= 0 || C < 1 || C > 4 || OH <= 0 || OW <= 0 || RH < OH || RW < OW) {
    zoo_set_last_error("serving_preproc: bad arguments (need 1<=C<=4, resize >= crop)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  PreprocArgs a{};
  a.src = (const uint8_t*)src; a.offsets = offsets; a.hw = hw; a.out = out;
  a.B = B; a.C = C; a.OH = OH; a.OW = OW; a.RH = RH; a.RW = RW; a.swap_rb = swap_rb; a.chw = chw; a.out_bf16 = out_bf16;
  for (int c = 0; c < 4; ++c) { a.mean[c] = (mean && c < C) ? mean[c] : 0.f; a.inv_std[c] = (stdv && c < C) ? 1.f / stdv[c] : 1.f; }
  const int npix = OH * OW;
  dim3 grid((unsigned)((npix + 255) / 256), (unsigned)B);
  serving_preproc_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_serving_topk(const void* x, long long ldx, int rows, int n, int k, int dtype, int softmax, int* out_idx,
                             float* out_val, void* stream) {
  if (!x || !out_idx || !out_val || rows <= 0 || n <= 0 || k < 1 || k > kMaxTopK) {
    zoo_set_last_error("serving_topk: bad arguments (1 <= k <= 8)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const unsigned blocks = (unsigned)((rows + 3) / 4);
  if (dtype == 0) serving_topk_kernel<float><<<blocks, 128, 0, (cudaStream_t)stream>>>((const float*)x, ldx, rows, n, k, softmax, out_idx, out_val);
  else serving_topk_kernel<__nv_bfloat16><<<blocks, 128, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, ldx, rows, n, k, softmax, out_idx, out_val);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_pair_score(const void* probs, long long ldp, long long P, int C, int dtype, int* cls, float* prob,
                           long long* rank_key, void* stream) {
  if (!probs || !cls || !prob || P < 0 || C < 1) { zoo_set_last_error("pair_score: bad arguments", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (P == 0) return ZOO_OK;
  const unsigned blocks = (unsigned)((P + 255) / 256);
  if (dtype == 0) pair_score_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>((const float*)probs, ldp, P, C, cls, prob, rank_key);
  else pair_score_kernel<__nv_bfloat16><<<blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)probs, ldp, P, C, cls, prob, rank_key);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_segment_head(const long long* keys, long long P, int n, unsigned char* keep, void* stream) {
  if (!keys || !keep || P < 0 || n < 1) { zoo_set_last_error("segment_head: bad arguments", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (P == 0) return ZOO_OK;
  segment_head_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)stream>>>(keys, P, n, keep);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
