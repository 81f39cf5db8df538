// FP8 (E4M3) quantisation of activations for the low-precision inference path (csrc/gemm/dense_tcgen05.cu, FP8 mode).
// Stands where the reference quantises to int8 inside OpenVINO / BigDL (OpenVinoInferenceSupportive.scala:35-57,
// docs/docs/wp-bigdl.md:192-196): per-tensor scale, saturating round-to-nearest conversion.
#include "../common/zoo_common.cuh"

namespace zoo {

__device__ __forceinline__ uint32_t q_e4m3x4(float a, float b, float c, float d) {
  uint16_t lo, hi;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(lo) : "f"(a), "f"(b));
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(hi) : "f"(c), "f"(d));
  return (uint32_t)lo | ((uint32_t)hi << 16);
}

// out[r, c] = e4m3(in[r, c] * qscale) for c < cols, 0 for cols <= c < ldo (row padding up to the 16-byte granule of the GEMM)
template <typename T>
__global__ void __launch_bounds__(256)
quantize_e4m3_kernel(const T* __restrict__ in, long long ldi, uint8_t* __restrict__ out, long long ldo, long long rows, int cols, float qscale) {
  const int groups = (int)(ldo >> 2);                       // 4 output bytes per thread
  const long long total = rows * groups;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / groups;
    const int c = (int)(i - r * groups) << 2;
    float v[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) v[t] = (c + t < cols) ? (float)in[r * ldi + c + t] * qscale : 0.f;
    *reinterpret_cast<uint32_t*>(out + r * ldo + c) = q_e4m3x4(v[0], v[1], v[2], v[3]);
  }
}

// absmax of a tensor (calibration): one atomicMax on the bit pattern of the non-negative float per CTA
template <typename T>
__global__ void __launch_bounds__(256)
absmax_kernel(const T* __restrict__ in, long long n, float* __restrict__ out) {
  float m = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    m = fmaxf(m, fabsf((float)in[i]));
  m = warp_max(m);
  __shared__ float s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < 8; ++k) m = fmaxf(m, s[k]);
    atomicMax(reinterpret_cast<int*>(out), __float_as_int(m));
  }
}

}  // namespace zoo

using namespace zoo;

// in: [rows, cols] fp32 (dtype 0) or bf16 (dtype 1) with leading dimension ldi; out: E4M3 bytes [rows, ldo], ldo % 16 == 0, ldo >= cols.
ZOO_API int zoo_quantize_e4m3(const void* in, long long ldi, int dtype, void* out, long long ldo, long long rows, int cols, float qscale,
                              void* stream) {
  if (!in || !out || rows < 0 || cols <= 0 || ldo < cols || (ldo % 16) || ((uintptr_t)out & 3)) {
    zoo_set_last_error("quantize_e4m3: bad arguments (ldo must be a multiple of 16 and >= cols)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (rows == 0) return ZOO_OK;
  long long blocks = (rows * (ldo >> 2) + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == 0) quantize_e4m3_kernel<float><<<(unsigned)blocks, 256, 0, s>>>((const float*)in, ldi, (uint8_t*)out, ldo, rows, cols, qscale);
  else quantize_e4m3_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, s>>>((const __nv_bfloat16*)in, ldi, (uint8_t*)out, ldo, rows, cols, qscale);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

// *out = max(*out, max |in[i]|); out: device fp32, set it to 0 before the first call.
ZOO_API int zoo_absmax(const void* in, long long n, int dtype, float* out, void* stream) {
  if (!in || !out || n < 0) return ZOO_ERR_ARG;
  if (n == 0) return ZOO_OK;
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == 0) absmax_kernel<float><<<(unsigned)blocks, 256, 0, s>>>((const float*)in, n, out);
  else absmax_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, s>>>((const __nv_bfloat16*)in, n, out);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
