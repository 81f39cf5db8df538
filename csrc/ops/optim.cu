// K1 (local part) + utilities: fused dense optimizers over the flat parameter arena.
//
// The reference runs Adam as eight separate MKL level-1 passes over the weight slice
// (zoo/.../pipeline/api/keras/optimizers/Adam.scala:59-106) and AdamWeightDecay as a similar
// chain (AdamWeightDecay.scala:75-135); SGD / Adagrad / Ftrl come from BigDL OptimMethods that
// Estimator assigns per sub-module (Estimator.scala:65-68, CensusWideAndDeep.scala:122-132).
// Here every optimizer is ONE pass: 16 B/param read + 12 B/param written for Adam.
// The cross-GPU version of the same pass (reduce-scatter + optimizer + all-gather over NVLink
// peer memory) lives in csrc/comm/rs_opt_ag.cu and shares zoo_opt_update() below.
#include "../common/zoo_common.cuh"
#include "optim_math.cuh"

namespace zoo {

// p/g/m/v: fp32; hyper: device fp32[ZOO_HYPER_LEN]; p16: optional bf16 shadow of the weights.
template <bool VEC>
__global__ void __launch_bounds__(256)
dense_opt_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                 __nv_bfloat16* __restrict__ p16, long long n, int opt, const float* __restrict__ hyper,
                 int zero_grad, float* __restrict__ gmut) {
  pdl_wait();
  pdl_launch_dependents();
  const OptHyper h = load_hyper(hyper, opt);
  const long long stride = (long long)gridDim.x * blockDim.x;
  if (VEC) {
    const long long n4 = n >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
      float4 pp = reinterpret_cast<float4*>(p)[i];
      const float4 gg = reinterpret_cast<const float4*>(g)[i];
      float4 mm = m ? reinterpret_cast<float4*>(m)[i] : make_float4(0, 0, 0, 0);
      float4 vv = v ? reinterpret_cast<float4*>(v)[i] : make_float4(0, 0, 0, 0);
      zoo_opt_update(pp.x, gg.x, mm.x, vv.x, h);
      zoo_opt_update(pp.y, gg.y, mm.y, vv.y, h);
      zoo_opt_update(pp.z, gg.z, mm.z, vv.z, h);
      zoo_opt_update(pp.w, gg.w, mm.w, vv.w, h);
      reinterpret_cast<float4*>(p)[i] = pp;
      if (m) reinterpret_cast<float4*>(m)[i] = mm;
      if (v) reinterpret_cast<float4*>(v)[i] = vv;
      if (p16) {
        uint2 pk; pk.x = pack_bf16(pp.x, pp.y); pk.y = pack_bf16(pp.z, pp.w);
        reinterpret_cast<uint2*>(p16)[i] = pk;
      }
      if (zero_grad) reinterpret_cast<float4*>(gmut)[i] = make_float4(0, 0, 0, 0);
    }
    // scalar tail
    for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      float pp = p[i], mm = m ? m[i] : 0.f, vv = v ? v[i] : 0.f;
      zoo_opt_update(pp, g[i], mm, vv, h);
      p[i] = pp; if (m) m[i] = mm; if (v) v[i] = vv;
      if (p16) p16[i] = __float2bfloat16(pp);
      if (zero_grad) gmut[i] = 0.f;
    }
  } else {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      float pp = p[i], mm = m ? m[i] : 0.f, vv = v ? v[i] : 0.f;
      zoo_opt_update(pp, g[i], mm, vv, h);
      p[i] = pp; if (m) m[i] = mm; if (v) v[i] = vv;
      if (p16) p16[i] = __float2bfloat16(pp);
      if (zero_grad) gmut[i] = 0.f;
    }
  }
}

// hyper[H_STEP] += 1 (keeps the bias-correction step on the device so the launch sequence is
// identical every iteration and can be replayed from a CUDA graph).
__global__ void hyper_step_kernel(float* hyper, const float* __restrict__ src) {
  // Optional refresh of every slot but the step from a staging copy of the host-side values.
  if (src && threadIdx.x < ZOO_HYPER_LEN && threadIdx.x != H_STEP) hyper[threadIdx.x] = src[threadIdx.x];
  if (threadIdx.x == 0) hyper[H_STEP] += 1.f;
}

// Sum of squares with a scratch array of per-block partials, finished by the last block.
// scratch layout: float partial[kNormBlocks]; unsigned counter (after the partials).
constexpr int kNormBlocks = 1024;

__global__ void __launch_bounds__(256)
sumsq_kernel(const float* __restrict__ x, long long n, float* __restrict__ partial, unsigned* __restrict__ counter,
             float* __restrict__ out, int accumulate) {
  __shared__ float wsum[8];
  __shared__ bool last;
  float acc = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float t = x[i];
    acc += t * t;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += wsum[i];
    partial[blockIdx.x] = s;
    __threadfence();
    const unsigned done = atomicAdd(counter, 1u);
    last = (done == gridDim.x - 1);
  }
  __syncthreads();
  if (last) {
    // deterministic final reduction in block order
    float s = 0.f;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) s += __ldcg(partial + i);
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int i = 0; i < 8; ++i) t += wsum[i];
      out[0] = accumulate ? out[0] + t : t;
      *counter = 0;  // leave the scratch ready for the next call
    }
  }
}

// g *= min(1, max_norm / (sqrt(sumsq) + 1e-6))  (L2NormClipping) or clamp to [lo, hi] (ConstantClipping)
__global__ void __launch_bounds__(256)
clip_kernel(float* __restrict__ g, long long n, const float* __restrict__ sumsq, float max_norm, int constant_mode,
            float lo, float hi) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  float scale = 1.f;
  if (!constant_mode) {
    const float norm = sqrtf(sumsq[0]);
    scale = norm > max_norm ? max_norm / (norm + 1e-6f) : 1.f;
    if (scale == 1.f) return;
  }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    g[i] = constant_mode ? fminf(fmaxf(g[i], lo), hi) : g[i] * scale;
  }
}

// out[c] += sum_r x[r, c]   (bias gradient of a Dense layer; x bf16 with leading dim ldx, out fp32)
// Each thread owns 8 consecutive columns (one 16-byte load per row); `lpr` lanes cover a row
// and a warp walks 32/lpr rows at a time.  Requires ldx % 8 == 0 (checked by the launcher).
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, int rows, int cols, float* __restrict__ out,
                   int rows_per_block, int lpr, int col_base) {
  __shared__ float red[8][32][9];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int sub = lane % lpr;              // which 8-column group of the row
  const int rsel = lane / lpr;             // which row of the warp's row bundle
  const int rows_per_warp = 32 / lpr;
  const int c0 = col_base + blockIdx.x * (lpr * 8) + sub * 8;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (c0 < cols) {
    const bool full = c0 + 8 <= cols;
    for (int r = r0 + wid * rows_per_warp + rsel; r < r1; r += 8 * rows_per_warp) {
      const __nv_bfloat16* p = x + (long long)r * ldx + c0;
      if (full) {
        const uint4 u = ld_nc_u4(p);
        const float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y), c = unpack_bf16(u.z), d = unpack_bf16(u.w);
        acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y; acc[4] += c.x; acc[5] += c.y; acc[6] += d.x; acc[7] += d.y;
      } else {
        for (int t = 0; t < 8; ++t) if (c0 + t < cols) acc[t] += __bfloat162float(p[t]);
      }
    }
  }
#pragma unroll
  for (int t = 0; t < 8; ++t) red[wid][lane][t] = acc[t];
  __syncthreads();
  // thread (t8 = threadIdx.x % 8 column, g = threadIdx.x / 8 -> 8-column group `sub`) finishes the reduction
  const int t8 = threadIdx.x & 7, sg = threadIdx.x >> 3;
  if (sg < lpr) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w)
      for (int rr = 0; rr < rows_per_warp; ++rr) s += red[w][rr * lpr + sg][t8];
    const int c = col_base + blockIdx.x * (lpr * 8) + sg * 8 + t8;
    if (c < cols) atomicAdd(out + c, s);
  }
}

template <typename TI, typename TO>
__global__ void __launch_bounds__(256) cast_kernel(const TI* __restrict__ in, TO* __restrict__ out, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = (TO)(float)in[i];
}

static int grid_for(long long n, int per_thread) {
  long long b = (n / per_thread + 255) / 256;
  if (b < 1) b = 1;
  if (b > 148 * 8) b = 148 * 8;
  return (int)b;
}

}  // namespace zoo

using namespace zoo;

// One fused optimizer pass over n fp32 parameters.  opt: see optim_math.cuh (OPT_*).
ZOO_API int zoo_dense_opt_step(float* p, float* g, float* m, float* v, void* p16, long long n, int opt,
                               const float* hyper, int zero_grad, void* stream) {
  if (!p || !g || !hyper || n < 0) { zoo_set_last_error("dense_opt_step: bad arguments", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (n == 0) return ZOO_OK;
  if ((opt_needs_m(opt) && !m) || (opt_needs_v(opt) && !v)) {
    zoo_set_last_error("dense_opt_step: optimizer state missing", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (!opt_needs_v(opt)) v = nullptr;
  const bool vec = !(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) && !((uintptr_t)p16 & 7);
  cudaStream_t s = (cudaStream_t)stream;
  if (vec) ZOO_CUDA_CHECK(zoo_launch_pdl(dense_opt_kernel<true>, dim3(grid_for(n, 4)), dim3(256), 0, s, p, g, m, v, (__nv_bfloat16*)p16, n, opt, hyper, zero_grad, g));
  else     dense_opt_kernel<false><<<grid_for(n, 1), 256, 0, s>>>(p, g, m, v, (__nv_bfloat16*)p16, n, opt, hyper, zero_grad, g);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_hyper_step(float* hyper, const float* src, void* stream) {
  if (!hyper) return ZOO_ERR_ARG;
  hyper_step_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(hyper, src);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

// out[0] (=|+=) sum(x^2).  scratch: at least zoo_sumsq_scratch_bytes() bytes, zero-initialised once.
ZOO_API long long zoo_sumsq_scratch_bytes() { return (long long)(kNormBlocks * sizeof(float) + 256); }

ZOO_API int zoo_sumsq(const float* x, long long n, void* scratch, float* out, int accumulate, void* stream) {
  if (!x || !scratch || !out || n < 0) { zoo_set_last_error("sumsq: bad arguments", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  int blocks = grid_for(n, 4);
  if (blocks > kNormBlocks) blocks = kNormBlocks;
  float* partial = (float*)scratch;
  unsigned* counter = (unsigned*)((char*)scratch + kNormBlocks * sizeof(float));
  sumsq_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, n, partial, counter, out, accumulate);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_clip_l2(float* g, long long n, const float* sumsq, float max_norm, void* stream) {
  if (!g || !sumsq) return ZOO_ERR_ARG;
  if (n == 0) return ZOO_OK;
  clip_kernel<<<grid_for(n, 1), 256, 0, (cudaStream_t)stream>>>(g, n, sumsq, max_norm, 0, 0.f, 0.f);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_clip_const(float* g, long long n, float lo, float hi, void* stream) {
  if (!g) return ZOO_ERR_ARG;
  if (n == 0) return ZOO_OK;
  clip_kernel<<<grid_for(n, 1), 256, 0, (cudaStream_t)stream>>>(g, n, nullptr, 0.f, 1, lo, hi);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_colsum_bf16(const void* x, long long ldx, int rows, int cols, float* out, void* stream) {
  if (!x || !out || rows < 0 || cols <= 0 || (ldx % 8) || ((uintptr_t)x & 15)) {
    zoo_set_last_error("colsum_bf16: bad arguments (rows need 16-byte alignment)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (rows == 0) return ZOO_OK;
  int groups = (cols + 7) / 8;             // 8-column groups per row
  int lpr = 1;
  while (lpr < groups && lpr < 32) lpr <<= 1;
  const int col_tiles = (groups + lpr - 1) / lpr;
  int chunks = (148 * 4) / col_tiles;
  if (chunks < 1) chunks = 1;
  int rpb = (rows + chunks - 1) / chunks;
  const int min_rows = 8 * (32 / lpr) * 4;
  if (rpb < min_rows) rpb = min_rows;
  chunks = (rows + rpb - 1) / rpb;
  colsum_bf16_kernel<<<dim3(col_tiles, chunks), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, ldx, rows, cols, out, rpb, lpr, 0);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream) {
  if (!in || !out) return ZOO_ERR_ARG;
  if (n == 0) return ZOO_OK;
  cast_kernel<float, __nv_bfloat16><<<grid_for(n, 1), 256, 0, (cudaStream_t)stream>>>(in, (__nv_bfloat16*)out, n);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_cast_bf16_to_f32(const void* in, float* out, long long n, void* stream) {
  if (!in || !out) return ZOO_ERR_ARG;
  if (n == 0) return ZOO_OK;
  cast_kernel<__nv_bfloat16, float><<<grid_for(n, 1), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)in, out, n);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
