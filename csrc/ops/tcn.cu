// K7 -- temporal-convolution block kernels: dilated causal conv1d with a fused epilogue.
//
// No TCN exists in the reference (north-star extension for the time-series forecasters);
// the same kernel is the CNN front-end of MTNet (pyzoo/zoo/automl/model/MTNet_keras.py:398-406).
//   forward :  a   = dropout(relu(conv(x, w, dilation d) + bias))          (causal: taps reach back only)
//              out = residual ? relu(a + res) : a
//   backward:  tcn_prep_grad turns d(out) into the gradient of the conv's pre-activation WITHOUT a saved
//              mask: relu/dropout masks are recovered from the saved output (a == 0  <=>  masked), and
//              the same tap kernel run with mirrored offsets and transposed weights gives d(x).
// Layout: x [B, Cin, T], w [Cout, Cin, K], out [B, Cout, T], all fp32, T contiguous.
// Mapping: thread = one time step, 8 output channels in registers; grid = (T tiles, channel blocks, B).
// The weight slab of the CTA's 8 channels sits in shared memory as [ic][tap][8] so each (ic, tap)
// costs one coalesced global load of x and two broadcast 128-bit shared loads for 8 FMAs.
#include "../common/zoo_common.cuh"

namespace zoo {

constexpr int kOcPerThread = 8;
constexpr int kTcnThreads = 128;

__device__ __forceinline__ float hash_uniform(unsigned long long seed, unsigned long long idx) {
  unsigned long long z = idx + seed * 0x9E3779B97F4A7C15ull + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}

struct TapArgs {
  const float* in;      // [B, IC, T]
  const float* w;       // weight tensor; element (oc, ic, j) at oc*s_oc + ic*s_ic + j*s_j
  long long s_oc, s_ic, s_j;
  const float* bias;    // [OC] or null
  const float* res;     // [B, OC, T] or null (fused residual add + relu)
  float* out;           // [B, OC, T]
  int B, IC, OC, T, K, dil;
  int sign;             // -1: taps look back (forward conv); +1: taps look ahead (input gradient)
  int relu;
  float drop_p;
  unsigned long long seed;
};

template <bool EPI>
__global__ void __launch_bounds__(kTcnThreads) conv_taps_kernel(TapArgs a) {
  extern __shared__ __align__(16) float ws[];      // [IC][K][8]
  const int oc0 = blockIdx.y * kOcPerThread, b = blockIdx.z;
  const int t = blockIdx.x * kTcnThreads + threadIdx.x;
  for (int i = threadIdx.x; i < a.IC * a.K * kOcPerThread; i += kTcnThreads) {
    const int o = i % kOcPerThread, ij = i / kOcPerThread, j = ij % a.K, ic = ij / a.K;
    const int oc = oc0 + o;
    ws[i] = oc < a.OC ? a.w[oc * a.s_oc + ic * a.s_ic + j * a.s_j] : 0.f;
  }
  __syncthreads();
  if (t >= a.T) return;
  float acc[kOcPerThread];
#pragma unroll
  for (int o = 0; o < kOcPerThread; ++o) acc[o] = 0.f;
  const float* inb = a.in + (size_t)b * a.IC * a.T;
  for (int j = 0; j < a.K; ++j) {
    const int tt = t + a.sign * (a.K - 1 - j) * a.dil;
    if (tt < 0 || tt >= a.T) continue;             // causal zero padding / end of sequence
    const float* wj = ws + j * kOcPerThread;
    for (int ic = 0; ic < a.IC; ++ic) {
      const float xv = __ldg(inb + (size_t)ic * a.T + tt);
      const float4 w0 = *reinterpret_cast<const float4*>(wj + (size_t)ic * a.K * kOcPerThread);
      const float4 w1 = *reinterpret_cast<const float4*>(wj + (size_t)ic * a.K * kOcPerThread + 4);
      acc[0] = fmaf(w0.x, xv, acc[0]); acc[1] = fmaf(w0.y, xv, acc[1]);
      acc[2] = fmaf(w0.z, xv, acc[2]); acc[3] = fmaf(w0.w, xv, acc[3]);
      acc[4] = fmaf(w1.x, xv, acc[4]); acc[5] = fmaf(w1.y, xv, acc[5]);
      acc[6] = fmaf(w1.z, xv, acc[6]); acc[7] = fmaf(w1.w, xv, acc[7]);
    }
  }
  const float keep_scale = a.drop_p > 0.f ? 1.f / (1.f - a.drop_p) : 1.f;
#pragma unroll
  for (int o = 0; o < kOcPerThread; ++o) {
    const int oc = oc0 + o;
    if (oc >= a.OC) break;
    const size_t idx = ((size_t)b * a.OC + oc) * a.T + t;
    float v = acc[o];
    if (EPI) {
      if (a.bias) v += a.bias[oc];
      if (a.relu) v = fmaxf(v, 0.f);
      if (a.drop_p > 0.f) v = hash_uniform(a.seed, idx) < a.drop_p ? 0.f : v * keep_scale;
      if (a.res) v = fmaxf(v + a.res[idx], 0.f);
    }
    a.out[idx] = v;
  }
}

// d(out) -> gradient of the conv pre-activation (g) and of the residual input (dres).
__global__ void __launch_bounds__(256) tcn_prep_grad_kernel(const float* __restrict__ dout, const float* __restrict__ out,
                                                            const float* __restrict__ res, float* __restrict__ g,
                                                            float* __restrict__ dres, long long n, int relu, float keep_scale) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float d = dout[i];
    float a = out[i];
    if (res) {                       // out = relu(a + res)
      d = a > 0.f ? d : 0.f;
      if (dres) dres[i] = d;
      a = a - res[i];                // only meaningful where out > 0; elsewhere d is already 0
    }
    if (relu) d = a > 0.f ? d * keep_scale : 0.f;   // a == 0 <=> relu- or dropout-masked
    else d *= keep_scale;
    g[i] = d;
  }
}

// Weight gradient: dW[o][i][j] = sum over (b, t) of g[b][o][t] * x[b][i][t - (K-1-j)*dil].
// A small-output, huge-reduction GEMM (M = Cout, N = Cin per tap, reduction = B*T): each CTA owns a 64 x 64 block of
// (o, i), walks a strided share of the (b, t-tile) work items with a 4 x 4 register tile per thread, and adds its partial
// sum into dW with fp32 reductions at the end.  Taps are the outer loop (the g tile is re-read from L2, which is cheap
// next to one pass over x per tap).
constexpr int kWgTile = 64, kWgTT = 32;

__global__ void __launch_bounds__(256) tcn_wgrad_kernel(const float* __restrict__ g, const float* __restrict__ x, float* __restrict__ dw,
                                                        int B, int Cin, int Cout, int T, int K, int dil) {
  __shared__ __align__(16) float gs[kWgTT][kWgTile + 4];
  __shared__ __align__(16) float xs[kWgTT][kWgTile + 4];
  const int o0 = blockIdx.y * kWgTile, i0 = blockIdx.z * kWgTile;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;          // thread tile: o = ty*4.., i = tx*4..
  const int ttiles = (T + kWgTT - 1) / kWgTT;
  const long long items = (long long)B * ttiles;
  for (int j = 0; j < K; ++j) {
    const int shift = (K - 1 - j) * dil;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[a][c] = 0.f;
    for (long long it = blockIdx.x; it < items; it += gridDim.x) {
      const int b = (int)(it / ttiles), t0 = (int)(it - (long long)b * ttiles) * kWgTT;
      // load g[b][o0..o0+63][t0..t0+31] and x[b][i0..i0+63][t0-shift ..] transposed into [t][channel]
      for (int l = tid; l < kWgTile * kWgTT; l += 256) {
        const int ch = l / kWgTT, tt = l - ch * kWgTT, t = t0 + tt;
        const int o = o0 + ch, i = i0 + ch, ts = t - shift;
        gs[tt][ch] = (o < Cout && t < T) ? __ldg(g + ((size_t)b * Cout + o) * T + t) : 0.f;
        xs[tt][ch] = (i < Cin && t < T && ts >= 0) ? __ldg(x + ((size_t)b * Cin + i) * T + ts) : 0.f;
      }
      __syncthreads();
#pragma unroll 8
      for (int tt = 0; tt < kWgTT; ++tt) {
        const float4 a4 = *reinterpret_cast<const float4*>(&gs[tt][ty * 4]);
        const float4 b4 = *reinterpret_cast<const float4*>(&xs[tt][tx * 4]);
        const float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[a][c] = fmaf(av[a], bv[c], acc[a][c]);
      }
      __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int o = o0 + ty * 4 + a, i = i0 + tx * 4 + c;
        if (o < Cout && i < Cin && acc[a][c] != 0.f) atomicAdd(dw + ((size_t)o * Cin + i) * K + j, acc[a][c]);
      }
  }
}

}  // namespace zoo

using namespace zoo;

static int launch_taps(TapArgs& a, bool epi, void* stream) {
  const size_t sm = (size_t)a.IC * a.K * kOcPerThread * sizeof(float);
  if (sm > 200 * 1024) { zoo_set_last_error("tcn: Cin*K too large for the weight slab (Cin*K <= 6400)", __FILE__, __LINE__); return ZOO_ERR_TOO_LARGE; }
  dim3 grid((unsigned)((a.T + kTcnThreads - 1) / kTcnThreads), (unsigned)((a.OC + kOcPerThread - 1) / kOcPerThread), (unsigned)a.B);
  cudaError_t e = cudaSuccess;
  if (epi) {
    if (sm > 48 * 1024) e = cudaFuncSetAttribute(conv_taps_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    if (e == cudaSuccess) conv_taps_kernel<true><<<grid, kTcnThreads, sm, (cudaStream_t)stream>>>(a);
  } else {
    if (sm > 48 * 1024) e = cudaFuncSetAttribute(conv_taps_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    if (e == cudaSuccess) conv_taps_kernel<false><<<grid, kTcnThreads, sm, (cudaStream_t)stream>>>(a);
  }
  if (e != cudaSuccess) { zoo_set_last_error(cudaGetErrorString(e), __FILE__, __LINE__); return ZOO_ERR_CUDA; }
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

/* out = [relu(. + res)] dropout(relu(causal_conv1d(x, w, dilation) + bias)) */
ZOO_API int zoo_tcn_conv_forward(const float* x, const float* w, const float* bias, const float* res, float* out, int B,
                                 int Cin, int Cout, int T, int K, int dilation, int relu, float drop_p,
                                 unsigned long long seed, void* stream) {
  if (!x || !w || !out || B <= 0 || Cin <= 0 || Cout <= 0 || T <= 0 || K <= 0 || dilation <= 0 || drop_p < 0.f || drop_p >= 1.f) {
    zoo_set_last_error("tcn_conv_forward: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  TapArgs a{x, w, (long long)Cin * K, (long long)K, 1, bias, res, out, B, Cin, Cout, T, K, dilation, -1, relu, drop_p, seed};
  return launch_taps(a, true, stream);
}

/* dx[b,ci,s] = sum_co sum_j w[co,ci,j] * g[b,co,s + (K-1-j)*dilation] */
ZOO_API int zoo_tcn_conv_dgrad(const float* g, const float* w, float* dx, int B, int Cin, int Cout, int T, int K,
                               int dilation, void* stream) {
  if (!g || !w || !dx || B <= 0 || Cin <= 0 || Cout <= 0 || T <= 0 || K <= 0 || dilation <= 0) {
    zoo_set_last_error("tcn_conv_dgrad: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  // roles swap: "output channel" = ci, "input channel" = co; element (ci, co, j) of w[co][ci][j]
  TapArgs a{g, w, (long long)K, (long long)Cin * K, 1, nullptr, nullptr, dx, B, Cout, Cin, T, K, dilation, +1, 0, 0.f, 0ull};
  return launch_taps(a, false, stream);
}

ZOO_API int zoo_tcn_prep_grad(const float* dout, const float* out, const float* res, float* g, float* dres, long long n,
                              int relu, float drop_p, void* stream) {
  if (!dout || !out || !g || n < 0) { zoo_set_last_error("tcn_prep_grad: bad arguments", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (n == 0) return ZOO_OK;
  const long long blocks = (n + 255) / 256;
  tcn_prep_grad_kernel<<<(unsigned)(blocks > 148 * 8 ? 148 * 8 : blocks), 256, 0, (cudaStream_t)stream>>>(
      dout, out, res, g, dres, n, relu, drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

/* dw[Cout,Cin,K] += sum_{b,t} g[b,o,t] * x[b,i,t-(K-1-j)*dilation]   (dw must be zero-initialised by the caller) */
ZOO_API int zoo_tcn_conv_wgrad(const float* g, const float* x, float* dw, int B, int Cin, int Cout, int T, int K, int dilation,
                               void* stream) {
  if (!g || !x || !dw || B <= 0 || Cin <= 0 || Cout <= 0 || T <= 0 || K <= 0 || dilation <= 0) {
    zoo_set_last_error("tcn_conv_wgrad: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const long long items = (long long)B * ((T + kWgTT - 1) / kWgTT);
  const unsigned ob = (Cout + kWgTile - 1) / kWgTile, ib = (Cin + kWgTile - 1) / kWgTile;
  long long gx = (148LL * 2) / (ob * ib);
  if (gx < 1) gx = 1;
  if (gx > items) gx = items;
  tcn_wgrad_kernel<<<dim3((unsigned)gx, ob, ib), 256, 0, (cudaStream_t)stream>>>(g, x, dw, B, Cin, Cout, T, K, dilation);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
