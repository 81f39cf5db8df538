// K7, tensor-core form -- the dilated causal conv1d of csrc/ops/tcn.cu as an implicit GEMM on `mma.sync.m16n8k8` TF32
// (fp32 in / fp32 accumulate, operands rounded to TF32 with cvt.rna: the precision cuDNN's convolutions use by default).
//
// Why warp-level MMA and not tcgen05 here: the B operand of a tap is the input tile shifted by (K-1-j)*dilation time steps.
// With time contiguous in memory ([B, C, T]) that shift is an arbitrary element offset along the operand's contiguous
// dimension, which a UMMA shared-memory descriptor cannot express (its start address moves in 16-byte, swizzle-aligned
// steps); register fragments read with plain `ld.shared` can start anywhere.  At C = 64 the whole layer is 3.8 GFLOP per
// direction, far from the tensor-pipe limit either way -- the kernel is bound by staging the tiles.
//
//   conv_taps_mma_kernel   forward (sign -1, fused bias / ReLU / dropout / residual epilogue) and input gradient (sign +1):
//                          out[m][t] = sum_j sum_k Wt[j][m][k] * in[k][t + sign*(K-1-j)*dil]
//                          CTA = one sample, 64 output channels, TT time steps; the input chunk [KC][TT + halo] stays in
//                          shared memory for all taps, the taps' 64 x KC weight slabs go through a 3-deep cp.async ring.
//                          A (weights) by ldmatrix, B (shifted input) by conflict-free ld.shared.
//   tcn_wgrad_mma_kernel   dW[o][i][j] = sum_{b,t} g[b][o][t] * x[b][i][t - (K-1-j)*dil]: CTA = one tap, a 64 x 64 block of
//                          (o, i), a strided share of the (sample, time tile) items; both operands by ldmatrix.
//   tcn_prep_weights_kernel  w -> Wt[j][m][k] (zero padded, TF32-rounded), so the slabs are contiguous 16-byte rows.
// Reference: the time-series forecasters' CNN front-end, pyzoo/zoo/automl/model/MTNet_keras.py:398-406.
#include "../common/zoo_common.cuh"

namespace zoo {

__device__ __forceinline__ float tcn_hash_uniform(unsigned long long seed, unsigned long long idx) {   // same stream as tcn.cu
  unsigned long long z = idx + seed * 0x9E3779B97F4A7C15ull + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ float to_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src));
}

__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src));
}

// Round-to-nearest to TF32 for a fragment register: the tensor core reads the upper 19 bits, so adding half a TF32 ulp to the
// magnitude rounds (ties away from zero, like cvt.rna).  `cvt.rna.tf32.f32` itself is two instructions on sm_100 (an inf / nan
// test around the same add); the test is dropped here -- an inf operand turns into nan instead of staying inf.
__device__ __forceinline__ uint32_t tf32_bits(float v) { return __float_as_uint(v) + 0x1000u; }

// One warp stages `ncols` floats of a time row into shared memory with cp.async (nothing waits here: every request of the tile is
// in flight at once -- a load / convert / store loop costs one L2 round trip per iteration).  Column c holds time tstart + c;
// columns outside [0, T), or a null row (channel beyond the tensor), are zero.  `vec`: tstart and T are multiples of 4, so chunks
// of 4 columns are 16-byte aligned on both sides.  The values stay fp32 in shared memory; fragments are rounded to TF32 when read.
__device__ __forceinline__ void stage_row(float* dst, const float* row, int tstart, int ncols, int T, bool vec, int lane) {
  const uint32_t d32 = smem_u32(dst);
  if (vec) {
    for (int c = lane * 4; c < ncols; c += 128) {
      const int t = tstart + c;
      if (row && t >= 0 && t + 3 < T && c + 3 < ncols) {
        cp_async16(d32 + c * 4u, row + t);
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (c + e >= ncols) break;
          if (row && t + e >= 0 && t + e < T) cp_async4(d32 + (c + e) * 4u, row + t + e);
          else dst[c + e] = 0.f;
        }
      }
    }
  } else {
    for (int c = lane; c < ncols; c += 32) {
      const int t = tstart + c;
      if (row && t >= 0 && t < T) cp_async4(d32 + c * 4u, row + t);
      else dst[c] = 0.f;
    }
  }
}

constexpr int kMmaThreads = 256;   // conv_taps: 8 warps = 4 time quarters x 2 halves of the 64 output channels
constexpr int kMmaM = 64;          // output channels per CTA
constexpr int kSlabs = 3;          // weight-slab ring depth (one tap's MMAs are shorter than an L2 round trip)

// w (element (m, k, j) at m*s_m + k*s_k + j*s_j) -> wt[j][Mp][Kp], zero beyond (M, Kd), TF32-rounded
__global__ void __launch_bounds__(256) tcn_prep_weights_kernel(const float* __restrict__ w, float* __restrict__ wt, long long s_m,
                                                               long long s_k, long long s_j, int M, int Kd, int Mp, int Kp, int K) {
  const long long n = (long long)K * Mp * Kp, stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int k = (int)(i % Kp), m = (int)((i / Kp) % Mp), j = (int)(i / ((long long)Kp * Mp));
    wt[i] = (m < M && k < Kd) ? to_tf32(w[m * s_m + k * s_k + j * s_j]) : 0.f;
  }
}

struct MmaTapArgs {
  const float* in;      // [B, IC, T]
  const float* wt;      // [K][Mp][Kp]  (Mp = OC rounded up to 64, Kp = IC rounded up to KC)
  const float* bias;    // [OC] or null
  const float* res;     // [B, OC, T] or null
  float* out;           // [B, OC, T]
  int B, IC, OC, T, K, dil;
  int Mp, Kp, KC;       // KC: input channels staged per pass (multiple of 8, <= 64)
  int XS;               // row stride of the staged input tile in floats (== 8 mod 32: the B-fragment loads hit 32 distinct banks)
  int sign, relu;
  float drop_p;
  unsigned long long seed;
};

// NT = 8-wide time tiles per warp; the CTA covers TT = 4 * NT * 8 time steps.
template <int NT, bool EPI>
__global__ void __launch_bounds__(kMmaThreads) conv_taps_mma_kernel(MmaTapArgs a) {
  constexpr int TT = 4 * NT * 8;
  extern __shared__ __align__(16) float smem[];
  const int WS = a.KC + 4;                          // weight slab row stride (== 4 mod 32 rows of 16-byte chunks on distinct bank groups)
  float* ws = smem;                                 // [kSlabs][64][WS]: ring of tap slabs, fetched two taps ahead
  float* xs = smem + kSlabs * kMmaM * WS;           // [KC][XS]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t4 = lane & 3;
  const int wt = warp & 3, wm = warp >> 2;          // time quarter, half of the output channels (m-tiles 2*wm, 2*wm + 1)
  const int b = blockIdx.z, m0 = blockIdx.y * kMmaM, t0 = blockIdx.x * TT;
  const int halo = (a.K - 1) * a.dil;
  const int tbase = a.sign < 0 ? t0 - halo : t0;    // time of column 0 of the staged tile
  const int cols = TT + halo;
  const float* inb = a.in + (size_t)b * a.IC * a.T;
  const bool vec = (a.T & 3) == 0 && (halo & 3) == 0 && (reinterpret_cast<uintptr_t>(a.in) & 15) == 0;   // 16-byte chunks line up

  float acc[2][NT][4];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[mt][nt][c] = 0.f;

  const uint32_t ws_u32 = smem_u32(ws);
  // ldmatrix row of this lane inside a 16 x 8 (tf32) A tile: matrices 0..3 = (rows 0-7, k 0-3), (rows 8-15, k 0-3), (rows 0-7, k 4-7), (rows 8-15, k 4-7)
  const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 4;
  const int chunks16 = a.KC / 4;                    // 16-byte chunks per slab row

  for (int kc0 = 0; kc0 < a.Kp; kc0 += a.KC) {
    __syncthreads();                                // everybody is done with the previous input chunk and weight slabs
    auto fetch_slab = [&](int j) {                  // one commit group per call, empty past the last tap
      if (j < a.K) {
        const float* src = a.wt + ((size_t)j * a.Mp + m0) * a.Kp + kc0;
        const uint32_t dstb = ws_u32 + (uint32_t)((j % kSlabs) * kMmaM * WS) * 4u;
        for (int i = tid; i < kMmaM * chunks16; i += kMmaThreads) {
          const int r = i / chunks16, c = i - r * chunks16;
          cp_async16(dstb + (uint32_t)(r * WS + c * 4) * 4u, src + (size_t)r * a.Kp + c * 4);
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // the input chunk (same commit group as slab 0): zero outside [0, T) and beyond IC
    for (int r = warp; r < a.KC; r += kMmaThreads / 32) {
      const int ic = kc0 + r;
      stage_row(xs + (size_t)r * a.XS, ic < a.IC ? inb + (size_t)ic * a.T : nullptr, tbase, cols, a.T, vec, lane);
    }
    fetch_slab(0);
    fetch_slab(1);
    for (int j = 0; j < a.K; ++j) {
      asm volatile("cp.async.wait_group 1;" ::: "memory");     // slab j landed (slab j+1 may still be in flight)
      __syncthreads();                              // slab j (and the input chunk) visible; compute of tap j-1 finished everywhere
      fetch_slab(j + 2);                            // into the buffer tap j-1 used
      const uint32_t wj = ws_u32 + (uint32_t)((j % kSlabs) * kMmaM * WS) * 4u;
      const int shift = a.sign < 0 ? j * a.dil : (a.K - 1 - j) * a.dil;     // column of output time 0 for this tap
      const float* xcol = xs + wt * (NT * 8) + shift + g;
      for (int kk = 0; kk < a.KC; kk += 8) {
        uint32_t bf[NT][2];
        const float* x0 = xcol + (size_t)(kk + t4) * a.XS;
        const float* x1 = x0 + (size_t)4 * a.XS;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          bf[nt][0] = tf32_bits(x0[nt * 8]);
          bf[nt][1] = tf32_bits(x1[nt * 8]);
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          uint32_t af[4];
          ldsm_x4(af, wj + (uint32_t)(((wm * 2 + mt) * 16 + a_row) * WS + kk + a_col) * 4u);
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) mma_tf32(acc[mt][nt], af, bf[nt][0], bf[nt][1]);
        }
      }
    }
  }

  // epilogue: c0 = (row g, col 2*t4), c1 = (g, 2*t4 + 1), c2 = (g + 8, 2*t4), c3 = (g + 8, 2*t4 + 1)
  const float keep_scale = a.drop_p > 0.f ? 1.f / (1.f - a.drop_p) : 1.f;
  const bool pair_ok = (a.T & 1) == 0;
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int oc = m0 + (wm * 2 + mt) * 16 + g + h * 8;
      if (oc >= a.OC) continue;
      const float bv = (EPI && a.bias) ? a.bias[oc] : 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int t = t0 + wt * (NT * 8) + nt * 8 + 2 * t4;
        if (t >= a.T) continue;
        const size_t idx = ((size_t)b * a.OC + oc) * a.T + t;
        float v[2] = {acc[mt][nt][h * 2], acc[mt][nt][h * 2 + 1]};
        const int nv = (t + 1 < a.T) ? 2 : 1;
        if (EPI) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            if (e >= nv) break;
            float y = v[e] + bv;
            if (a.relu) y = fmaxf(y, 0.f);
            if (a.drop_p > 0.f) y = tcn_hash_uniform(a.seed, idx + e) < a.drop_p ? 0.f : y * keep_scale;
            if (a.res) y = fmaxf(y + a.res[idx + e], 0.f);
            v[e] = y;
          }
        }
        if (nv == 2 && pair_ok) *reinterpret_cast<float2*>(a.out + idx) = make_float2(v[0], v[1]);
        else { a.out[idx] = v[0]; if (nv == 2) a.out[idx + 1] = v[1]; }
      }
    }
}

// ---- weight gradient -----------------------------------------------------------------------------------------------------
constexpr int kWg1Threads = 128;     // single-tap kernel: 4 warps, one 16-row o tile each
constexpr int kWgT = 64;             // time steps per work item
constexpr int kWgS = kWgT + 4;       // row stride (== 4 mod 32)

__global__ void __launch_bounds__(kWg1Threads) tcn_wgrad_mma_kernel(const float* __restrict__ gsrc, const float* __restrict__ x,
                                                                    float* __restrict__ dw, int B, int Cin, int Cout, int T, int K,
                                                                    int dil, int iblocks) {
  __shared__ __align__(16) float gs[kMmaM * kWgS];     // [o][t]
  __shared__ __align__(16) float xs[kMmaM * kWgS];     // [i][t - shift]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t4 = lane & 3;
  const int j = blockIdx.y, o0 = (blockIdx.z / iblocks) * kMmaM, i0 = (blockIdx.z % iblocks) * kMmaM;
  const int shift = (K - 1 - j) * dil;
  const int ttiles = (T + kWgT - 1) / kWgT;
  const long long items = (long long)B * ttiles;
  float acc[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[nt][c] = 0.f;
  const uint32_t gs_u32 = smem_u32(gs), xs_u32 = smem_u32(xs);
  const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 4;      // A: (rows 0-7,k 0-3) (rows 8-15,k 0-3) (rows 0-7,k 4-7) (rows 8-15,k 4-7)
  const int b_row = (lane & 7) + (lane >> 4) * 8, b_col = ((lane >> 3) & 1) * 4;      // B: (n 0-7,k 0-3) (n 0-7,k 4-7) (n 8-15,k 0-3) (n 8-15,k 4-7)
  const bool vec_g = (T & 3) == 0 && (reinterpret_cast<uintptr_t>(gsrc) & 15) == 0;
  const bool vec_x = (T & 3) == 0 && (shift & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
  for (long long it = blockIdx.x; it < items; it += gridDim.x) {
    const int b = (int)(it / ttiles), t0 = (int)(it - (long long)b * ttiles) * kWgT;
    __syncthreads();
    for (int ch = warp; ch < kMmaM; ch += 4) {
      const int o = o0 + ch, i = i0 + ch;
      stage_row(gs + ch * kWgS, o < Cout ? gsrc + ((size_t)b * Cout + o) * T : nullptr, t0, kWgT, T, vec_g, lane);
      stage_row(xs + ch * kWgS, i < Cin ? x + ((size_t)b * Cin + i) * T : nullptr, t0 - shift, kWgT, T, vec_x, lane);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kWgT; kk += 8) {
      uint32_t af[4];
      ldsm_x4(af, gs_u32 + (uint32_t)((warp * 16 + a_row) * kWgS + kk + a_col) * 4u);
#pragma unroll
      for (int e = 0; e < 4; ++e) af[e] = tf32_bits(__uint_as_float(af[e]));
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        uint32_t bf[4];
        ldsm_x4(bf, xs_u32 + (uint32_t)((np * 16 + b_row) * kWgS + kk + b_col) * 4u);
#pragma unroll
        for (int e = 0; e < 4; ++e) bf[e] = tf32_bits(__uint_as_float(bf[e]));
        mma_tf32(acc[np * 2], af, bf[0], bf[1]);
        mma_tf32(acc[np * 2 + 1], af, bf[2], bf[3]);
      }
    }
  }
#pragma unroll
  for (int nt = 0; nt < 8; ++nt)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int o = o0 + warp * 16 + g + (c >> 1) * 8, i = i0 + nt * 8 + 2 * t4 + (c & 1);
      if (o < Cout && i < Cin && acc[nt][c] != 0.f) atomicAdd(dw + ((size_t)o * Cin + i) * K + j, acc[nt][c]);
    }
}

// Several taps per CTA: the g tile and ONE input tile with a (TAPS-1)*dil halo serve TAPS taps, so the operands cross L2 -> SM
// K/TAPS times instead of K times; the tiles arrive by cp.async into a two-stage ring, so the next item loads while this one is
// multiplied.  8 warps = 4 (o) x 2 (i halves); a warp holds TAPS x 16 x 32 outputs.  B fragments by ld.shared
// (tap offsets q*dil are arbitrary element offsets); the partial sums leave through shared memory in dW's own (i, j) order so the
// reductions of a warp fall into a few sectors.
constexpr int kWgThreads2 = 256;

template <int TAPS>
__global__ void __launch_bounds__(kWgThreads2) tcn_wgrad_mma_taps_kernel(const float* __restrict__ gsrc, const float* __restrict__ x,
                                                                         float* __restrict__ dw, int B, int Cin, int Cout, int T, int K,
                                                                         int dil, int iblocks, int XSW) {
  extern __shared__ __align__(16) float smem[];
  const int stage_floats = kMmaM * kWgS + kMmaM * XSW;      // one stage: gs [64 o][kWgS] then xs [64 i][XSW] (column c <-> time t0 - shift0 + c)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t4 = lane & 3;
  const int wo = warp & 3, wi = warp >> 2;
  const int j0 = blockIdx.y * TAPS, ntaps = (K - j0) < TAPS ? (K - j0) : TAPS;
  const int o0 = (blockIdx.z / iblocks) * kMmaM, i0 = (blockIdx.z % iblocks) * kMmaM;
  const int shift0 = (K - 1 - j0) * dil;                 // the group's largest look-back
  const int xcols = kWgT + (TAPS - 1) * dil;
  const int ttiles = (T + kWgT - 1) / kWgT;
  const long long items = (long long)B * ttiles;
  const bool vec_g = (T & 3) == 0 && (reinterpret_cast<uintptr_t>(gsrc) & 15) == 0;
  const bool vec_x = (T & 3) == 0 && (shift0 & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
  float acc[TAPS][4][4];
#pragma unroll
  for (int q = 0; q < TAPS; ++q)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[q][nt][c] = 0.f;
  const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 4;

  auto issue = [&](long long it, int buf) {               // one commit group per call (empty past the last item)
    if (it < items) {
      const int b = (int)(it / ttiles), t0 = (int)(it - (long long)b * ttiles) * kWgT;
      float* gs = smem + (size_t)buf * stage_floats;
      float* xs = gs + kMmaM * kWgS;
      for (int ch = warp; ch < kMmaM; ch += 8) {
        const int o = o0 + ch, i = i0 + ch;
        stage_row(gs + ch * kWgS, o < Cout ? gsrc + ((size_t)b * Cout + o) * T : nullptr, t0, kWgT, T, vec_g, lane);
        stage_row(xs + (size_t)ch * XSW, i < Cin ? x + ((size_t)b * Cin + i) * T : nullptr, t0 - shift0, xcols, T, vec_x, lane);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  issue(blockIdx.x, 0);
  int n = 0;
  for (long long it = blockIdx.x; it < items; it += gridDim.x, ++n) {
    issue(it + gridDim.x, (n + 1) & 1);                   // the next item streams in while this one is multiplied
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    const float* gs = smem + (size_t)(n & 1) * stage_floats;
    const float* xs = gs + kMmaM * kWgS;
    const uint32_t gs_u32 = smem_u32(gs);
#pragma unroll 2
    for (int kk = 0; kk < kWgT; kk += 8) {
      uint32_t af[4];
      ldsm_x4(af, gs_u32 + (uint32_t)((wo * 16 + a_row) * kWgS + kk + a_col) * 4u);
#pragma unroll
      for (int e = 0; e < 4; ++e) af[e] = tf32_bits(__uint_as_float(af[e]));
      const float* xb = xs + (size_t)(wi * 32 + g) * XSW + kk + t4;
#pragma unroll
      for (int q = 0; q < TAPS; ++q) {                 // taps past K (last group) multiply staged columns too; their sums are dropped
        const float* xq = xb + q * dil;
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          const float* p = xq + (size_t)(nt * 8) * XSW;
          mma_tf32(acc[q][nt], af, tf32_bits(p[0]), tf32_bits(p[4]));
        }
      }
    }
    __syncthreads();                                      // this buffer is refilled by the next iteration's issue
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  // reduction into dW through shared memory, 32 rows of o at a time: stage[o_l][i][q]
  float* stage = smem;
#pragma unroll 1
  for (int h = 0; h < 2; ++h) {
    __syncthreads();
    if ((wo >> 1) == h) {
#pragma unroll
      for (int q = 0; q < TAPS; ++q)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int ol = (wo & 1) * 16 + g + (c >> 1) * 8, i = wi * 32 + nt * 8 + 2 * t4 + (c & 1);
            stage[(ol * kMmaM + i) * TAPS + q] = acc[q][nt][c];
          }
    }
    __syncthreads();
    for (int idx = tid; idx < 32 * kMmaM * TAPS; idx += kWgThreads2) {
      const int q = idx % TAPS, i = (idx / TAPS) % kMmaM, ol = idx / (TAPS * kMmaM);
      const int o = o0 + h * 32 + ol;
      const float v = stage[idx];
      if (q < ntaps && o < Cout && i0 + i < Cin && v != 0.f) atomicAdd(dw + ((size_t)o * Cin + i0 + i) * K + j0 + q, v);
    }
  }
}

static inline int round_up_i(int v, int m) { return (v + m - 1) / m * m; }

// Picks KC / XS / NT for a problem; returns the dynamic shared-memory bytes, or 0 when the tensor-core form does not apply.
static size_t mma_plan(int IC, int T, int K, int dil, int* KC, int* XS, int* NT) {
  const int nt = T > 64 ? 4 : 2, TT = 4 * nt * 8;
  const long long halo = (long long)(K - 1) * dil;
  if (halo > 4096) return 0;
  int xs = TT + (int)halo;
  xs += ((8 - (xs % 32)) + 32) % 32;                 // == 8 mod 32
  const int icp8 = round_up_i(IC, 8);
  for (int kc = 64; kc >= 8; kc >>= 1) {
    if (kc > icp8 && kc > 8) continue;
    const size_t bytes = ((size_t)kSlabs * kMmaM * (kc + 4) + (size_t)kc * xs) * sizeof(float);
    if (bytes <= 200 * 1024) { *KC = kc; *XS = xs; *NT = nt; return bytes; }
  }
  return 0;
}

}  // namespace zoo

using namespace zoo;

/* 1 when the TF32 tensor-core kernels take this problem (else the CUDA-core kernels of tcn.cu run). */
ZOO_API int zoo_tcn_mma_supported(int Cin, int Cout, int T, int K, int dilation) {
  int kc, xs, nt;
  if (Cin < 16 || Cout < 16 || T < 32) return 0;      // tiny layers: padding to 64 x 8 tiles wastes the tensor pipe
  return mma_plan(Cin, T, K, dilation, &kc, &xs, &nt) && mma_plan(Cout, T, K, dilation, &kc, &xs, &nt) ? 1 : 0;
}

/* Floats of scratch the forward / dgrad entry points need for the re-laid-out weights. */
ZOO_API long long zoo_tcn_mma_scratch_elems(int Cin, int Cout, int K) {
  const long long a = (long long)K * round_up_i(Cout, kMmaM) * round_up_i(Cin, 64);
  const long long b = (long long)K * round_up_i(Cin, kMmaM) * round_up_i(Cout, 64);
  return a > b ? a : b;
}

static int launch_mma_taps(const float* in, const float* w, float* wt, long long s_m, long long s_k, const float* bias,
                           const float* res, float* out, int B, int IC, int OC, int T, int K, int dil, int sign, int relu,
                           float drop_p, unsigned long long seed, bool epi, cudaStream_t st) {
  MmaTapArgs a{};
  int nt = 0;
  const size_t sm = mma_plan(IC, T, K, dil, &a.KC, &a.XS, &nt);
  if (!sm) { zoo_set_last_error("tcn mma: problem does not fit the tensor-core kernel", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  a.in = in; a.wt = wt; a.bias = bias; a.res = res; a.out = out;
  a.B = B; a.IC = IC; a.OC = OC; a.T = T; a.K = K; a.dil = dil;
  a.Mp = round_up_i(OC, kMmaM); a.Kp = round_up_i(IC, a.KC);
  a.sign = sign; a.relu = relu; a.drop_p = drop_p; a.seed = seed;
  const long long n = (long long)K * a.Mp * a.Kp;
  tcn_prep_weights_kernel<<<(unsigned)((n + 255) / 256 > 592 ? 592 : (n + 255) / 256), 256, 0, st>>>(w, wt, s_m, s_k, 1, OC, IC, a.Mp, a.Kp, K);
  const int TT = 4 * nt * 8;
  dim3 grid((unsigned)((T + TT - 1) / TT), (unsigned)(a.Mp / kMmaM), (unsigned)B);
  cudaError_t e = cudaSuccess;
#define ZOO_TCN_LAUNCH(NTV, EPIV)                                                                                         \
  do {                                                                                                                    \
    e = cudaFuncSetAttribute(conv_taps_mma_kernel<NTV, EPIV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);      \
    if (e == cudaSuccess) conv_taps_mma_kernel<NTV, EPIV><<<grid, kMmaThreads, sm, st>>>(a);                              \
  } while (0)
  if (nt == 4) { if (epi) ZOO_TCN_LAUNCH(4, true); else ZOO_TCN_LAUNCH(4, false); }
  else         { if (epi) ZOO_TCN_LAUNCH(2, true); else ZOO_TCN_LAUNCH(2, false); }
#undef ZOO_TCN_LAUNCH
  if (e != cudaSuccess) { zoo_set_last_error(cudaGetErrorString(e), __FILE__, __LINE__); return ZOO_ERR_CUDA; }
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_tcn_conv_forward_mma(const float* x, const float* w, float* wt_scratch, const float* bias, const float* res,
                                     float* out, int B, int Cin, int Cout, int T, int K, int dilation, int relu, float drop_p,
                                     unsigned long long seed, void* stream) {
  if (!x || !w || !wt_scratch || !out || B <= 0 || Cin <= 0 || Cout <= 0 || T <= 0 || K <= 0 || dilation <= 0 || drop_p < 0.f || drop_p >= 1.f) {
    zoo_set_last_error("tcn_conv_forward_mma: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  // Wt[j][oc][ic] = w[oc][ic][j]
  return launch_mma_taps(x, w, wt_scratch, (long long)Cin * K, (long long)K, bias, res, out, B, Cin, Cout, T, K, dilation, -1, relu,
                         drop_p, seed, true, (cudaStream_t)stream);
}

ZOO_API int zoo_tcn_conv_dgrad_mma(const float* g, const float* w, float* wt_scratch, float* dx, int B, int Cin, int Cout, int T,
                                   int K, int dilation, void* stream) {
  if (!g || !w || !wt_scratch || !dx || B <= 0 || Cin <= 0 || Cout <= 0 || T <= 0 || K <= 0 || dilation <= 0) {
    zoo_set_last_error("tcn_conv_dgrad_mma: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  // roles swap: rows = ci, reduction = co;  Wt[j][ci][co] = w[co][ci][j]
  return launch_mma_taps(g, w, wt_scratch, (long long)K, (long long)Cin * K, nullptr, nullptr, dx, B, Cout, Cin, T, K, dilation, +1, 0,
                         0.f, 0ull, false, (cudaStream_t)stream);
}

/* dw[Cout,Cin,K] += ...  (dw zero-initialised by the caller) */
ZOO_API int zoo_tcn_conv_wgrad_mma(const float* g, const float* x, float* dw, int B, int Cin, int Cout, int T, int K, int dilation,
                                   void* stream) {
  if (!g || !x || !dw || B <= 0 || Cin <= 0 || Cout <= 0 || T <= 0 || K <= 0 || dilation <= 0) {
    zoo_set_last_error("tcn_conv_wgrad_mma: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const int ob = (Cout + kMmaM - 1) / kMmaM, ib = (Cin + kMmaM - 1) / kMmaM;
  const long long items = (long long)B * ((T + kWgT - 1) / kWgT);
  constexpr int TAPS = 4;
  int xsw = kWgT + (TAPS - 1) * dilation;
  xsw += ((4 - (xsw % 32)) + 32) % 32;               // == 4 mod 32: the B-fragment loads hit 32 distinct banks
  size_t sm = 2 * ((size_t)kMmaM * kWgS + (size_t)kMmaM * xsw) * sizeof(float);      // two stages
  if (sm < (size_t)32 * kMmaM * TAPS * sizeof(float)) sm = (size_t)32 * kMmaM * TAPS * sizeof(float);
  if (K >= 2 && sm <= 110 * 1024) {
    const int groups = (K + TAPS - 1) / TAPS;
    long long gx = (148LL * 2) / ((long long)groups * ob * ib);
    if (gx < 1) gx = 1;
    if (gx > items) gx = items;
    cudaError_t e = cudaFuncSetAttribute(tcn_wgrad_mma_taps_kernel<TAPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    if (e != cudaSuccess) { zoo_set_last_error(cudaGetErrorString(e), __FILE__, __LINE__); return ZOO_ERR_CUDA; }
    tcn_wgrad_mma_taps_kernel<TAPS><<<dim3((unsigned)gx, (unsigned)groups, (unsigned)(ob * ib)), kWgThreads2, sm, (cudaStream_t)stream>>>(
        g, x, dw, B, Cin, Cout, T, K, dilation, ib, xsw);
    ZOO_CUDA_CHECK(cudaGetLastError());
    return ZOO_OK;
  }
  long long gx = (148LL * 4) / ((long long)K * ob * ib);          // wide dilation: one tap per CTA, no halo
  if (gx < 1) gx = 1;
  if (gx > items) gx = items;
  tcn_wgrad_mma_kernel<<<dim3((unsigned)gx, (unsigned)K, (unsigned)(ob * ib)), kWg1Threads, 0, (cudaStream_t)stream>>>(
      g, x, dw, B, Cin, Cout, T, K, dilation, ib);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
