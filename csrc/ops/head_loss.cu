// K5b: the classification head in one pass -- last Dense layer + softmax + NLL + the head's whole backward.
//
// The last layer of the recommendation towers is skinny: NeuralCF ends in Dense(128 -> 5) + SoftMax + ClassNLL
// (zoo/.../models/recommendation/NeuralCF.scala:89-96; criterion ZooClassNLLCriterion.scala:48-121), Wide&Deep in Dense(-> 2)
// (WideAndDeep.scala:146-157).  As separate launches that head costs five kernels per step (forward GEMM, loss, dgrad, wgrad,
// plus the small fills around the loss), each one all fixed cost: a 65,536 x 128 x 5 GEMM is 84 MFLOP.  Here one streaming kernel
// reads the head's input X once and leaves everything the step needs:
//   logits = X W^T + b  ->  p = softmax(logits)  ->  loss += -log clip(p[t])            (forward)
//   dlogits = (p - onehot(t)) * scale                                                   (loss backward)
//   dX = dlogits W, with the ReLU mask of the columns of X that are ReLU outputs        (dgrad of the head)
//   dW += dlogits^T X,  db += sum dlogits                                               (wgrad of the head)
// X is read once (bf16), dX written once (bf16); W (C x K, C <= 8) sits in shared memory, dW is accumulated in registers and
// leaves through one shared-memory reduction per CTA.  A group of G lanes owns a row (8 columns per lane, one 16-byte load),
// two rows are in flight per group.
#include "../common/zoo_common.cuh"

namespace zoo {

struct HeadArgs {
  const __nv_bfloat16* x; long long ldx;     // [B, K] input of the head
  const __nv_bfloat16* w; long long ldw;     // [C, K] bf16 shadow of the weight
  const float* bias;                         // [C] or null
  const long long* labels;                   // [B]
  int B, K, C;
  int label_offset;                          // 0: zero-based labels, 1: one-based
  long long padding_value; int has_padding;  // rows whose label equals padding_value contribute nothing
  const float* valid_count;                  // device scalar: number of non-padding rows (mean divisor), or null -> B
  int size_average;
  float eps;
  int relu_cols;                             // columns [0, relu_cols) of X are ReLU outputs: dX is masked where X <= 0
  float* loss_out;                           // [1] += loss
  __nv_bfloat16* dx; long long ldd;          // [B, K]
  float* dw; long long lddw;                 // [C, K] fp32, accumulated
  float* db;                                 // [C] fp32, accumulated (or null)
  float* logits_out; long long ldl;          // optional [B, C] fp32
};

__global__ void __launch_bounds__(256)
count_valid_labels_kernel(const long long* __restrict__ labels, int B, long long padding_value, int label_offset, int C,
                          float* __restrict__ out) {
  pdl_wait();
  pdl_launch_dependents();
  float n = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < B; i += gridDim.x * blockDim.x) {
    const long long raw = labels[i];
    const long long t = raw - label_offset;
    n += (raw != padding_value && t >= 0 && t < C) ? 1.f : 0.f;
  }
  n = warp_sum(n);
  __shared__ float part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += part[i];
    if (s != 0.f) atomicAdd(out, s);
  }
}

// Packed fp32 FMA (FFMA2 on sm_100a): two lanes of a float2 per instruction -- the head is instruction-bound, not memory-bound.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b),
                     rc = *reinterpret_cast<unsigned long long*>(&c), rd;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  return *reinterpret_cast<float2*>(&rd);
}

// Shared-memory layout of a [C, KP] fp32 matrix whose lane l owns columns 8l .. 8l+7: the lane's first four columns sit at
// 4l and the other four at KP/2 + 4l of the row, so a 16-byte access of consecutive lanes touches consecutive banks.
template <int KP>
__device__ __forceinline__ int head_perm(int c, int k) { return c * KP + ((k >> 2) & 1) * (KP / 2) + (k >> 3) * 4 + (k & 3); }

template <int C, int G>
__global__ void __launch_bounds__(256, (C <= 5) ? 2 : 1)
head_softmax_nll_kernel(const HeadArgs a) {
  constexpr int KP = G * 8;              // padded K
  constexpr int R = 2;                   // rows in flight per group
  __shared__ __align__(16) float sW[C * KP];
  __shared__ __align__(16) float sdW[C * KP];
  __shared__ float sdb[8][8];
  __shared__ float sloss[8];
  const int tid = threadIdx.x;
  for (int i = tid; i < C * KP; i += 256) {
    const int c = i / KP, k = i - c * KP;
    // (the weights were written by the previous optimizer step, long before the predecessor kernel)
    sW[head_perm<KP>(c, k)] = k < a.K ? __bfloat162float(a.w[(long long)c * a.ldw + k]) : 0.f;
    sdW[i] = 0.f;
  }
  pdl_wait();
  pdl_launch_dependents();
  __syncthreads();

  const int l = tid % G;                                  // lane inside the row group
  const int col = l * 8;
  const bool col_ok = col < a.K;                          // K % 8 == 0 is required by the host wrapper
  const int group = (blockIdx.x * 256 + tid) / G;
  const int cta_group0 = blockIdx.x * (256 / G);          // first group of this CTA: the trip count below is uniform over the CTA
  const int ngroups = gridDim.x * 256 / G;
  float denom = (float)max(a.B, 1);
  if (a.valid_count) denom = fmaxf(*a.valid_count, 1.f);
  const float scale = a.size_average ? 1.f / denom : 1.f;
  float bias[C];
#pragma unroll
  for (int c = 0; c < C; ++c) bias[c] = a.bias ? a.bias[c] : 0.f;
  const float* w_lo = sW + 4 * l;                         // + c * KP: this lane's columns 8l .. 8l+3
  const float* w_hi = sW + KP / 2 + 4 * l;                //           and 8l+4 .. 8l+7

  float2 dwacc[C][4];
#pragma unroll
  for (int c = 0; c < C; ++c)
#pragma unroll
    for (int j = 0; j < 4; ++j) dwacc[c][j] = make_float2(0.f, 0.f);
  float dbacc[C];
#pragma unroll
  for (int c = 0; c < C; ++c) dbacc[c] = 0.f;
  float loss = 0.f;

  for (int it = 0; (cta_group0 + it * ngroups) * R < a.B; ++it) {      // rows past B are masked by row_ok (full-mask shuffles inside)
    const int row0 = (group + it * ngroups) * R;
    uint4 xv[R];
    long long lab[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {                          // issue the loads of both rows first
      const int row = row0 + r;
      xv[r] = make_uint4(0u, 0u, 0u, 0u);
      lab[r] = 0;
      if (row < a.B) {
        if (col_ok) xv[r] = ld_nc_u4(reinterpret_cast<const uint4*>(a.x + (long long)row * a.ldx + col));
        lab[r] = a.labels[row];
      }
    }
    float2 x[R][4];
    float lg[R][C];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      x[r][0] = unpack_bf16(xv[r].x); x[r][1] = unpack_bf16(xv[r].y); x[r][2] = unpack_bf16(xv[r].z); x[r][3] = unpack_bf16(xv[r].w);
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {                          // one weight read serves both rows
      const float4 w0 = *reinterpret_cast<const float4*>(w_lo + c * KP);
      const float4 w1 = *reinterpret_cast<const float4*>(w_hi + c * KP);
      const float2 wa = make_float2(w0.x, w0.y), wb = make_float2(w0.z, w0.w), wc = make_float2(w1.x, w1.y), wd = make_float2(w1.z, w1.w);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        float2 acc = make_float2(x[r][0].x * wa.x, x[r][0].y * wa.y);
        acc = ffma2(x[r][1], wb, acc);
        acc = ffma2(x[r][2], wc, acc);
        acc = ffma2(x[r][3], wd, acc);
        lg[r][c] = acc.x + acc.y;
      }
    }
#pragma unroll
    for (int off = G / 2; off >= 1; off >>= 1)
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < C; ++c) lg[r][c] += __shfl_xor_sync(0xffffffffu, lg[r][c], off);
    float dl[R][C];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = row0 + r;
      const bool row_ok = row < a.B;
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < C; ++c) { lg[r][c] += bias[c]; mx = fmaxf(mx, lg[r][c]); }
      if (a.logits_out && row_ok && l < C) {
#pragma unroll
        for (int c = 0; c < C; ++c) if (c == l) a.logits_out[(long long)row * a.ldl + c] = lg[r][c];
      }
      float p[C], sum = 0.f;
#pragma unroll
      for (int c = 0; c < C; ++c) { p[c] = __expf(lg[r][c] - mx); sum += p[c]; }
      const float inv = 1.f / sum;
      const long long t = lab[r] - a.label_offset;
      const bool valid = row_ok && !(a.has_padding && lab[r] == a.padding_value) && t >= 0 && t < C;
      float pt = 1.f;
#pragma unroll
      for (int c = 0; c < C; ++c) { p[c] *= inv; if (c == (int)t) pt = p[c]; }
      const bool clipped = (pt < a.eps) || (pt > 1.f - a.eps);
      pt = fminf(fmaxf(pt, a.eps), 1.f - a.eps);
      if (valid && l == 0) loss += -__logf(pt);
      const float gs = (valid && !clipped) ? scale : 0.f;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        dl[r][c] = gs * (p[c] - (c == (int)t ? 1.f : 0.f));
        dbacc[c] += dl[r][c];                              // every lane of the group holds the same value; lane 0's copy is used
      }
    }
    // head dgrad + wgrad for this lane's 8 columns
    float2 dxv[R][4];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int j = 0; j < 4; ++j) dxv[r][j] = make_float2(0.f, 0.f);
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const float4 w0 = *reinterpret_cast<const float4*>(w_lo + c * KP);
      const float4 w1 = *reinterpret_cast<const float4*>(w_hi + c * KP);
      const float2 wq[4] = {make_float2(w0.x, w0.y), make_float2(w0.z, w0.w), make_float2(w1.x, w1.y), make_float2(w1.z, w1.w)};
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float2 d2 = make_float2(dl[r][c], dl[r][c]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          dxv[r][j] = ffma2(d2, wq[j], dxv[r][j]);
          dwacc[c][j] = ffma2(d2, x[r][j], dwacc[c][j]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = row0 + r;
      if (col < a.relu_cols) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (col + 2 * j < a.relu_cols && !(x[r][j].x > 0.f)) dxv[r][j].x = 0.f;
          if (col + 2 * j + 1 < a.relu_cols && !(x[r][j].y > 0.f)) dxv[r][j].y = 0.f;
        }
      }
      if (row < a.B && col_ok && a.dx) {
        uint4 pk;
        pk.x = pack_bf16(dxv[r][0].x, dxv[r][0].y); pk.y = pack_bf16(dxv[r][1].x, dxv[r][1].y);
        pk.z = pack_bf16(dxv[r][2].x, dxv[r][2].y); pk.w = pack_bf16(dxv[r][3].x, dxv[r][3].y);
        *reinterpret_cast<uint4*>(a.dx + (long long)row * a.ldd + col) = pk;
      }
    }
  }

  // ---- CTA reduction of dW / db / loss without shared-memory atomics (fp32 shared atomics are compare-and-swap loops) ----------
  // 1. the groups of a warp hold the same columns: fold them with shuffles; 2. the eight warps add their sums into sdW one after
  // the other; 3. one global atomic per element per CTA.
#pragma unroll
  for (int off = G; off < 32; off <<= 1)
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        dwacc[c][j].x += __shfl_xor_sync(0xffffffffu, dwacc[c][j].x, off);
        dwacc[c][j].y += __shfl_xor_sync(0xffffffffu, dwacc[c][j].y, off);
      }
#pragma unroll
  for (int off = G; off < 32; off <<= 1)
#pragma unroll
    for (int c = 0; c < C; ++c) dbacc[c] += __shfl_xor_sync(0xffffffffu, dbacc[c], off);
  loss = warp_sum(loss);
  const int warp = tid >> 5, lane = tid & 31;
  if (lane == 0) {
    sloss[warp] = loss;
#pragma unroll
    for (int c = 0; c < C; ++c) sdb[warp][c] = dbacc[c];
  }
  for (int wv = 0; wv < 8; ++wv) {
    if (warp == wv && lane < G && col_ok) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        float4* lo = reinterpret_cast<float4*>(sdW + c * KP + 4 * l);
        float4* hi = reinterpret_cast<float4*>(sdW + c * KP + KP / 2 + 4 * l);
        float4 v0 = *lo, v1 = *hi;
        v0.x += dwacc[c][0].x; v0.y += dwacc[c][0].y; v0.z += dwacc[c][1].x; v0.w += dwacc[c][1].y;
        v1.x += dwacc[c][2].x; v1.y += dwacc[c][2].y; v1.z += dwacc[c][3].x; v1.w += dwacc[c][3].y;
        *lo = v0; *hi = v1;
      }
    }
    __syncthreads();
  }
  for (int i = tid; i < C * KP; i += 256) {
    const int c = i / KP, k = i - c * KP;
    if (k < a.K && a.dw) atomicAdd(a.dw + (long long)c * a.lddw + k, sdW[head_perm<KP>(c, k)]);
  }
  if (tid < C && a.db) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += sdb[i][tid];
    atomicAdd(a.db + tid, s);
  }
  if (tid == 0 && a.loss_out) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += sloss[i];
    atomicAdd(a.loss_out, s * scale);
  }
}

// ---------------------------------------------------------------------------------------------------------------------------
// Tensor-core form.  The SIMT kernel above is instruction-bound (14.7 M warp instructions for 65,536 x 128 x 5: the dot products
// are spread over 16 lanes per row, so every row pays a 4-step shuffle reduction and 16 redundant softmaxes).  Here a warp owns a
// 16-row tile of X in shared memory and the three small GEMMs of the head run on the tensor cores through warp-level MMA
// (m16n8k16, bf16 in, fp32 accumulate -- 84 MFLOP in total, so the issue rate of `mma.sync` is irrelevant; what matters is that
// the per-row instruction count drops ~5x and the kernel becomes a memory stream):
//   logits[16, 8]   = X[16, K]  W^T[K, 8]            A: ldmatrix of the X tile; B: W fragments held in registers
//   dW^T[K, 8]     += X^T[K, 16] dlogits[16, 8]       A: ldmatrix.trans of the same tile; B: dlogits through movmatrix.trans
//   dX[16, K]       = dlogits[16, 8(16)] W[8(16), K]  A: the logits accumulator layout IS the A layout; B: W fragments in registers
// The softmax of a row lives in the 4 lanes of a quad (2 classes each).  dlogits are rounded to bf16 for the two backward GEMMs,
// like the layer-by-layer path (whose dlogits tensor is bf16).
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t v) {
  uint32_t o;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(o) : "r"(v));
  return o;
}
__device__ __forceinline__ float quad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

template <int KT>                        // K padded to KT * 16 columns
__global__ void __launch_bounds__(256, KT <= 8 ? 2 : 1)
head_softmax_nll_mma_kernel(const HeadArgs a) {
  constexpr int KP = KT * 16;
  constexpr int STRIDE = KP + 8;         // elements; (STRIDE * 2 / 16) is odd: ldmatrix rows fall on distinct 16-byte bank groups
  constexpr int CPR = KP / 8;            // 16-byte chunks per row
  constexpr int NLD = 16 * CPR / 32;     // chunks per lane per tile
  extern __shared__ __align__(16) uint8_t head_smem[];
  __shared__ float sdb[8][8];
  __shared__ float sloss[8];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int qr = lane >> 2, qc = lane & 3;           // quad row (0..7) and position inside the quad
  const int n0 = qc * 2, n1 = n0 + 1;                // the two classes this lane holds
  __nv_bfloat16* sx = reinterpret_cast<__nv_bfloat16*>(head_smem) + warp * 16 * STRIDE;
  const uint32_t sx_u32 = smem_u32(sx);
  const int C = a.C;

  // ---- W fragments (the weights were written by the previous optimizer step, long before the predecessor kernel) ----
  uint32_t wb1[KT][2];                   // logits GEMM, k-tile kk: B[k][n = qr] = W[qr][kk*16 + 2 qc (+8) ..+1]
  uint32_t wb2[2 * KT];                  // dX GEMM, n-tile j: B[k = class][n] = (W[n0][8j + qr], W[n1][8j + qr])
  const uint16_t* w16 = reinterpret_cast<const uint16_t*>(a.w);
#pragma unroll
  for (int kk = 0; kk < KT; ++kk) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = kk * 16 + h * 8 + qc * 2;
      uint32_t v = 0u;
      if (qr < C && k < a.K) v = *reinterpret_cast<const uint32_t*>(w16 + (long long)qr * a.ldw + k);      // K % 8 == 0: k + 1 < K too
      wb1[kk][h] = v;
    }
  }
  // the 8 x 8 block (classes x columns 8j .. 8j+7) of W that wb1 holds row-wise is what the dX GEMM needs column-wise
#pragma unroll
  for (int j = 0; j < 2 * KT; ++j) wb2[j] = movmatrix_trans(wb1[j >> 1][j & 1]);
  const float bias0 = (a.bias && n0 < C) ? a.bias[n0] : 0.f, bias1 = (a.bias && n1 < C) ? a.bias[n1] : 0.f;
  pdl_wait();
  pdl_launch_dependents();

  float denom = (float)max(a.B, 1);
  if (a.valid_count) denom = fmaxf(*a.valid_count, 1.f);
  const float scale = a.size_average ? 1.f / denom : 1.f;

  float acc[KT][4];                      // dW^T: [col = 16 mi + qr (+8)][class n0, n1]
#pragma unroll
  for (int mi = 0; mi < KT; ++mi) { acc[mi][0] = acc[mi][1] = acc[mi][2] = acc[mi][3] = 0.f; }
  float db0 = 0.f, db1 = 0.f, loss = 0.f;

  const int ntiles = (a.B + 15) >> 4;
  const int nwarps = gridDim.x * 8;
  for (int tile = blockIdx.x * 8 + warp; tile < ntiles; tile += nwarps) {
    const int row0 = tile << 4;
    // ---- 1. X tile -> shared memory (coalesced 16-byte chunks; rows / columns past the matrix are zero) ----
    uint4 xv[NLD];
#pragma unroll
    for (int i = 0; i < NLD; ++i) {
      const int q = i * 32 + lane, r = q / CPR, ch = q - r * CPR;
      xv[i] = make_uint4(0u, 0u, 0u, 0u);
      if (row0 + r < a.B && ch * 8 < a.K) xv[i] = ld_nc_u4(reinterpret_cast<const uint4*>(a.x + (long long)(row0 + r) * a.ldx + ch * 8));
    }
    const int rA = row0 + qr, rB = rA + 8;
    long long tA = 0, tB = 0;
    if (rA < a.B) tA = a.labels[rA];
    if (rB < a.B) tB = a.labels[rB];
#pragma unroll
    for (int i = 0; i < NLD; ++i) {
      const int q = i * 32 + lane, r = q / CPR, ch = q - r * CPR;
      *reinterpret_cast<uint4*>(sx + r * STRIDE + ch * 8) = xv[i];
    }
    __syncwarp();
    // ---- 2. logits ----
    float lg[4] = {0.f, 0.f, 0.f, 0.f};
    {
      const uint32_t base = sx_u32 + (uint32_t)(((lane & 7) + 8 * ((lane >> 3) & 1)) * STRIDE + 8 * (lane >> 4)) * 2u;
#pragma unroll
      for (int kk = 0; kk < KT; ++kk) {
        uint32_t af[4];
        ldmatrix_x4(af, base + kk * 32);
        mma_16816(lg, af, wb1[kk][0], wb1[kk][1]);
      }
    }
    // ---- 3. softmax + NLL per row, inside the quad ----
    float dl[4];
    {
      float v[4] = {lg[0] + bias0, lg[1] + bias1, lg[2] + bias0, lg[3] + bias1};
      if (a.logits_out) {
        if (rA < a.B) { if (n0 < C) a.logits_out[(long long)rA * a.ldl + n0] = v[0]; if (n1 < C) a.logits_out[(long long)rA * a.ldl + n1] = v[1]; }
        if (rB < a.B) { if (n0 < C) a.logits_out[(long long)rB * a.ldl + n0] = v[2]; if (n1 < C) a.logits_out[(long long)rB * a.ldl + n1] = v[3]; }
      }
      if (n0 >= C) { v[0] = -INFINITY; v[2] = -INFINITY; }
      if (n1 >= C) { v[1] = -INFINITY; v[3] = -INFINITY; }
      const float mA = quad_max(fmaxf(v[0], v[1])), mB = quad_max(fmaxf(v[2], v[3]));
      float e[4] = {__expf(v[0] - mA), __expf(v[1] - mA), __expf(v[2] - mB), __expf(v[3] - mB)};
      const float iA = 1.f / quad_sum(e[0] + e[1]), iB = 1.f / quad_sum(e[2] + e[3]);
      e[0] *= iA; e[1] *= iA; e[2] *= iB; e[3] *= iB;
      const long long cA = tA - a.label_offset, cB = tB - a.label_offset;
      const bool okA = rA < a.B && !(a.has_padding && tA == a.padding_value) && cA >= 0 && cA < C;
      const bool okB = rB < a.B && !(a.has_padding && tB == a.padding_value) && cB >= 0 && cB < C;
      const float hA0 = (okA && cA == n0) ? 1.f : 0.f, hA1 = (okA && cA == n1) ? 1.f : 0.f;
      const float hB0 = (okB && cB == n0) ? 1.f : 0.f, hB1 = (okB && cB == n1) ? 1.f : 0.f;
      float ptA = quad_sum(hA0 * e[0] + hA1 * e[1]), ptB = quad_sum(hB0 * e[2] + hB1 * e[3]);      // p[label], known to the whole quad
      const bool clipA = (ptA < a.eps) || (ptA > 1.f - a.eps), clipB = (ptB < a.eps) || (ptB > 1.f - a.eps);
      ptA = fminf(fmaxf(ptA, a.eps), 1.f - a.eps);
      ptB = fminf(fmaxf(ptB, a.eps), 1.f - a.eps);
      if (qc == 0) {
        if (okA) loss += -__logf(ptA);
        if (okB) loss += -__logf(ptB);
      }
      const float gA = (okA && !clipA) ? scale : 0.f, gB = (okB && !clipB) ? scale : 0.f;
      dl[0] = gA * (e[0] - hA0); dl[1] = gA * (e[1] - hA1); dl[2] = gB * (e[2] - hB0); dl[3] = gB * (e[3] - hB1);
      if (n0 >= C) { dl[0] = 0.f; dl[2] = 0.f; }
      if (n1 >= C) { dl[1] = 0.f; dl[3] = 0.f; }
      db0 += dl[0] + dl[2];
      db1 += dl[1] + dl[3];
    }
    const uint32_t pA = pack_bf16(dl[0], dl[1]), pB = pack_bf16(dl[2], dl[3]);
    // ---- 4. dW^T += X^T dlogits ----
    {
      const uint32_t b0 = movmatrix_trans(pA), b1 = movmatrix_trans(pB);
      const uint32_t base = sx_u32 + (uint32_t)(((lane & 7) + 8 * (lane >> 4)) * STRIDE + 8 * ((lane >> 3) & 1)) * 2u;
#pragma unroll
      for (int mi = 0; mi < KT; ++mi) {
        uint32_t af[4];
        ldmatrix_x4_trans(af, base + mi * 32);
        mma_16816(acc[mi], af, b0, b1);
      }
    }
    __syncwarp();
    // ---- 5. dX = dlogits W, ReLU mask from the X values, written over the X tile ----
    {
      const uint32_t af[4] = {pA, pB, 0u, 0u};
      __nv_bfloat16* rowA = sx + qr * STRIDE + qc * 2;
      __nv_bfloat16* rowB = rowA + 8 * STRIDE;
#pragma unroll
      for (int j = 0; j < 2 * KT; ++j) {
        float d[4] = {0.f, 0.f, 0.f, 0.f};
        mma_16816(d, af, wb2[j], 0u);
        const int c = 8 * j + qc * 2;
        if (c < a.relu_cols) {               // X is a ReLU output here: non-negative, so "x > 0" is "bits != 0"
          const uint32_t xa = *reinterpret_cast<const uint32_t*>(rowA + 8 * j), xb = *reinterpret_cast<const uint32_t*>(rowB + 8 * j);
          if (!(xa & 0x7fffu)) d[0] = 0.f;
          if (!(xb & 0x7fffu)) d[2] = 0.f;
          if (c + 1 < a.relu_cols) {
            if (!(xa & 0x7fff0000u)) d[1] = 0.f;
            if (!(xb & 0x7fff0000u)) d[3] = 0.f;
          }
        }
        *reinterpret_cast<uint32_t*>(rowA + 8 * j) = pack_bf16(d[0], d[1]);
        *reinterpret_cast<uint32_t*>(rowB + 8 * j) = pack_bf16(d[2], d[3]);
      }
    }
    __syncwarp();
    // ---- 6. dX tile -> global (the chunk mapping of step 1) ----
    if (a.dx) {
#pragma unroll
      for (int i = 0; i < NLD; ++i) {
        const int q = i * 32 + lane, r = q / CPR, ch = q - r * CPR;
        if (row0 + r < a.B && ch * 8 < a.K)
          *reinterpret_cast<uint4*>(a.dx + (long long)(row0 + r) * a.ldd + ch * 8) = *reinterpret_cast<const uint4*>(sx + r * STRIDE + ch * 8);
      }
    }
    __syncwarp();
  }

  // ---- CTA reduction: the eight warps add their dW^T into shared memory one after the other; one global atomic per element ----
  db0 += __shfl_xor_sync(0xffffffffu, db0, 4); db0 += __shfl_xor_sync(0xffffffffu, db0, 8); db0 += __shfl_xor_sync(0xffffffffu, db0, 16);
  db1 += __shfl_xor_sync(0xffffffffu, db1, 4); db1 += __shfl_xor_sync(0xffffffffu, db1, 8); db1 += __shfl_xor_sync(0xffffffffu, db1, 16);
  loss = warp_sum(loss);
  if (lane < 4) { sdb[warp][n0] = db0; sdb[warp][n1] = db1; }
  if (lane == 0) sloss[warp] = loss;
  // every warp parks its dW^T in its own tile memory ([8 classes][KP] fp32 fits: 8 KP 4 <= 16 (KP + 8) 2), then all threads
  // add the eight copies and issue one global atomic per element
  {
    float* mine = reinterpret_cast<float*>(sx);
#pragma unroll
    for (int mi = 0; mi < KT; ++mi) {
      const int col = mi * 16 + qr;
      mine[n0 * KP + col] = acc[mi][0];
      mine[n1 * KP + col] = acc[mi][1];
      mine[n0 * KP + col + 8] = acc[mi][2];
      mine[n1 * KP + col + 8] = acc[mi][3];
    }
  }
  __syncthreads();
  if (a.dw) {
    const float* base = reinterpret_cast<const float*>(head_smem);
    constexpr int WSTRIDE = 16 * STRIDE / 2;                 // floats between the tile memories of consecutive warps
    for (int i = tid; i < C * KP; i += 256) {
      const int c = i / KP, k = i - c * KP;
      if (k < a.K) {
        float v = 0.f;
#pragma unroll
        for (int wv = 0; wv < 8; ++wv) v += base[wv * WSTRIDE + i];
        atomicAdd(a.dw + (long long)c * a.lddw + k, v);
      }
    }
  }
  if (tid < C && a.db) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += sdb[i][tid];
    atomicAdd(a.db + tid, s);
  }
  if (tid == 0 && a.loss_out) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += sloss[i];
    atomicAdd(a.loss_out, s * scale);
  }
}

template <int KT>
static int launch_head_mma(const HeadArgs& a, cudaStream_t s) {
  constexpr int KP = KT * 16;
  constexpr int smem = 8 * 16 * (KP + 8) * 2 > 8 * KP * 4 ? 8 * 16 * (KP + 8) * 2 : 8 * KP * 4;
  static bool attr_set = false;
  if (!attr_set) {
    ZOO_CUDA_CHECK(cudaFuncSetAttribute(head_softmax_nll_mma_kernel<KT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  const int ntiles = (a.B + 15) / 16;
  int grid = (ntiles + 7) / 8;
  const int cap = (KT <= 8 ? 2 : 1) * 148;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  zoo_launch_pdl(head_softmax_nll_mma_kernel<KT>, dim3(grid), dim3(256), (size_t)smem, s, a);
  return ZOO_OK;
}

template <int C>
static int launch_head(const HeadArgs& a, int grid, cudaStream_t s) {
  if (a.K <= 64) zoo_launch_pdl(head_softmax_nll_kernel<C, 8>, dim3(grid), dim3(256), 0, s, a);
  else if (a.K <= 128) zoo_launch_pdl(head_softmax_nll_kernel<C, 16>, dim3(grid), dim3(256), 0, s, a);
  else zoo_launch_pdl(head_softmax_nll_kernel<C, 32>, dim3(grid), dim3(256), 0, s, a);
  return ZOO_OK;
}

}  // namespace zoo

// Head of a classifier in one pass (see the top of this file).  x: [B, K] bf16 (K % 8 == 0, K <= 256, 16-byte aligned rows);
// w: [C, K] bf16 (C <= 8); labels: int64 [B].  loss_out[0] += loss; dx: [B, K] bf16; dw (fp32 [C, K]) and db (fp32 [C]) are
// ACCUMULATED.  padding: has_padding != 0 makes rows whose label equals padding_value contribute nothing, and the mean divisor
// the number of remaining rows -- counted here first into valid_scratch[0] (a zeroed device float the caller provides).
// impl: 0 = warp-MMA kernel (default), 1 = SIMT kernel.
ZOO_API int zoo_head_softmax_nll(const void* x, long long ldx, const void* w, long long ldw, const float* bias,
                                 const long long* labels, int B, int K, int C, int zero_based, long long padding_value,
                                 int has_padding, int size_average, float eps, int relu_cols, float* loss_out, void* dx,
                                 long long ldd, float* dw, long long lddw, float* db, float* logits_out, long long ldl,
                                 float* valid_scratch, int impl, void* stream) {
  using namespace zoo;
  if (!x || !w || !labels || B < 0 || K <= 0 || K > 256 || (K & 7) || C <= 0 || C > 8 || (ldx & 7) || (dx && (ldd & 7))) {
    zoo_set_last_error("head_softmax_nll: bad arguments (K % 8 == 0, K <= 256, C <= 8, 16-byte rows)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (has_padding && size_average && !valid_scratch) {
    zoo_set_last_error("head_softmax_nll: padding with size_average needs valid_scratch", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (B == 0) return ZOO_OK;
  cudaStream_t s = (cudaStream_t)stream;
  HeadArgs a;
  a.x = (const __nv_bfloat16*)x; a.ldx = ldx; a.w = (const __nv_bfloat16*)w; a.ldw = ldw; a.bias = bias; a.labels = labels;
  a.B = B; a.K = K; a.C = C; a.label_offset = zero_based ? 0 : 1; a.padding_value = padding_value; a.has_padding = has_padding;
  a.valid_count = nullptr; a.size_average = size_average; a.eps = eps; a.relu_cols = relu_cols; a.loss_out = loss_out;
  a.dx = (__nv_bfloat16*)dx; a.ldd = ldd; a.dw = dw; a.lddw = lddw; a.db = db; a.logits_out = logits_out; a.ldl = ldl;
  if (has_padding && size_average) {
    const int cb = (B + 255) / 256 < 148 ? (B + 255) / 256 : 148;
    zoo_launch_pdl(count_valid_labels_kernel, dim3(cb), dim3(256), 0, s, labels, B, padding_value, a.label_offset, C, valid_scratch);
    a.valid_count = valid_scratch;
  }
  if (impl == 0) {       // tensor-core form; impl == 1: the SIMT kernel (second implementation, kept for the tests)
    int rc;
    if (K <= 64) rc = launch_head_mma<4>(a, s);
    else if (K <= 128) rc = launch_head_mma<8>(a, s);
    else rc = launch_head_mma<16>(a, s);
    if (rc != ZOO_OK) return rc;
    ZOO_CUDA_CHECK(cudaGetLastError());
    return ZOO_OK;
  }
  const int G = K <= 64 ? 8 : (K <= 128 ? 16 : 32);
  const long long groups_needed = ((long long)B + 1) / 2;
  long long want = (groups_needed * G + 255) / 256;
  int grid = (int)(want < 2 * 148 ? want : 2 * 148);
  if (grid < 1) grid = 1;
  switch (C) {
    case 1: launch_head<1>(a, grid, s); break;
    case 2: launch_head<2>(a, grid, s); break;
    case 3: launch_head<3>(a, grid, s); break;
    case 4: launch_head<4>(a, grid, s); break;
    case 5: launch_head<5>(a, grid, s); break;
    case 6: launch_head<6>(a, grid, s); break;
    case 7: launch_head<7>(a, grid, s); break;
    default: launch_head<8>(a, grid, s); break;
  }
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
