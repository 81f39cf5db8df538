// K6 -- persistent recurrent kernels: LSTM and GRU, forward and backward-through-time.
//
// Replaces the per-timestep cell loops of the reference's recurrent layers
// (zoo/.../pipeline/api/keras/layers/LSTM.scala, GRU.scala, InternalRecurrent.scala) and of its
// PyTorch AutoML models (pyzoo/zoo/automl/model/VanillaLSTM_pytorch.py:29-56,
// Seq2Seq_pytorch.py:31-101).  The shapes are small (H 8..128, T <= a few hundred, B <= a few
// thousand), so a step is latency-bound: launching one kernel per time step (what a cell loop or a
// library RNN does) costs more than the math.  Here ONE launch runs all T steps.
//
// Decomposition.  The recurrence is independent across batch rows, so a CTA owns RPT (=4) batch
// rows for the whole sequence and never talks to another CTA:
//   * W_hh lives in shared memory for the whole launch (fp32, or bf16 when 4H*H*4 B does not fit);
//   * h_{t-1} of the CTA's rows lives in shared memory, c_t in a register of the thread that owns
//     (row, unit);
//   * the input projection x_t W_ih^T + b for ALL steps is one big GEMM done beforehand (it has no
//     sequential dependence); this kernel reads that buffer one step ahead (register prefetch) and
//     overwrites it in place with the activated gates, which is exactly what backward needs.
// Thread layout: blockDim = round_up(4H, 32).  Phase 1 (thread = gate column j): 4 dot products of
// length H against the h tile.  Phase 2 (thread = (row, unit)): cell update.  Backward mirrors it:
// phase A (row, unit) turns dh/dc into pre-activation gate gradients (written in place over the
// saved gates), phase B (row, k) multiplies them by W_hh to get dh_{t-1}.
// The weight/input gradients (sums over B*T) are plain GEMMs over the dG buffer, done by the caller.
#include "../common/zoo_common.cuh"

namespace zoo {

constexpr int kRnnRows = 4;  // batch rows per CTA

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }
// gate activation and its derivative expressed through the activated value v
__device__ __forceinline__ float gate_act(float x, int hard) { return hard ? fminf(fmaxf(fmaf(0.2f, x, 0.5f), 0.f), 1.f) : sigmoidf_(x); }
__device__ __forceinline__ float gate_der(float v, int hard) { return hard ? ((v > 0.f && v < 1.f) ? 0.2f : 0.f) : v * (1.f - v); }
__device__ __forceinline__ float tanhf_(float x) {
  // tanh(x) = 2*sigmoid(2x) - 1; exact enough in fp32 and NaN-free for large |x|
  return 2.f / (1.f + __expf(-2.f * x)) - 1.f;
}
template <typename WT> __device__ __forceinline__ float wload(const WT* p);
template <> __device__ __forceinline__ float wload<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float wload<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename WT> __device__ __forceinline__ WT wstore(float v);
template <> __device__ __forceinline__ float wstore<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 wstore<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

struct RnnArgs {
  float* gates;        // LSTM: [B,T,4H] in: x-projection (+biases), out: activated i,f,g,o
                       // GRU : [B,T,3H] in: x-projection (+b_ih, +b_hh for r,z) (read only)
  float* saved;        // GRU : [B,T,4H] out: r,z,n,hn (may be null at inference); LSTM: unused
  const float* whh;    // [G*H, H] row-major (G = 4 LSTM, 3 GRU)
  const float* bhn;    // GRU: [H] recurrent bias of the candidate gate (may be null)
  const float* h0;     // [B,H] or null
  const float* c0;     // [B,H] or null (LSTM)
  float* hout;         // [B,T,H]
  float* cout;         // [B,T,H] or null (LSTM; needed for backward)
  float* hT;           // [B,H] or null
  float* cT;           // [B,H] or null
  int B, T, H;
  int save;            // LSTM: write activated gates back into `gates`
  int hard;            // inner (gate) activation: 0 sigmoid (torch), 1 hard_sigmoid (the zoo keras default)
};

// ---------------------------------------------------------------- LSTM forward --
template <typename WT>
__global__ void lstm_fwd_kernel(RnnArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 4 * H, T = a.T;
  const int Hp = (H + 3) & ~3;                 // h tile rows padded to float4
  WT* Wt = reinterpret_cast<WT*>(smem_raw);    // [H][NC]  (transposed: column j contiguous over j)
  float* hs = reinterpret_cast<float*>(smem_raw + ((size_t)H * NC * sizeof(WT) + 15) / 16 * 16);  // [4][Hp]
  float* gs = hs + kRnnRows * Hp;              // [4][NC]
  const int tid = threadIdx.x;
  const int b0 = blockIdx.x * kRnnRows;
  for (int i = tid; i < NC * H; i += blockDim.x) {
    const int j = i / H, k = i - j * H;        // whh[j][k] -> Wt[k][j]
    Wt[(size_t)k * NC + j] = wstore<WT>(a.whh[i]);
  }
  for (int i = tid; i < kRnnRows * Hp; i += blockDim.x) {
    const int r = i / Hp, k = i - r * Hp, b = b0 + r;
    hs[i] = (a.h0 && b < a.B && k < H) ? a.h0[(size_t)b * H + k] : 0.f;
  }
  // phase-2 ownership: thread -> (row r2, unit u2); c lives in a register
  const int r2 = tid / H, u2 = tid - r2 * H;
  const bool own = tid < kRnnRows * H && (b0 + r2) < a.B;
  float c = (own && a.c0) ? a.c0[(size_t)(b0 + r2) * H + u2] : 0.f;
  const int j = tid;
  const bool col = j < NC;
  const bool is_g = col && j >= 2 * H && j < 3 * H;
  float gx[kRnnRows];
#pragma unroll
  for (int r = 0; r < kRnnRows; ++r) gx[r] = (col && b0 + r < a.B) ? a.gates[((size_t)(b0 + r) * T) * NC + j] : 0.f;
  __syncthreads();
  for (int t = 0; t < T; ++t) {
    float acc[kRnnRows];
#pragma unroll
    for (int r = 0; r < kRnnRows; ++r) acc[r] = gx[r];
    if (t + 1 < T) {                            // prefetch the next step's x-projection
#pragma unroll
      for (int r = 0; r < kRnnRows; ++r)
        gx[r] = (col && b0 + r < a.B) ? a.gates[((size_t)(b0 + r) * T + t + 1) * NC + j] : 0.f;
    }
    if (col) {
      const WT* wp = Wt + j;
      int k = 0;
      for (; k + 4 <= H; k += 4) {
        const float w0 = wload<WT>(wp + (size_t)k * NC), w1 = wload<WT>(wp + (size_t)(k + 1) * NC);
        const float w2 = wload<WT>(wp + (size_t)(k + 2) * NC), w3 = wload<WT>(wp + (size_t)(k + 3) * NC);
#pragma unroll
        for (int r = 0; r < kRnnRows; ++r) {
          const float4 h4 = *reinterpret_cast<const float4*>(hs + r * Hp + k);
          acc[r] = fmaf(w0, h4.x, fmaf(w1, h4.y, fmaf(w2, h4.z, fmaf(w3, h4.w, acc[r]))));
        }
      }
      for (; k < H; ++k) {
        const float w = wload<WT>(wp + (size_t)k * NC);
#pragma unroll
        for (int r = 0; r < kRnnRows; ++r) acc[r] = fmaf(w, hs[r * Hp + k], acc[r]);
      }
#pragma unroll
      for (int r = 0; r < kRnnRows; ++r) {
        const float v = is_g ? tanhf_(acc[r]) : gate_act(acc[r], a.hard);
        gs[r * NC + j] = v;
        if (a.save && b0 + r < a.B) a.gates[((size_t)(b0 + r) * T + t) * NC + j] = v;
      }
    }
    __syncthreads();
    if (own) {
      const float* g = gs + r2 * NC;
      const float ig = g[u2], fg = g[H + u2], gg = g[2 * H + u2], og = g[3 * H + u2];
      c = fmaf(fg, c, ig * gg);
      const float h = og * tanhf_(c);
      hs[r2 * Hp + u2] = h;
      const size_t o = ((size_t)(b0 + r2) * T + t) * H + u2;
      a.hout[o] = h;
      if (a.cout) a.cout[o] = c;
    }
    __syncthreads();
  }
  if (own) {
    const size_t o = (size_t)(b0 + r2) * H + u2;
    if (a.hT) a.hT[o] = hs[r2 * Hp + u2];
    if (a.cT) a.cT[o] = c;
  }
}

struct RnnBwdArgs {
  float* gates;        // LSTM: [B,T,4H] in: activated gates, out: pre-activation gate grads (in place)
                       // GRU : [B,T,4H] in: saved r,z,n,hn; out: dpr,dpz,dpn,dhn (in place)
  const float* whh;    // [G*H, H]
  const float* hout;   // [B,T,H] forward outputs (GRU needs h_{t-1})
  const float* cout;   // [B,T,H] (LSTM)
  const float* h0;     // [B,H] or null
  const float* c0;     // [B,H] or null
  const float* dy;     // [B,T,H] or null (gradient w.r.t. every output step)
  long long dy_sb, dy_st;  // strides of dy in elements (batch, time); unit stride 1
  const float* dhT;    // [B,H] or null
  const float* dcT;    // [B,H] or null
  float* dh0;          // [B,H] or null
  float* dc0;          // [B,H] or null
  int B, T, H;
  int hard;
};

// --------------------------------------------------------------- LSTM backward --
template <typename WT>
__global__ void lstm_bwd_kernel(RnnBwdArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 4 * H, T = a.T;
  WT* W = reinterpret_cast<WT*>(smem_raw);                 // [NC][H] natural layout
  float* dgs = reinterpret_cast<float*>(smem_raw + ((size_t)H * NC * sizeof(WT) + 15) / 16 * 16);  // [4][NC]
  float* dhs = dgs + kRnnRows * NC;                         // [4][H]
  const int tid = threadIdx.x, b0 = blockIdx.x * kRnnRows;
  for (int i = tid; i < NC * H; i += blockDim.x) W[i] = wstore<WT>(a.whh[i]);
  const int r = tid / H, u = tid - r * H, b = b0 + r;
  const bool own = tid < kRnnRows * H && b < a.B;
  if (tid < kRnnRows * H) dhs[r * H + u] = (own && a.dhT) ? a.dhT[(size_t)b * H + u] : 0.f;
  float dc_next = (own && a.dcT) ? a.dcT[(size_t)b * H + u] : 0.f;
  // registers for the current step (prefetched one step ahead)
  float ig = 0, fg = 0, gg = 0, og = 0, c = 0, cprev = 0, dyv = 0;
  auto load_step = [&](int t) {
    if (!own) return;
    const size_t gb = ((size_t)b * T + t) * NC;
    ig = a.gates[gb + u]; fg = a.gates[gb + H + u]; gg = a.gates[gb + 2 * H + u]; og = a.gates[gb + 3 * H + u];
    c = a.cout[((size_t)b * T + t) * H + u];
    cprev = t > 0 ? a.cout[((size_t)b * T + t - 1) * H + u] : (a.c0 ? a.c0[(size_t)b * H + u] : 0.f);
    dyv = a.dy ? a.dy[(size_t)b * a.dy_sb + (size_t)t * a.dy_st + u] : 0.f;
  };
  load_step(T - 1);
  __syncthreads();
  for (int t = T - 1; t >= 0; --t) {
    if (own) {
      const float dh = dyv + dhs[r * H + u];
      const float tc = tanhf_(c);
      const float dc = fmaf(dh * og, 1.f - tc * tc, dc_next);
      const float dpi = dc * gg * gate_der(ig, a.hard);
      const float dpf = dc * cprev * gate_der(fg, a.hard);
      const float dpg = dc * ig * (1.f - gg * gg);
      const float dpo = dh * tc * gate_der(og, a.hard);
      dc_next = dc * fg;
      float* d = dgs + r * NC;
      d[u] = dpi; d[H + u] = dpf; d[2 * H + u] = dpg; d[3 * H + u] = dpo;
      const size_t gb = ((size_t)b * T + t) * NC;
      a.gates[gb + u] = dpi; a.gates[gb + H + u] = dpf; a.gates[gb + 2 * H + u] = dpg; a.gates[gb + 3 * H + u] = dpo;
    } else if (tid < kRnnRows * H) {
      float* d = dgs + r * NC;
      d[u] = 0.f; d[H + u] = 0.f; d[2 * H + u] = 0.f; d[3 * H + u] = 0.f;
    }
    if (t > 0) load_step(t - 1);                // overlaps the mat-vec below
    __syncthreads();
    if (tid < kRnnRows * H) {                   // dh_{t-1}[r][u] = sum_j dG[r][j] * W[j][u]
      const float* d = dgs + r * NC;
      const WT* wp = W + u;
      float s0 = 0.f, s1 = 0.f;
      int j2 = 0;
      for (; j2 + 2 <= NC; j2 += 2) {
        s0 = fmaf(d[j2], wload<WT>(wp + (size_t)j2 * H), s0);
        s1 = fmaf(d[j2 + 1], wload<WT>(wp + (size_t)(j2 + 1) * H), s1);
      }
      dhs[r * H + u] = s0 + s1;                 // safe: dhs was consumed before the barrier above
    }
    __syncthreads();
  }
  if (own) {
    if (a.dh0) a.dh0[(size_t)b * H + u] = dhs[r * H + u];
    if (a.dc0) a.dc0[(size_t)b * H + u] = dc_next;
  }
}

// ----------------------------------------------------------------- GRU forward --
template <typename WT>
__global__ void gru_fwd_kernel(RnnArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 3 * H, T = a.T;
  const int Hp = (H + 3) & ~3;
  WT* Wt = reinterpret_cast<WT*>(smem_raw);    // [H][NC]
  float* hs = reinterpret_cast<float*>(smem_raw + ((size_t)H * NC * sizeof(WT) + 15) / 16 * 16);  // [4][Hp]
  float* rs = hs + kRnnRows * Hp;              // [4][NC] recurrent pre-activations
  float* xs = rs + kRnnRows * NC;              // [4][NC] x-projection of the step
  const int tid = threadIdx.x, b0 = blockIdx.x * kRnnRows;
  for (int i = tid; i < NC * H; i += blockDim.x) {
    const int j = i / H, k = i - j * H;
    Wt[(size_t)k * NC + j] = wstore<WT>(a.whh[i]);
  }
  for (int i = tid; i < kRnnRows * Hp; i += blockDim.x) {
    const int r = i / Hp, k = i - r * Hp, b = b0 + r;
    hs[i] = (a.h0 && b < a.B && k < H) ? a.h0[(size_t)b * H + k] : 0.f;
  }
  const int r2 = tid / H, u2 = tid - r2 * H;
  const bool own = tid < kRnnRows * H && (b0 + r2) < a.B;
  const float bhn = (own && a.bhn) ? a.bhn[u2] : 0.f;
  const int j = tid;
  const bool col = j < NC;
  float gx[kRnnRows];
#pragma unroll
  for (int r = 0; r < kRnnRows; ++r) gx[r] = (col && b0 + r < a.B) ? a.gates[((size_t)(b0 + r) * T) * NC + j] : 0.f;
  __syncthreads();
  for (int t = 0; t < T; ++t) {
    if (col) {
      float acc[kRnnRows];
#pragma unroll
      for (int r = 0; r < kRnnRows; ++r) { acc[r] = 0.f; xs[r * NC + j] = gx[r]; }
      if (t + 1 < T) {
#pragma unroll
        for (int r = 0; r < kRnnRows; ++r)
          gx[r] = (b0 + r < a.B) ? a.gates[((size_t)(b0 + r) * T + t + 1) * NC + j] : 0.f;
      }
      const WT* wp = Wt + j;
      int k = 0;
      for (; k + 4 <= H; k += 4) {
        const float w0 = wload<WT>(wp + (size_t)k * NC), w1 = wload<WT>(wp + (size_t)(k + 1) * NC);
        const float w2 = wload<WT>(wp + (size_t)(k + 2) * NC), w3 = wload<WT>(wp + (size_t)(k + 3) * NC);
#pragma unroll
        for (int r = 0; r < kRnnRows; ++r) {
          const float4 h4 = *reinterpret_cast<const float4*>(hs + r * Hp + k);
          acc[r] = fmaf(w0, h4.x, fmaf(w1, h4.y, fmaf(w2, h4.z, fmaf(w3, h4.w, acc[r]))));
        }
      }
      for (; k < H; ++k) {
        const float w = wload<WT>(wp + (size_t)k * NC);
#pragma unroll
        for (int r = 0; r < kRnnRows; ++r) acc[r] = fmaf(w, hs[r * Hp + k], acc[r]);
      }
#pragma unroll
      for (int r = 0; r < kRnnRows; ++r) rs[r * NC + j] = acc[r];
    }
    __syncthreads();
    if (own) {
      const float* rr = rs + r2 * NC;
      const float* xx = xs + r2 * NC;
      const float rg = gate_act(xx[u2] + rr[u2], a.hard);
      const float zg = gate_act(xx[H + u2] + rr[H + u2], a.hard);
      const float hn = rr[2 * H + u2] + bhn;
      const float ng = tanhf_(fmaf(rg, hn, xx[2 * H + u2]));
      const float hp = hs[r2 * Hp + u2];
      const float h = fmaf(zg, hp - ng, ng);     // (1-z)*n + z*h
      const size_t o = ((size_t)(b0 + r2) * T + t);
      a.hout[o * H + u2] = h;
      if (a.saved) {
        float* s = a.saved + o * 4 * H;
        s[u2] = rg; s[H + u2] = zg; s[2 * H + u2] = ng; s[3 * H + u2] = hn;
      }
      hs[r2 * Hp + u2] = h;
    }
    __syncthreads();
  }
  if (own && a.hT) a.hT[(size_t)(b0 + r2) * H + u2] = hs[r2 * Hp + u2];
}

// ---------------------------------------------------------------- GRU backward --
template <typename WT>
__global__ void gru_bwd_kernel(RnnBwdArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 3 * H, T = a.T;
  WT* W = reinterpret_cast<WT*>(smem_raw);                 // [NC][H]
  float* dgs = reinterpret_cast<float*>(smem_raw + ((size_t)H * NC * sizeof(WT) + 15) / 16 * 16);  // [4][NC]: dpr,dpz,dhn
  float* dhs = dgs + kRnnRows * NC;                         // [4][H]
  const int tid = threadIdx.x, b0 = blockIdx.x * kRnnRows;
  for (int i = tid; i < NC * H; i += blockDim.x) W[i] = wstore<WT>(a.whh[i]);
  const int r = tid / H, u = tid - r * H, b = b0 + r;
  const bool own = tid < kRnnRows * H && b < a.B;
  if (tid < kRnnRows * H) dhs[r * H + u] = (own && a.dhT) ? a.dhT[(size_t)b * H + u] : 0.f;
  float rg = 0, zg = 0, ng = 0, hn = 0, hprev = 0, dyv = 0;
  auto load_step = [&](int t) {
    if (!own) return;
    const float* s = a.gates + ((size_t)b * T + t) * 4 * H;
    rg = s[u]; zg = s[H + u]; ng = s[2 * H + u]; hn = s[3 * H + u];
    hprev = t > 0 ? a.hout[((size_t)b * T + t - 1) * H + u] : (a.h0 ? a.h0[(size_t)b * H + u] : 0.f);
    dyv = a.dy ? a.dy[(size_t)b * a.dy_sb + (size_t)t * a.dy_st + u] : 0.f;
  };
  load_step(T - 1);
  __syncthreads();
  for (int t = T - 1; t >= 0; --t) {
    float dh_direct = 0.f;
    if (own) {
      const float dh = dyv + dhs[r * H + u];
      const float dn = dh * (1.f - zg);
      const float dpz = dh * (hprev - ng) * gate_der(zg, a.hard);
      const float dpn = dn * (1.f - ng * ng);
      const float dpr = dpn * hn * gate_der(rg, a.hard);
      const float dhn = dpn * rg;
      dh_direct = dh * zg;
      float* d = dgs + r * NC;
      d[u] = dpr; d[H + u] = dpz; d[2 * H + u] = dhn;
      float* s = a.gates + ((size_t)b * T + t) * 4 * H;
      s[u] = dpr; s[H + u] = dpz; s[2 * H + u] = dpn; s[3 * H + u] = dhn;
    } else if (tid < kRnnRows * H) {
      float* d = dgs + r * NC;
      d[u] = 0.f; d[H + u] = 0.f; d[2 * H + u] = 0.f;
    }
    if (t > 0) load_step(t - 1);
    __syncthreads();
    if (tid < kRnnRows * H) {
      const float* d = dgs + r * NC;
      const WT* wp = W + u;
      float s0 = 0.f, s1 = 0.f;
      int j2 = 0;
      for (; j2 + 2 <= NC; j2 += 2) {
        s0 = fmaf(d[j2], wload<WT>(wp + (size_t)j2 * H), s0);
        s1 = fmaf(d[j2 + 1], wload<WT>(wp + (size_t)(j2 + 1) * H), s1);
      }
      for (; j2 < NC; ++j2) s0 = fmaf(d[j2], wload<WT>(wp + (size_t)j2 * H), s0);
      dhs[r * H + u] = s0 + s1 + dh_direct;
    }
    __syncthreads();
  }
  if (own && a.dh0) a.dh0[(size_t)b * H + u] = dhs[r * H + u];
}

}  // namespace zoo

using namespace zoo;

namespace {
constexpr size_t kSmemMax = 200 * 1024;

// shared memory bytes for weights of type WT plus the float tiles
template <typename WT> size_t rnn_smem(int H, int G, int tiles_floats) {
  return ((size_t)H * G * H * sizeof(WT) + 15) / 16 * 16 + (size_t)tiles_floats * sizeof(float);
}

template <typename K> int set_smem(K kernel, size_t bytes) {
  if (bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) { zoo_set_last_error(cudaGetErrorString(e), __FILE__, __LINE__); return ZOO_ERR_CUDA; }
  }
  return ZOO_OK;
}

int check_shape(int B, int T, int H) {
  if (B <= 0 || T <= 0 || H <= 0) { zoo_set_last_error("rnn: B, T, H must be positive", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (H > 128) { zoo_set_last_error("rnn: hidden size above 128 is not supported by the persistent kernel", __FILE__, __LINE__); return ZOO_ERR_TOO_LARGE; }
  return ZOO_OK;
}
}  // namespace

/* Returns 1 when the persistent kernels can run this hidden size, else 0. */
ZOO_API int zoo_rnn_supported(int H) { return H > 0 && H <= 128; }

ZOO_API int zoo_lstm_forward(float* gates, const float* whh, const float* h0, const float* c0, float* hout, float* cout,
                             float* hT, float* cT, int B, int T, int H, int save, int hard, void* stream) {
  int rc = check_shape(B, T, H);
  if (rc) return rc;
  if (!gates || !whh || !hout) { zoo_set_last_error("lstm_forward: null pointer", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  RnnArgs a{gates, nullptr, whh, nullptr, h0, c0, hout, cout, hT, cT, B, T, H, save, hard};
  const int Hp = (H + 3) & ~3, NC = 4 * H;
  const int tiles = kRnnRows * Hp + kRnnRows * NC;
  const unsigned grid = (B + kRnnRows - 1) / kRnnRows, block = (NC + 31) / 32 * 32;
  cudaStream_t s = (cudaStream_t)stream;
  size_t sm = rnn_smem<float>(H, 4, tiles);
  if (sm <= kSmemMax) {
    if ((rc = set_smem(lstm_fwd_kernel<float>, sm))) return rc;
    lstm_fwd_kernel<float><<<grid, block, sm, s>>>(a);
  } else {
    sm = rnn_smem<__nv_bfloat16>(H, 4, tiles);
    if ((rc = set_smem(lstm_fwd_kernel<__nv_bfloat16>, sm))) return rc;
    lstm_fwd_kernel<__nv_bfloat16><<<grid, block, sm, s>>>(a);
  }
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_lstm_backward(float* gates, const float* whh, const float* hout, const float* cout, const float* h0,
                              const float* c0, const float* dy, long long dy_sb, long long dy_st, const float* dhT,
                              const float* dcT, float* dh0, float* dc0, int B, int T, int H, int hard, void* stream) {
  int rc = check_shape(B, T, H);
  if (rc) return rc;
  if (!gates || !whh || !cout) { zoo_set_last_error("lstm_backward: null pointer", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  RnnBwdArgs a{gates, whh, hout, cout, h0, c0, dy, dy_sb, dy_st, dhT, dcT, dh0, dc0, B, T, H, hard};
  const int NC = 4 * H;
  const int tiles = kRnnRows * NC + kRnnRows * H;
  const unsigned grid = (B + kRnnRows - 1) / kRnnRows, block = (NC + 31) / 32 * 32;
  cudaStream_t s = (cudaStream_t)stream;
  size_t sm = rnn_smem<float>(H, 4, tiles);
  if (sm <= kSmemMax) {
    if ((rc = set_smem(lstm_bwd_kernel<float>, sm))) return rc;
    lstm_bwd_kernel<float><<<grid, block, sm, s>>>(a);
  } else {
    sm = rnn_smem<__nv_bfloat16>(H, 4, tiles);
    if ((rc = set_smem(lstm_bwd_kernel<__nv_bfloat16>, sm))) return rc;
    lstm_bwd_kernel<__nv_bfloat16><<<grid, block, sm, s>>>(a);
  }
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_gru_forward(const float* gx, float* saved, const float* whh, const float* bhn, const float* h0,
                            float* hout, float* hT, int B, int T, int H, int hard, void* stream) {
  int rc = check_shape(B, T, H);
  if (rc) return rc;
  if (!gx || !whh || !hout) { zoo_set_last_error("gru_forward: null pointer", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  RnnArgs a{const_cast<float*>(gx), saved, whh, bhn, h0, nullptr, hout, nullptr, hT, nullptr, B, T, H, 0, hard};
  const int Hp = (H + 3) & ~3, NC = 3 * H;
  const int tiles = kRnnRows * Hp + 2 * kRnnRows * NC;
  const unsigned grid = (B + kRnnRows - 1) / kRnnRows, block = (4 * H + 31) / 32 * 32;
  cudaStream_t s = (cudaStream_t)stream;
  size_t sm = rnn_smem<float>(H, 3, tiles);
  if (sm <= kSmemMax) {
    if ((rc = set_smem(gru_fwd_kernel<float>, sm))) return rc;
    gru_fwd_kernel<float><<<grid, block, sm, s>>>(a);
  } else {
    sm = rnn_smem<__nv_bfloat16>(H, 3, tiles);
    if ((rc = set_smem(gru_fwd_kernel<__nv_bfloat16>, sm))) return rc;
    gru_fwd_kernel<__nv_bfloat16><<<grid, block, sm, s>>>(a);
  }
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_gru_backward(float* saved, const float* whh, const float* hout, const float* h0, const float* dy,
                             long long dy_sb, long long dy_st, const float* dhT, float* dh0, int B, int T, int H,
                             int hard, void* stream) {
  int rc = check_shape(B, T, H);
  if (rc) return rc;
  if (!saved || !whh || !hout) { zoo_set_last_error("gru_backward: null pointer", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  RnnBwdArgs a{saved, whh, hout, nullptr, h0, nullptr, dy, dy_sb, dy_st, dhT, nullptr, dh0, nullptr, B, T, H, hard};
  const int NC = 3 * H;
  const int tiles = kRnnRows * NC + kRnnRows * H;
  const unsigned grid = (B + kRnnRows - 1) / kRnnRows, block = (4 * H + 31) / 32 * 32;
  cudaStream_t s = (cudaStream_t)stream;
  size_t sm = rnn_smem<float>(H, 3, tiles);
  if (sm <= kSmemMax) {
    if ((rc = set_smem(gru_bwd_kernel<float>, sm))) return rc;
    gru_bwd_kernel<float><<<grid, block, sm, s>>>(a);
  } else {
    sm = rnn_smem<__nv_bfloat16>(H, 3, tiles);
    if ((rc = set_smem(gru_bwd_kernel<__nv_bfloat16>, sm))) return rc;
    gru_bwd_kernel<__nv_bfloat16><<<grid, block, sm, s>>>(a);
  }
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
