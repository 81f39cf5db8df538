// K6 -- persistent recurrent kernels: LSTM and GRU, forward and backward-through-time.
//
// Replaces the per-timestep cell loops of the reference's recurrent layers
// (zoo/.../pipeline/api/keras/layers/LSTM.scala, GRU.scala, InternalRecurrent.scala) and of its
// PyTorch AutoML models (pyzoo/zoo/automl/model/VanillaLSTM_pytorch.py:29-56,
// Seq2Seq_pytorch.py:31-101).  The shapes are small (H 8..128, T <= a few hundred, B <= a few
// thousand), so a step is latency-bound: launching one kernel per time step (what a cell loop or a
// library RNN does) costs more than the math.  Here ONE launch runs all T steps.
//
// Decomposition.  The recurrence is independent across batch rows, so a CTA owns R batch rows (1, 2 or 4)
// for the whole sequence and never talks to another CTA.  A step is a dependent chain (dot products -> activation -> barrier ->
// cell update -> barrier) that a CTA cannot shorten by being bigger, so R is the SMALLEST tile for which every CTA of the
// launch is resident at once: B 1024, H 32 runs as 1024 CTAs of one row (28 warps per SM) instead of 256 CTAs of four rows
// (7 warps per SM, each thread doing four rows' worth of serial work per step).
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

constexpr int kRnnRows = 4;  // batch rows per CTA, upper bound (kernels are instantiated for 1, 2 and 4)

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
template <typename WT, int R>
__global__ void lstm_fwd_kernel(RnnArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 4 * H, T = a.T;
  const int Hp = (H + 3) & ~3;                 // h tile rows padded to float4
  WT* Wt = reinterpret_cast<WT*>(smem_raw);    // [H][NC]  (transposed: column j contiguous over j)
  float* hs = reinterpret_cast<float*>(smem_raw + ((size_t)H * NC * sizeof(WT) + 15) / 16 * 16);  // [4][Hp]
  float* gs = hs + R * Hp;              // [4][NC]
  const int tid = threadIdx.x;
  const int b0 = blockIdx.x * R;
  for (int i = tid; i < NC * H; i += blockDim.x) {
    const int j = i / H, k = i - j * H;        // whh[j][k] -> Wt[k][j]
    Wt[(size_t)k * NC + j] = wstore<WT>(a.whh[i]);
  }
  for (int i = tid; i < R * Hp; i += blockDim.x) {
    const int r = i / Hp, k = i - r * Hp, b = b0 + r;
    hs[i] = (a.h0 && b < a.B && k < H) ? a.h0[(size_t)b * H + k] : 0.f;
  }
  // phase-2 ownership: thread -> (row r2, unit u2); c lives in a register
  const int r2 = tid / H, u2 = tid - r2 * H;
  const bool own = tid < R * H && (b0 + r2) < a.B;
  float c = (own && a.c0) ? a.c0[(size_t)(b0 + r2) * H + u2] : 0.f;
  const int j = tid;
  const bool col = j < NC;
  const bool is_g = col && j >= 2 * H && j < 3 * H;
  float gx[R];
#pragma unroll
  for (int r = 0; r < R; ++r) gx[r] = (col && b0 + r < a.B) ? a.gates[((size_t)(b0 + r) * T) * NC + j] : 0.f;
  __syncthreads();
  for (int t = 0; t < T; ++t) {
    float acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = gx[r];
    if (t + 1 < T) {                            // prefetch the next step's x-projection
#pragma unroll
      for (int r = 0; r < R; ++r)
        gx[r] = (col && b0 + r < a.B) ? a.gates[((size_t)(b0 + r) * T + t + 1) * NC + j] : 0.f;
    }
    if (col) {
      const WT* wp = Wt + j;
      int k = 0;
      for (; k + 4 <= H; k += 4) {
        const float w0 = wload<WT>(wp + (size_t)k * NC), w1 = wload<WT>(wp + (size_t)(k + 1) * NC);
        const float w2 = wload<WT>(wp + (size_t)(k + 2) * NC), w3 = wload<WT>(wp + (size_t)(k + 3) * NC);
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float4 h4 = *reinterpret_cast<const float4*>(hs + r * Hp + k);
          acc[r] = fmaf(w0, h4.x, fmaf(w1, h4.y, fmaf(w2, h4.z, fmaf(w3, h4.w, acc[r]))));
        }
      }
      for (; k < H; ++k) {
        const float w = wload<WT>(wp + (size_t)k * NC);
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = fmaf(w, hs[r * Hp + k], acc[r]);
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
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
  const float* dy;     // [B,T,H] through (dy_sb, dy_st), or null (gradient w.r.t. every output step)
  long long dy_sb, dy_st;  // strides of dy in elements (batch, time); unit stride 1
  const float* dhT;    // [B,H] or null
  const float* dcT;    // [B,H] or null
  float* dh0;          // [B,H] or null
  float* dc0;          // [B,H] or null
  int B, T, H;
  int hard;
};

// --------------------------------------------------------------- LSTM backward --
template <typename WT, int R>
__global__ void lstm_bwd_kernel(RnnBwdArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 4 * H, T = a.T;
  WT* W = reinterpret_cast<WT*>(smem_raw);                 // [NC][H] natural layout
  float* dgs = reinterpret_cast<float*>(smem_raw + ((size_t)H * NC * sizeof(WT) + 15) / 16 * 16);  // [4][NC]
  float* dhs = dgs + R * NC;                         // [4][H]
  const int tid = threadIdx.x, b0 = blockIdx.x * R;
  for (int i = tid; i < NC * H; i += blockDim.x) W[i] = wstore<WT>(a.whh[i]);
  const int r = tid / H, u = tid - r * H, b = b0 + r;
  const bool own = tid < R * H && b < a.B;
  if (tid < R * H) dhs[r * H + u] = (own && a.dhT) ? a.dhT[(size_t)b * H + u] : 0.f;
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
    } else if (tid < R * H) {
      float* d = dgs + r * NC;
      d[u] = 0.f; d[H + u] = 0.f; d[2 * H + u] = 0.f; d[3 * H + u] = 0.f;
    }
    if (t > 0) load_step(t - 1);                // overlaps the mat-vec below
    __syncthreads();
    if (tid < R * H) {                   // dh_{t-1}[r][u] = sum_j dG[r][j] * W[j][u]
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
template <typename WT, int R>
__global__ void gru_fwd_kernel(RnnArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 3 * H, T = a.T;
  const int Hp = (H + 3) & ~3;
  WT* Wt = reinterpret_cast<WT*>(smem_raw);    // [H][NC]
  float* hs = reinterpret_cast<float*>(smem_raw + ((size_t)H * NC * sizeof(WT) + 15) / 16 * 16);  // [4][Hp]
  float* rs = hs + R * Hp;              // [4][NC] recurrent pre-activations
  float* xs = rs + R * NC;              // [4][NC] x-projection of the step
  const int tid = threadIdx.x, b0 = blockIdx.x * R;
  for (int i = tid; i < NC * H; i += blockDim.x) {
    const int j = i / H, k = i - j * H;
    Wt[(size_t)k * NC + j] = wstore<WT>(a.whh[i]);
  }
  for (int i = tid; i < R * Hp; i += blockDim.x) {
    const int r = i / Hp, k = i - r * Hp, b = b0 + r;
    hs[i] = (a.h0 && b < a.B && k < H) ? a.h0[(size_t)b * H + k] : 0.f;
  }
  const int r2 = tid / H, u2 = tid - r2 * H;
  const bool own = tid < R * H && (b0 + r2) < a.B;
  const float bhn = (own && a.bhn) ? a.bhn[u2] : 0.f;
  const int j = tid;
  const bool col = j < NC;
  float gx[R];
#pragma unroll
  for (int r = 0; r < R; ++r) gx[r] = (col && b0 + r < a.B) ? a.gates[((size_t)(b0 + r) * T) * NC + j] : 0.f;
  __syncthreads();
  for (int t = 0; t < T; ++t) {
    if (col) {
      float acc[R];
#pragma unroll
      for (int r = 0; r < R; ++r) { acc[r] = 0.f; xs[r * NC + j] = gx[r]; }
      if (t + 1 < T) {
#pragma unroll
        for (int r = 0; r < R; ++r)
          gx[r] = (b0 + r < a.B) ? a.gates[((size_t)(b0 + r) * T + t + 1) * NC + j] : 0.f;
      }
      const WT* wp = Wt + j;
      int k = 0;
      for (; k + 4 <= H; k += 4) {
        const float w0 = wload<WT>(wp + (size_t)k * NC), w1 = wload<WT>(wp + (size_t)(k + 1) * NC);
        const float w2 = wload<WT>(wp + (size_t)(k + 2) * NC), w3 = wload<WT>(wp + (size_t)(k + 3) * NC);
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float4 h4 = *reinterpret_cast<const float4*>(hs + r * Hp + k);
          acc[r] = fmaf(w0, h4.x, fmaf(w1, h4.y, fmaf(w2, h4.z, fmaf(w3, h4.w, acc[r]))));
        }
      }
      for (; k < H; ++k) {
        const float w = wload<WT>(wp + (size_t)k * NC);
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = fmaf(w, hs[r * Hp + k], acc[r]);
      }
#pragma unroll
      for (int r = 0; r < R; ++r) rs[r * NC + j] = acc[r];
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
template <typename WT, int R>
__global__ void gru_bwd_kernel(RnnBwdArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 3 * H, T = a.T;
  WT* W = reinterpret_cast<WT*>(smem_raw);                 // [NC][H]
  float* dgs = reinterpret_cast<float*>(smem_raw + ((size_t)H * NC * sizeof(WT) + 15) / 16 * 16);  // [4][NC]: dpr,dpz,dhn
  float* dhs = dgs + R * NC;                         // [4][H]
  const int tid = threadIdx.x, b0 = blockIdx.x * R;
  for (int i = tid; i < NC * H; i += blockDim.x) W[i] = wstore<WT>(a.whh[i]);
  const int r = tid / H, u = tid - r * H, b = b0 + r;
  const bool own = tid < R * H && b < a.B;
  if (tid < R * H) dhs[r * H + u] = (own && a.dhT) ? a.dhT[(size_t)b * H + u] : 0.f;
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
    } else if (tid < R * H) {
      float* d = dgs + r * NC;
      d[u] = 0.f; d[H + u] = 0.f; d[2 * H + u] = 0.f;
    }
    if (t > 0) load_step(t - 1);
    __syncthreads();
    if (tid < R * H) {
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
// ------------------------------------------------------------------------------------------------------------------------
// Warp-per-row forward kernels (H <= 64).  In the CTA-tiled kernels above a step is: dot products spread over 4H threads ->
// shared-memory hand-over -> barrier -> cell update on H threads -> barrier, ~2,800 cycles per step at H = 32 however small the
// tile, because the chain is latency (two block barriers, shared-memory round trips), not math.  Here ONE WARP owns a batch
// row for the whole sequence: lane u computes all gates of hidden unit u (and u + 32 when H > 32), h_{t-1} lives in registers
// and reaches the other lanes through shuffles, c_t lives in a register -- a step has no barrier and touches no shared memory
// at H <= 32 (the lane's 4 x 32 recurrent weights sit in registers); for 32 < H <= 64 the weights are read from shared memory
// as one float4 = (i, f, g, o) per (k, unit).  The buffers written for backward are the same as the tiled kernels'.
// The x-projection is prefetched into registers kAhead steps ahead and shifted down one slot per step.  Measured at B 1024, T 50,
// H 32 (per layer): one step ahead 43.4 us, two steps with the shift 39.8 us; unrolling the time loop by the depth instead of
// shifting was SLOWER (2x: 57.7 us, 4x: 55 us -- the step body is ~300 instructions and the unrolled loop no longer sits in the
// instruction cache of a scheduler that runs one or two warps), a cp.async ring through shared memory 47.6 us, time-major
// buffers no change.  cuDNN's persistent kernel takes 43.7 us on the same shape.
template <int UPL> struct RnnAhead { static constexpr int value = UPL == 1 ? 2 : 1; };   // deeper unrolling of the H > 32 body overflows ptxas
// RW batch rows per warp: the recurrent weights (registers / shared memory) and every shuffle index are shared by the rows, and
// the rows' FMA chains are independent -- a single row leaves the warp waiting on its own 4-cycle FMA latency most of the time.
template <int UPL, int NG, int RW>        // NG gates: 4 LSTM, 3 GRU
__device__ __forceinline__ void warp_dots(const float (&hreg)[RW][UPL], const float* __restrict__ wreg /* UPL == 1: [NG][32] */,
                                          const float4* __restrict__ wsm /* UPL > 1: [32 * UPL][32 * UPL] */, int lane,
                                          float (&acc)[RW][UPL][NG]) {
#pragma unroll
  for (int kk = 0; kk < UPL; ++kk) {
    // all 32 broadcasts of this chunk first, then the FMAs: issue is in order, so a shuffle placed between FMAs that wait on
    // the previous shuffle exposes its ~25-cycle latency once per k (measured: the step time did not depend on the FMA count)
    float hk[RW][32];
#pragma unroll
    for (int kl = 0; kl < 32; ++kl)
#pragma unroll
      for (int r = 0; r < RW; ++r) hk[r][kl] = __shfl_sync(0xffffffffu, hreg[r][kk], kl);
#pragma unroll
    for (int kl = 0; kl < 32; ++kl) {
      if (UPL == 1) {
#pragma unroll
        for (int g = 0; g < NG; ++g)
#pragma unroll
          for (int r = 0; r < RW; ++r) acc[r][0][g] = fmaf(wreg[g * 32 + kl], hk[r][kl], acc[r][0][g]);
      } else {
#pragma unroll
        for (int i = 0; i < UPL; ++i) {
          const float4 w = wsm[(kk * 32 + kl) * (32 * UPL) + lane + 32 * i];
#pragma unroll
          for (int r = 0; r < RW; ++r) {
            acc[r][i][0] = fmaf(w.x, hk[r][kl], acc[r][i][0]);
            acc[r][i][1] = fmaf(w.y, hk[r][kl], acc[r][i][1]);
            acc[r][i][2] = fmaf(w.z, hk[r][kl], acc[r][i][2]);
            if (NG == 4) acc[r][i][3] = fmaf(w.w, hk[r][kl], acc[r][i][3]);
          }
        }
      }
    }
  }
}

// Load the lane's recurrent weights.  whh is [NG * H, H] row-major (gate-major rows).  UPL == 1: wreg[g][k] = whh[g*H + lane][k];
// UPL > 1: shared float4 table [k][unit] = (gate 0..3 of that unit at input k), zero outside H.
template <int UPL, int NG>
__device__ __forceinline__ void warp_load_w(const float* __restrict__ whh, int H, int lane, float* wreg, float4* wsm) {
  if (UPL == 1) {
#pragma unroll
    for (int g = 0; g < NG; ++g)
#pragma unroll
      for (int k = 0; k < 32; ++k) wreg[g * 32 + k] = (lane < H && k < H) ? whh[((size_t)g * H + lane) * H + k] : 0.f;
  } else {
    constexpr int HP = 32 * UPL;
    for (int i = threadIdx.x; i < HP * HP; i += blockDim.x) {
      const int k = i / HP, u = i - k * HP;
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (k < H && u < H) {
        w.x = whh[((size_t)0 * H + u) * H + k];
        w.y = whh[((size_t)1 * H + u) * H + k];
        w.z = whh[((size_t)2 * H + u) * H + k];
        if (NG == 4) w.w = whh[((size_t)3 * H + u) * H + k];
      }
      wsm[i] = w;
    }
    __syncthreads();
  }
}

template <int UPL, int RW>
__global__ void __launch_bounds__(128)
lstm_fwd_warp_kernel(RnnArgs a) {
  constexpr int kAhead = RnnAhead<UPL>::value;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 4 * H, T = a.T;
  const int lane = threadIdx.x & 31;
  const int b0 = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RW;
  float wreg[UPL == 1 ? 128 : 1];
  float4* wsm = reinterpret_cast<float4*>(smem_raw);
  warp_load_w<UPL, 4>(a.whh, H, lane, wreg, wsm);
  if (b0 >= a.B) return;
  float hreg[RW][UPL], creg[RW][UPL], gx[kAhead][RW][UPL][4];
  bool ok[RW][UPL];
#pragma unroll
  for (int r = 0; r < RW; ++r) {
    const int b = b0 + r;
    const float* grow = a.gates + (size_t)b * T * NC;
#pragma unroll
    for (int i = 0; i < UPL; ++i) {
      const int u = lane + 32 * i;
      ok[r][i] = u < H && b < a.B;
      hreg[r][i] = (ok[r][i] && a.h0) ? a.h0[(size_t)b * H + u] : 0.f;
      creg[r][i] = (ok[r][i] && a.c0) ? a.c0[(size_t)b * H + u] : 0.f;
#pragma unroll
      for (int p = 0; p < kAhead; ++p)
#pragma unroll
        for (int g = 0; g < 4; ++g) gx[p][r][i][g] = (ok[r][i] && p < T) ? grow[(size_t)p * NC + g * H + u] : 0.f;
    }
  }
  for (int t = 0; t < T; ++t) {
   {
    float acc[RW][UPL][4];
#pragma unroll
    for (int r = 0; r < RW; ++r) {
      const float* grow = a.gates + (size_t)(b0 + r) * T * NC;
#pragma unroll
      for (int i = 0; i < UPL; ++i)
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          acc[r][i][g] = gx[0][r][i][g];
#pragma unroll
          for (int p = 0; p + 1 < kAhead; ++p) gx[p][r][i][g] = gx[p + 1][r][i][g];
          if (ok[r][i] && t + kAhead < T) gx[kAhead - 1][r][i][g] = grow[(size_t)(t + kAhead) * NC + g * H + lane + 32 * i];
        }
    }
    warp_dots<UPL, 4, RW>(hreg, wreg, wsm, lane, acc);
#pragma unroll
    for (int r = 0; r < RW; ++r) {
#pragma unroll
      for (int i = 0; i < UPL; ++i) {
        const int u = lane + 32 * i;
        const float ig = gate_act(acc[r][i][0], a.hard), fg = gate_act(acc[r][i][1], a.hard), gg = tanhf_(acc[r][i][2]),
                    og = gate_act(acc[r][i][3], a.hard);
        creg[r][i] = fmaf(fg, creg[r][i], ig * gg);
        const float h = og * tanhf_(creg[r][i]);
        hreg[r][i] = ok[r][i] ? h : 0.f;
        if (ok[r][i]) {
          const size_t o = ((size_t)(b0 + r) * T + t);
          a.hout[o * H + u] = h;
          if (a.cout) a.cout[o * H + u] = creg[r][i];
          if (a.save) {
            float* gsv = a.gates + o * NC;
            gsv[u] = ig; gsv[H + u] = fg; gsv[2 * H + u] = gg; gsv[3 * H + u] = og;
          }
        }
      }
    }
   }
  }
#pragma unroll
  for (int r = 0; r < RW; ++r)
#pragma unroll
    for (int i = 0; i < UPL; ++i) {
      if (!ok[r][i]) continue;
      const size_t o = (size_t)(b0 + r) * H + lane + 32 * i;
      if (a.hT) a.hT[o] = hreg[r][i];
      if (a.cT) a.cT[o] = creg[r][i];
    }
}

template <int UPL, int RW>
__global__ void __launch_bounds__(128)
gru_fwd_warp_kernel(RnnArgs a) {
  constexpr int kAhead = RnnAhead<UPL>::value;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int H = a.H, NC = 3 * H, T = a.T;
  const int lane = threadIdx.x & 31;
  const int b0 = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RW;
  float wreg[UPL == 1 ? 96 : 1];
  float4* wsm = reinterpret_cast<float4*>(smem_raw);
  warp_load_w<UPL, 3>(a.whh, H, lane, wreg, wsm);
  if (b0 >= a.B) return;
  float hreg[RW][UPL], bhn[UPL], gx[kAhead][RW][UPL][3];
  bool ok[RW][UPL];
#pragma unroll
  for (int i = 0; i < UPL; ++i) bhn[i] = (lane + 32 * i < H && a.bhn) ? a.bhn[lane + 32 * i] : 0.f;
#pragma unroll
  for (int r = 0; r < RW; ++r) {
    const int b = b0 + r;
    const float* grow = a.gates + (size_t)b * T * NC;
#pragma unroll
    for (int i = 0; i < UPL; ++i) {
      const int u = lane + 32 * i;
      ok[r][i] = u < H && b < a.B;
      hreg[r][i] = (ok[r][i] && a.h0) ? a.h0[(size_t)b * H + u] : 0.f;
#pragma unroll
      for (int p = 0; p < kAhead; ++p)
#pragma unroll
        for (int g = 0; g < 3; ++g) gx[p][r][i][g] = (ok[r][i] && p < T) ? grow[(size_t)p * NC + g * H + u] : 0.f;
    }
  }
  for (int t = 0; t < T; ++t) {
   {
    float acc[RW][UPL][3], xg[RW][UPL][3];
#pragma unroll
    for (int r = 0; r < RW; ++r) {
      const float* grow = a.gates + (size_t)(b0 + r) * T * NC;
#pragma unroll
      for (int i = 0; i < UPL; ++i)
#pragma unroll
        for (int g = 0; g < 3; ++g) {
          acc[r][i][g] = 0.f;
          xg[r][i][g] = gx[0][r][i][g];
#pragma unroll
          for (int p = 0; p + 1 < kAhead; ++p) gx[p][r][i][g] = gx[p + 1][r][i][g];
          if (ok[r][i] && t + kAhead < T) gx[kAhead - 1][r][i][g] = grow[(size_t)(t + kAhead) * NC + g * H + lane + 32 * i];
        }
    }
    warp_dots<UPL, 3, RW>(hreg, wreg, wsm, lane, acc);
#pragma unroll
    for (int r = 0; r < RW; ++r) {
#pragma unroll
      for (int i = 0; i < UPL; ++i) {
        const int u = lane + 32 * i;
        const float rg = gate_act(xg[r][i][0] + acc[r][i][0], a.hard), zg = gate_act(xg[r][i][1] + acc[r][i][1], a.hard);
        const float hn = acc[r][i][2] + bhn[i];
        const float ng = tanhf_(fmaf(rg, hn, xg[r][i][2]));
        const float h = fmaf(zg, hreg[r][i] - ng, ng);     // (1-z)*n + z*h
        hreg[r][i] = ok[r][i] ? h : 0.f;
        if (ok[r][i]) {
          const size_t o = ((size_t)(b0 + r) * T + t);
          a.hout[o * H + u] = h;
          if (a.saved) {
            float* sv = a.saved + o * 4 * H;
            sv[u] = rg; sv[H + u] = zg; sv[2 * H + u] = ng; sv[3 * H + u] = hn;
          }
        }
      }
    }
   }
  }
#pragma unroll
  for (int r = 0; r < RW; ++r)
#pragma unroll
    for (int i = 0; i < UPL; ++i)
      if (ok[r][i] && a.hT) a.hT[(size_t)(b0 + r) * H + lane + 32 * i] = hreg[r][i];
}

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

// Rows per CTA: the smallest tile that keeps the whole launch resident (see the top of this file); 4 when nothing fits at once.
template <typename SmemFn>
int pick_rows(int B, unsigned threads, SmemFn smem_of_rows) {
  const int cands[3] = {1, 2, 4};
  for (int i = 0; i < 3; ++i) {
    const int r = cands[i];
    const size_t sm = smem_of_rows(r);
    if (sm > kSmemMax) continue;
    size_t by_smem = (227u * 1024u) / (sm + 1024u), by_thr = 2048u / threads;
    size_t resident = by_smem < by_thr ? by_smem : by_thr;
    if (resident > 32) resident = 32;
    if ((size_t)((B + r - 1) / r) <= resident * 148u) return r;
  }
  return 4;
}

#define ZOO_RNN_LAUNCH(KERNEL, WT, ROWS, GRID, BLOCK, SM, STREAM, ARGS)                                   \
  do {                                                                                                      \
    int rc_ = ZOO_OK;                                                                                       \
    if ((ROWS) == 1) { if ((rc_ = set_smem(KERNEL<WT, 1>, SM))) return rc_; KERNEL<WT, 1><<<GRID, BLOCK, SM, STREAM>>>(ARGS); } \
    else if ((ROWS) == 2) { if ((rc_ = set_smem(KERNEL<WT, 2>, SM))) return rc_; KERNEL<WT, 2><<<GRID, BLOCK, SM, STREAM>>>(ARGS); } \
    else { if ((rc_ = set_smem(KERNEL<WT, 4>, SM))) return rc_; KERNEL<WT, 4><<<GRID, BLOCK, SM, STREAM>>>(ARGS); } \
  } while (0)

// ZOO_RNN_WARP=0 keeps the CTA-tiled forward kernels for every H (A/B measurements and the second implementation in tests).
bool rnn_warp_path() {
  const char* e = getenv("ZOO_RNN_WARP");
  return !(e && e[0] == '0');
}

// Rows per warp of the warp-per-row kernels (ZOO_RNN_RW=1/2 overrides, for measurements).
int rnn_rows_per_warp(int B) {
  const char* e = getenv("ZOO_RNN_RW");
  if (e && (e[0] == '1' || e[0] == '2')) return e[0] - '0';
  (void)B;
  return 1;            // measured: the step is a latency chain, a second row per warp does not shorten it (138 us either way)
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
  if (H <= 64 && rnn_warp_path()) {           // one warp per batch row, no barriers (see lstm_fwd_warp_kernel)
    cudaStream_t s = (cudaStream_t)stream;
    const int rw = rnn_rows_per_warp(B);
    const unsigned grid = (B + 4 * rw - 1) / (4 * rw);
    if (H <= 32) {
      if (rw == 2) lstm_fwd_warp_kernel<1, 2><<<grid, 128, 0, s>>>(a);
      else lstm_fwd_warp_kernel<1, 1><<<grid, 128, 0, s>>>(a);
    } else {
      const size_t sm = (size_t)64 * 64 * sizeof(float4);
      if (rw == 2) {
        if ((rc = set_smem(lstm_fwd_warp_kernel<2, 2>, sm))) return rc;
        lstm_fwd_warp_kernel<2, 2><<<grid, 128, sm, s>>>(a);
      } else {
        if ((rc = set_smem(lstm_fwd_warp_kernel<2, 1>, sm))) return rc;
        lstm_fwd_warp_kernel<2, 1><<<grid, 128, sm, s>>>(a);
      }
    }
    ZOO_CUDA_CHECK(cudaGetLastError());
    return ZOO_OK;
  }
  const int Hp = (H + 3) & ~3, NC = 4 * H;
  const unsigned block = (NC + 31) / 32 * 32;
  cudaStream_t s = (cudaStream_t)stream;
  const bool f32 = rnn_smem<float>(H, 4, kRnnRows * (Hp + NC)) <= kSmemMax;
  auto smem_of = [&](int r) { return f32 ? rnn_smem<float>(H, 4, r * (Hp + NC)) : rnn_smem<__nv_bfloat16>(H, 4, r * (Hp + NC)); };
  const int rows = pick_rows(B, block, smem_of);
  const unsigned grid = (B + rows - 1) / rows;
  const size_t sm = smem_of(rows);
  if (f32) ZOO_RNN_LAUNCH(lstm_fwd_kernel, float, rows, grid, block, sm, s, a);
  else ZOO_RNN_LAUNCH(lstm_fwd_kernel, __nv_bfloat16, rows, grid, block, sm, s, a);
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
  const unsigned block = (NC + 31) / 32 * 32;
  cudaStream_t s = (cudaStream_t)stream;
  const bool f32 = rnn_smem<float>(H, 4, kRnnRows * (NC + H)) <= kSmemMax;
  auto smem_of = [&](int r) { return f32 ? rnn_smem<float>(H, 4, r * (NC + H)) : rnn_smem<__nv_bfloat16>(H, 4, r * (NC + H)); };
  const int rows = pick_rows(B, block, smem_of);
  const unsigned grid = (B + rows - 1) / rows;
  const size_t sm = smem_of(rows);
  if (f32) ZOO_RNN_LAUNCH(lstm_bwd_kernel, float, rows, grid, block, sm, s, a);
  else ZOO_RNN_LAUNCH(lstm_bwd_kernel, __nv_bfloat16, rows, grid, block, sm, s, a);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_gru_forward(const float* gx, float* saved, const float* whh, const float* bhn, const float* h0,
                            float* hout, float* hT, int B, int T, int H, int hard, void* stream) {
  int rc = check_shape(B, T, H);
  if (rc) return rc;
  if (!gx || !whh || !hout) { zoo_set_last_error("gru_forward: null pointer", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  RnnArgs a{const_cast<float*>(gx), saved, whh, bhn, h0, nullptr, hout, nullptr, hT, nullptr, B, T, H, 0, hard};
  if (H <= 64 && rnn_warp_path()) {
    cudaStream_t s = (cudaStream_t)stream;
    const int rw = rnn_rows_per_warp(B);
    const unsigned grid = (B + 4 * rw - 1) / (4 * rw);
    if (H <= 32) {
      if (rw == 2) gru_fwd_warp_kernel<1, 2><<<grid, 128, 0, s>>>(a);
      else gru_fwd_warp_kernel<1, 1><<<grid, 128, 0, s>>>(a);
    } else {
      const size_t sm = (size_t)64 * 64 * sizeof(float4);
      if (rw == 2) {
        if ((rc = set_smem(gru_fwd_warp_kernel<2, 2>, sm))) return rc;
        gru_fwd_warp_kernel<2, 2><<<grid, 128, sm, s>>>(a);
      } else {
        if ((rc = set_smem(gru_fwd_warp_kernel<2, 1>, sm))) return rc;
        gru_fwd_warp_kernel<2, 1><<<grid, 128, sm, s>>>(a);
      }
    }
    ZOO_CUDA_CHECK(cudaGetLastError());
    return ZOO_OK;
  }
  const int Hp = (H + 3) & ~3, NC = 3 * H;
  const unsigned block = (4 * H + 31) / 32 * 32;
  cudaStream_t s = (cudaStream_t)stream;
  const bool f32 = rnn_smem<float>(H, 3, kRnnRows * (Hp + 2 * NC)) <= kSmemMax;
  auto smem_of = [&](int r) { return f32 ? rnn_smem<float>(H, 3, r * (Hp + 2 * NC)) : rnn_smem<__nv_bfloat16>(H, 3, r * (Hp + 2 * NC)); };
  const int rows = pick_rows(B, block, smem_of);
  const unsigned grid = (B + rows - 1) / rows;
  const size_t sm = smem_of(rows);
  if (f32) ZOO_RNN_LAUNCH(gru_fwd_kernel, float, rows, grid, block, sm, s, a);
  else ZOO_RNN_LAUNCH(gru_fwd_kernel, __nv_bfloat16, rows, grid, block, sm, s, a);
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
  const unsigned block = (4 * H + 31) / 32 * 32;
  cudaStream_t s = (cudaStream_t)stream;
  const bool f32 = rnn_smem<float>(H, 3, kRnnRows * (NC + H)) <= kSmemMax;
  auto smem_of = [&](int r) { return f32 ? rnn_smem<float>(H, 3, r * (NC + H)) : rnn_smem<__nv_bfloat16>(H, 3, r * (NC + H)); };
  const int rows = pick_rows(B, block, smem_of);
  const unsigned grid = (B + rows - 1) / rows;
  const size_t sm = smem_of(rows);
  if (f32) ZOO_RNN_LAUNCH(gru_bwd_kernel, float, rows, grid, block, sm, s, a);
  else ZOO_RNN_LAUNCH(gru_bwd_kernel, __nv_bfloat16, rows, grid, block, sm, s, a);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
