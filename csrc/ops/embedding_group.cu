// K2 / K4, grouped form: every lookup of a step in one gather launch, every id column routed to the owners of its rows in one
// launch, and the sparse backward of ALL tables in three launches (elect / accumulate duplicates / apply).
//
// Why a second form next to csrc/ops/embedding.cu: per-table launches cost the headline step twice.  (1) The four tables
// of NeuralCF use two id columns, so the election of one representative entry per touched row, the winner list and the
// duplicate bookkeeping were computed twice per column.  (2) Each of these kernels is latency-bound (random 256..512-byte
// rows, over NVLink when the table is sharded), and one launch at a time leaves the memory system and the links half
// idle; independent groups in ONE launch (blockIdx.y) keep all of their loads in flight together.  (3) With row-sharded
// tables the owner used to scan every rank's ids (world x B of them, most not its own); now each source ROUTES its entries
// in forward -- (local row, entry index) pairs pushed into the owner's mailbox with peer stores -- and the owner's passes
// run over the ~B entries it actually owns.
//
// Reference semantics replaced: LookupTable inside zoo Embedding (zoo/.../pipeline/api/keras/layers/Embedding.scala:82-101),
// the GMF product of NeuralCF.scala:83-89, and the dense all-reduce + full-table optimizer pass of embedding gradients
// (zoo/.../pipeline/api/keras/models/Topology.scala:1187-1206); optimizer math as keras/optimizers/Adam.scala:59-106 on the
// touched rows only.
#include "../common/zoo_common.cuh"
#include "zoo_emb_group.h"
#include <limits.h>

namespace zoo {

typedef ZooEgTab EgTab;
typedef ZooEgGroup EgGroup;

// ---- entry enumeration ---------------------------------------------------------------------------------------------------
// world == 1: entry i = (sample i / bag, slot i % bag) of this rank's own id matrix.
// world  > 1: the owner's entries are the concatenation over sources s of pairs[s * cap .. + counts[s]).
struct EgIndex {
  int pre[ZOO_EG_MAX_PEERS + 1];
  int total;
};

__device__ __forceinline__ EgIndex eg_index(const EgGroup& g) {
  EgIndex ix;
  if (g.world == 1) {
    ix.pre[0] = 0;
#pragma unroll
    for (int s = 1; s <= ZOO_EG_MAX_PEERS; ++s) ix.pre[s] = g.E_local;
    ix.total = g.E_local;
    return ix;
  }
  int acc = 0;
#pragma unroll
  for (int s = 0; s < ZOO_EG_MAX_PEERS; ++s) {
    ix.pre[s] = acc;
    if (s < g.world) acc += __ldg(g.counts + s);
  }
  ix.pre[ZOO_EG_MAX_PEERS] = acc;
  ix.total = acc;
  return ix;
}

// entry i -> (source rank, entry index at the source, local row); false if the id is out of range (world == 1 only)
__device__ __forceinline__ bool eg_decode(const EgGroup& g, const EgIndex& ix, int i, int& src, int& ent, int& local) {
  if (g.world == 1) {
    const int b = i / g.bag, j = i - b * g.bag;
    const long long id = g.ids[(long long)b * g.id_stride + g.id_col + j];
    src = 0; ent = i; local = (int)id;
    return id >= 0 && id < g.rows;
  }
  int s = 0, first = 0;
#pragma unroll
  for (int k = 1; k < ZOO_EG_MAX_PEERS; ++k)
    if (i >= ix.pre[k]) { s = k; first = ix.pre[k]; }
  src = s;
  const int2 pr = reinterpret_cast<const int2*>(g.pairs)[(long long)s * g.cap + (i - first)];
  local = pr.x; ent = pr.y;
  return true;
}

// 4 columns of the gradient row of (src, ent) for table t, starting at column c
__device__ __forceinline__ void eg_grad(const EgTab& t, int bag, int src, int ent, int c, float (&gr)[4]) {
  const int b = ent / bag;
  const __nv_bfloat16* gp = reinterpret_cast<const __nv_bfloat16*>(t.dout[src]) + (long long)b * t.ldg + c;
  const uint2 u = *reinterpret_cast<const uint2*>(gp);
  const float2 a = unpack_bf16(u.x), d = unpack_bf16(u.y);
  gr[0] = a.x; gr[1] = a.y; gr[2] = d.x; gr[3] = d.y;
  if (t.mul[0]) {
    const __nv_bfloat16* mp = reinterpret_cast<const __nv_bfloat16*>(t.mul[src]) + (long long)b * t.ldmul + c;
    const uint2 q = *reinterpret_cast<const uint2*>(mp);
    const float2 e = unpack_bf16(q.x), f = unpack_bf16(q.y);
    gr[0] *= e.x; gr[1] *= e.y; gr[2] *= f.x; gr[3] *= f.y;
  }
  gr[0] *= t.grad_scale; gr[1] *= t.grad_scale; gr[2] *= t.grad_scale; gr[3] *= t.grad_scale;
}

// ---- pass 1: elect the smallest entry index of every touched row ---------------------------------------------------------
__global__ void __launch_bounds__(256)
eg_mark_kernel(const __grid_constant__ ZooEgParams p) {
  pdl_wait();
  pdl_launch_dependents();
  const EgGroup& g = p.g[blockIdx.y];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    *g.win_count = 0;
    for (int k = 0; k < g.ntabs; ++k)
      if (g.t[k].bump_step) g.t[k].hyper[5] += 1.f;
  }
  const EgIndex ix = eg_index(g);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ix.total; i += gridDim.x * blockDim.x) {
    int src, ent, local;
    if (eg_decode(g, ix, i, src, ent, local)) atomicMin(g.rep_of_row + local, i);
  }
}

// ---- pass 2: winners go to the compact list, losers add their gradient rows to the winner's accumulation rows -------------
__global__ void __launch_bounds__(256)
eg_accum_kernel(const __grid_constant__ ZooEgParams p) {
  pdl_wait();
  pdl_launch_dependents();
  const EgGroup& g = p.g[blockIdx.y];
  __shared__ int s_wcount[8];
  __shared__ int s_base;
  const EgIndex ix = eg_index(g);
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  int2* wl = reinterpret_cast<int2*>(g.win_list);
  for (int base = blockIdx.x * 256; base < ix.total; base += gridDim.x * 256) {    // uniform over the CTA
    const int i = base + threadIdx.x;
    int rep_mine = -1, src_mine = 0, ent_mine = 0, local_mine = 0;
    bool win = false;
    if (i < ix.total && eg_decode(g, ix, i, src_mine, ent_mine, local_mine)) {
      const int rep = g.rep_of_row[local_mine];
      if (rep != i) rep_mine = rep; else win = true;
    }
    const unsigned wmask = __ballot_sync(0xffffffffu, win);
    if (lane == 0) s_wcount[wib] = __popc(wmask);
    __syncthreads();
    if (threadIdx.x == 0) {        // one atomic per CTA on the list counter
      int tot = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) { const int c = s_wcount[k]; s_wcount[k] = tot; tot += c; }
      s_base = tot ? atomicAdd(g.win_count, tot) : 0;
    }
    __syncthreads();
    if (win) wl[s_base + s_wcount[wib] + __popc(wmask & ((1u << lane) - 1u))] = make_int2(i, local_mine);
    unsigned losers = __ballot_sync(0xffffffffu, rep_mine >= 0);
    while (losers) {
      const int sl = __ffs(losers) - 1;
      losers &= losers - 1;
      const int rep = __shfl_sync(0xffffffffu, rep_mine, sl);
      const int src = __shfl_sync(0xffffffffu, src_mine, sl);
      const int ent = __shfl_sync(0xffffffffu, ent_mine, sl);
      for (int k = 0; k < g.ntabs; ++k) {
        const EgTab& t = g.t[k];
        float* arow = t.accum + (long long)rep * t.D;
        for (int c = lane * 4; c < t.D; c += 128) {
          float gr[4];
          eg_grad(t, g.bag, src, ent, c, gr);
          red_add_f32x4(arow + c, gr[0], gr[1], gr[2], gr[3]);
        }
      }
      if (lane < g.ntabs) g.dup_flag[(long long)lane * g.max_entries + rep] = 1;
    }
    __syncthreads();               // s_wcount / s_base are reused by the next round
  }
}

// ---- pass 3: representatives apply the optimizer to their rows -----------------------------------------------------------------
// blockIdx.y = (group, table): every table of a group walks the group's ONE winner list.  The update is latency-bound -- a row
// is a handful of 16-byte loads (gradient row, possibly from a peer over NVLink, plus w / m / v) and a little math -- so the
// kernel is organised for loads in flight: a row uses only as many lanes as it has float4 columns (G = 8 / 16 / 32, two passes
// of 32 lanes for 256-wide rows), a warp covers 32 / G rows, and all loads of a row are issued before anything is computed.
// LOOP == false: one pass per warp over a grid sized for the entries (world == 1: the entry count is known on the host; no
// register cost of the loop, CTAs at different phases overlap their load and store phases).  LOOP == true: a grid sized for
// one GPU's share of the entries strides over what the owner really received.
// Duplicate flags are per table (dup_flag + table * max_entries) so tables never race on resetting them; table 0 restores the
// row map.
template <int ITERS, bool LOOP>
__global__ void __launch_bounds__(256, ITERS == 1 ? (LOOP ? 5 : 6) : 3)
eg_apply_kernel(const __grid_constant__ ZooEgParams p) {
  pdl_wait();
  pdl_launch_dependents();
  const EgGroup& g = p.g[blockIdx.y / ZOO_EG_MAX_TABS];
  const int k = blockIdx.y % ZOO_EG_MAX_TABS;
  if (k >= g.ntabs) return;
  const EgTab& t = g.t[k];
  const int count = *g.win_count;
  const int units = t.D >> 2;
  int G = 32;
  while (G > 8 && (G >> 1) >= units) G >>= 1;
  const int lane = threadIdx.x & 31;
  const int P = 32 / G, slot = lane / G, l = lane - slot * G;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  int base = warp * P;
  if (base >= count) return;
  const EgIndex ix = eg_index(g);
  const float* h = t.hyper;
  const float lr = h[0], b1 = h[1], b2 = h[2], eps = h[3], wd = h[4];
  float bc1 = 1.f, bc2 = 1.f;
  if (t.opt == 1) { bc1 = 1.f - powf(b1, h[5]); bc2 = 1.f - powf(b2, h[5]); }
  int* dupf = g.dup_flag + (long long)k * g.max_entries;
  do {
    const int idx = base + slot;
    const bool row_ok = idx < count;
    int i = 0, local = 0, src = 0, ent = 0;
    bool dup = false;
    if (row_ok) {
      const int2 wlv = reinterpret_cast<const int2*>(g.win_list)[idx];
      i = wlv.x; local = wlv.y;
      int lc;
      eg_decode(g, ix, i, src, ent, lc);
      dup = dupf[i] != 0;
    }
    float gr[ITERS][4];
    float4 w4[ITERS], m4[ITERS], v4[ITERS], a4[ITERS];
    bool ok[ITERS];
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {                      // ---- issue every load first
      const int c = (l + it * G) << 2;
      ok[it] = row_ok && c < t.D;
      if (ok[it]) {
        eg_grad(t, g.bag, src, ent, c, gr[it]);
        const long long o = (long long)local * t.D + c;
        w4[it] = *reinterpret_cast<const float4*>(t.w + o);
        if (t.m) m4[it] = *reinterpret_cast<const float4*>(t.m + o);
        if (t.v) v4[it] = *reinterpret_cast<const float4*>(t.v + o);
        if (dup) a4[it] = *reinterpret_cast<const float4*>(t.accum + (long long)i * t.D + c);
      }
    }
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {                      // ---- then compute and store
      if (!ok[it]) continue;
      const int c = (l + it * G) << 2;
      float w[4] = {w4[it].x, w4[it].y, w4[it].z, w4[it].w};
      float m[4] = {0.f, 0.f, 0.f, 0.f}, v[4] = {0.f, 0.f, 0.f, 0.f};
      if (t.m) { m[0] = m4[it].x; m[1] = m4[it].y; m[2] = m4[it].z; m[3] = m4[it].w; }
      if (t.v) { v[0] = v4[it].x; v[1] = v4[it].y; v[2] = v4[it].z; v[3] = v4[it].w; }
      if (dup) {
        gr[it][0] += a4[it].x; gr[it][1] += a4[it].y; gr[it][2] += a4[it].z; gr[it][3] += a4[it].w;
        *reinterpret_cast<float4*>(t.accum + (long long)i * t.D + c) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float gg = gr[it][q] + wd * w[q];
        if (t.opt == 1) {
          m[q] = b1 * m[q] + (1.f - b1) * gg;
          v[q] = b2 * v[q] + (1.f - b2) * gg * gg;
          w[q] -= lr * (m[q] / bc1) / (sqrtf(v[q] / bc2) + eps);
        } else if (t.opt == 2) {
          m[q] += gg * gg;
          w[q] -= lr * gg / (sqrtf(m[q]) + eps);
        } else {
          w[q] -= lr * gg;
        }
      }
      const long long o = (long long)local * t.D + c;
      *reinterpret_cast<float4*>(t.w + o) = make_float4(w[0], w[1], w[2], w[3]);
      if (t.m) *reinterpret_cast<float4*>(t.m + o) = make_float4(m[0], m[1], m[2], m[3]);
      if (t.v) *reinterpret_cast<float4*>(t.v + o) = make_float4(v[0], v[1], v[2], v[3]);
      if (t.w16) {
        uint2 pk; pk.x = pack_bf16(w[0], w[1]); pk.y = pack_bf16(w[2], w[3]);
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(t.w16) + o) = pk;
      }
    }
    if (row_ok && l == 0) {                                   // one lane per row restores the scratch
      if (k == 0) g.rep_of_row[local] = INT_MAX;
      if (dup) dupf[i] = 0;
    }
    base += nwarps * P;
  } while (LOOP && base < count);
}

// ---- forward: every lookup of the step in one launch ----------------------------------------------------------------------
template <typename T> struct EgVec;
template <> struct EgVec<float> { static constexpr int N = 4; };
template <> struct EgVec<__nv_bfloat16> { static constexpr int N = 8; };

template <typename T>
__device__ __forceinline__ void eg_load_vec(const T* p, float (&v)[EgVec<T>::N]);
template <>
__device__ __forceinline__ void eg_load_vec<float>(const float* p, float (&v)[4]) {
  const float4 f = *reinterpret_cast<const float4*>(p);
  v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
}
template <>
__device__ __forceinline__ void eg_load_vec<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y), c = unpack_bf16(u.z), d = unpack_bf16(u.w);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}

template <typename T>
__device__ __forceinline__ const T* eg_row(const void* const (&tabs)[ZOO_EG_MAX_PEERS], long long id, int world, int D) {
  if (world == 1) return reinterpret_cast<const T*>(tabs[0]) + id * (long long)D;
  return reinterpret_cast<const T*>(tabs[(int)(id % world)]) + (id / world) * (long long)D;
}

__device__ __forceinline__ uint32_t eg_mul_bf16x2(uint32_t a, uint32_t b) {
  __nv_bfloat162 x = *reinterpret_cast<__nv_bfloat162*>(&a), y = *reinterpret_cast<__nv_bfloat162*>(&b);
  __nv_bfloat162 z = __hmul2(x, y);          // fp32 product rounded once to bf16: the same value as multiply-then-pack
  return *reinterpret_cast<uint32_t*>(&z);
}

// bag == 1 (one row, or the product of two rows, per sample): no accumulator, no conversion for bf16 rows -- the kernel is a
// latency-bound copy, so the fewer registers it needs the more rows are in flight.
template <typename T>
__device__ __forceinline__ void eg_gather_one(const ZooEgGatherParams& p, const ZooEgSpec& s, int sample, int lane, int G) {
  constexpr int VN = EgVec<T>::N;
  const long long* idr = p.ids + (long long)sample * p.id_stride;
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(s.out) + (long long)sample * s.ldo;
  const bool mul = s.tb[0] != nullptr;
  const long long ia = idr[s.id_col];
  const bool oka = ia >= 0 && ia < s.rows_a;
  long long ib = 0;
  bool okb = false;
  if (mul) { ib = idr[s.id_col2]; okb = ib >= 0 && ib < s.rows_b; }
  const T* ra = oka ? eg_row<T>(s.ta, ia, p.world, s.D) : nullptr;
  const T* rb = okb ? eg_row<T>(s.tb, ib, p.world, s.D) : nullptr;
  for (int c = lane * VN; c < s.D; c += G * VN) {
    if (VN == 8) {
      uint4 a = make_uint4(0, 0, 0, 0), b = make_uint4(0, 0, 0, 0);
      if (ra) a = *reinterpret_cast<const uint4*>(ra + c);
      if (mul) {
        if (rb) b = *reinterpret_cast<const uint4*>(rb + c);
        if (s.save_a) {
          *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(s.save_a) + (long long)sample * s.D + c) = a;
          *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(s.save_b) + (long long)sample * s.D + c) = b;
        }
        a.x = eg_mul_bf16x2(a.x, b.x); a.y = eg_mul_bf16x2(a.y, b.y); a.z = eg_mul_bf16x2(a.z, b.z); a.w = eg_mul_bf16x2(a.w, b.w);
      }
      *reinterpret_cast<uint4*>(o + c) = a;
    } else {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ra) a = *reinterpret_cast<const float4*>(ra + c);
      if (mul) {
        if (rb) b = *reinterpret_cast<const float4*>(rb + c);
        if (s.save_a) {
          uint2 ua, ub;
          ua.x = pack_bf16(a.x, a.y); ua.y = pack_bf16(a.z, a.w);
          ub.x = pack_bf16(b.x, b.y); ub.y = pack_bf16(b.z, b.w);
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(s.save_a) + (long long)sample * s.D + c) = ua;
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(s.save_b) + (long long)sample * s.D + c) = ub;
        }
        a.x *= b.x; a.y *= b.y; a.z *= b.z; a.w *= b.w;
      }
      uint2 pk;
      pk.x = pack_bf16(a.x, a.y); pk.y = pack_bf16(a.z, a.w);
      *reinterpret_cast<uint2*>(o + c) = pk;
    }
  }
}

template <typename T>
__device__ __noinline__ void eg_gather_body(const ZooEgGatherParams& p, const ZooEgSpec& s, int sample, int lane, int G) {
  constexpr int VN = EgVec<T>::N;
  const long long* idr = p.ids + (long long)sample * p.id_stride;
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(s.out) + (long long)sample * s.ldo;
  const bool mul = s.tb[0] != nullptr;
  for (int c = lane * VN; c < s.D; c += G * VN) {
    float acc[VN];
#pragma unroll
    for (int q = 0; q < VN; ++q) acc[q] = 0.f;
    int cnt = 0;
    for (int j = 0; j < s.bag; ++j) {
      const long long ia = idr[s.id_col + j];
      if (ia < 0 || ia >= s.rows_a) continue;
      float va[VN];
      eg_load_vec<T>(eg_row<T>(s.ta, ia, p.world, s.D) + c, va);
      if (mul) {
        const long long ib = idr[s.id_col2 + j];
        float vb[VN];
        if (ib >= 0 && ib < s.rows_b) {
          eg_load_vec<T>(eg_row<T>(s.tb, ib, p.world, s.D) + c, vb);
        } else {
#pragma unroll
          for (int q = 0; q < VN; ++q) vb[q] = 0.f;
        }
        if (s.save_a) {       // both factors (bf16) for the product rule in backward
          __nv_bfloat16* pa = reinterpret_cast<__nv_bfloat16*>(s.save_a) + (long long)sample * s.D + c;
          __nv_bfloat16* pb = reinterpret_cast<__nv_bfloat16*>(s.save_b) + (long long)sample * s.D + c;
#pragma unroll
          for (int q = 0; q < VN; q += 4) {
            uint2 ua, ub;
            ua.x = pack_bf16(va[q], va[q + 1]); ua.y = pack_bf16(va[q + 2], va[q + 3]);
            ub.x = pack_bf16(vb[q], vb[q + 1]); ub.y = pack_bf16(vb[q + 2], vb[q + 3]);
            *reinterpret_cast<uint2*>(pa + q) = ua;
            *reinterpret_cast<uint2*>(pb + q) = ub;
          }
        }
#pragma unroll
        for (int q = 0; q < VN; ++q) va[q] *= vb[q];
      }
#pragma unroll
      for (int q = 0; q < VN; ++q) acc[q] += va[q];
      ++cnt;
    }
    if (s.mean && cnt > 1) {
      const float sc = 1.f / cnt;
#pragma unroll
      for (int q = 0; q < VN; ++q) acc[q] *= sc;
    }
    if (VN == 8) {
      uint4 pk;
      pk.x = pack_bf16(acc[0], acc[1]); pk.y = pack_bf16(acc[2], acc[3]);
      pk.z = pack_bf16(acc[4 % VN], acc[5 % VN]); pk.w = pack_bf16(acc[6 % VN], acc[7 % VN]);
      *reinterpret_cast<uint4*>(o + c) = pk;
    } else {
      uint2 pk;
      pk.x = pack_bf16(acc[0], acc[1]); pk.y = pack_bf16(acc[2], acc[3]);
      *reinterpret_cast<uint2*>(o + c) = pk;
    }
  }
}

struct EgLanes { int G[ZOO_EG_MAX_SPECS]; };

__global__ void __launch_bounds__(256, 6)
eg_gather_kernel(const __grid_constant__ ZooEgGatherParams p, const EgLanes lanes) {
  pdl_wait();
  pdl_launch_dependents();
  const ZooEgSpec& s = p.s[blockIdx.y];
  const int G = lanes.G[blockIdx.y];
  const int gl = threadIdx.x & 31;
  const int sample = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 / G) + gl / G;
  const int lane = gl - (gl / G) * G;
  if (sample >= p.B) return;
  if (s.bag == 1) {
    if (s.dtype == 0) eg_gather_one<float>(p, s, sample, lane, G);
    else eg_gather_one<__nv_bfloat16>(p, s, sample, lane, G);
  } else {
    if (s.dtype == 0) eg_gather_body<float>(p, s, sample, lane, G);
    else eg_gather_body<__nv_bfloat16>(p, s, sample, lane, G);
  }
}

// ---- forward, world > 1: route every entry of an id column to the owner of its row ---------------------------------------
__global__ void __launch_bounds__(256)
eg_route_kernel(const __grid_constant__ ZooEgRouteParams p) {
  pdl_wait();
  pdl_launch_dependents();
  const ZooEgRouteGroup& g = p.g[blockIdx.y];
  __shared__ int s_cnt[ZOO_EG_MAX_PEERS];
  __shared__ int s_last;
  const int t = threadIdx.x;
  if (t < ZOO_EG_MAX_PEERS) s_cnt[t] = 0;
  __syncthreads();
  const int E = p.B * g.bag;
  const int e = blockIdx.x * 256 + t;
  int o = -1, mypos = 0, local = 0;
  if (e < E) {
    const int b = e / g.bag, j = e - b * g.bag;
    const long long id = p.ids[(long long)b * p.id_stride + g.id_col + j];
    if (id >= 0 && id < g.rows) {
      o = (int)(id % p.world);
      local = (int)(id / p.world);
      mypos = atomicAdd(&s_cnt[o], 1);
    }
  }
  __syncthreads();
  if (t < p.world) {
    const int c = s_cnt[t];
    s_cnt[t] = c ? atomicAdd(g.counters + t, c) : 0;
  }
  __syncthreads();
  if (o >= 0) {
    const int pos = s_cnt[o] + mypos;
    if (pos >= p.cap) { printf("zoo eg_route: mailbox overflow (rank %d -> %d)\n", p.rank, o); __trap(); }
    reinterpret_cast<int2*>(g.pairs[o])[(long long)p.rank * p.cap + pos] = make_int2(local, e);
  }
  // the last CTA of this group publishes the per-owner counts and restores the local counters
  __threadfence();
  __syncthreads();
  if (t == 0) s_last = (atomicAdd(g.counters + ZOO_EG_MAX_PEERS, 1) == (int)gridDim.x - 1) ? 1 : 0;
  __syncthreads();
  if (s_last) {
    __threadfence();
    if (t < p.world) {
      const int c = atomicExch(g.counters + t, 0);
      g.counts[t][p.rank] = c;
    }
    if (t == 0) g.counters[ZOO_EG_MAX_PEERS] = 0;
    __threadfence_system();
  }
}

static int g_eg_sms = 0;
static int eg_sms() {
  if (!g_eg_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_eg_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_eg_sms <= 0) g_eg_sms = 148;
  }
  return g_eg_sms;
}

}  // namespace zoo

using namespace zoo;

ZOO_API int zoo_eg_gather(const ZooEgGatherParams* p, void* stream) {
  if (!p || !p->ids || p->nspecs < 1 || p->nspecs > ZOO_EG_MAX_SPECS || p->world < 1 || p->world > ZOO_EG_MAX_PEERS || p->B < 0) {
    zoo_set_last_error("eg_gather: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (p->B == 0) return ZOO_OK;
  EgLanes lanes{};
  long long max_blocks = 1;
  for (int k = 0; k < p->nspecs; ++k) {
    const ZooEgSpec& s = p->s[k];
    const int vn = s.dtype == 0 ? 4 : 8;
    const bool mul = s.tb[0] != nullptr;
    bool ok = s.out && s.D > 0 && s.bag >= 1 && (s.D % vn == 0) && (s.ldo % vn == 0) && (((uintptr_t)s.out & (vn * 2 - 1)) == 0) &&
              ((s.save_a != nullptr) == (s.save_b != nullptr)) && (!s.save_a || (mul && s.bag == 1));
    for (int r = 0; r < p->world && ok; ++r) {
      if (!s.ta[r] || ((uintptr_t)s.ta[r] & 15)) ok = false;
      if (mul && (!s.tb[r] || ((uintptr_t)s.tb[r] & 15))) ok = false;
    }
    if (!ok) {
      zoo_set_last_error("eg_gather: a lookup needs D % (16 bytes) == 0, aligned windows and complete shard pointer tables", __FILE__, __LINE__);
      return ZOO_ERR_ARG;
    }
    const int units = (s.D + vn - 1) / vn;
    int G = 32;
    while (G > 1 && G / 2 >= units) G /= 2;
    lanes.G[k] = G;
    const long long warps = ((long long)p->B + (32 / G) - 1) / (32 / G);
    const long long blocks = (warps * 32 + 255) / 256;
    if (blocks > max_blocks) max_blocks = blocks;
  }
  ZOO_CUDA_CHECK(zoo_launch_pdl(eg_gather_kernel, dim3((unsigned)max_blocks, (unsigned)p->nspecs), dim3(256), 0, (cudaStream_t)stream, *p, lanes));
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_eg_route(const ZooEgRouteParams* p, void* stream) {
  if (!p || !p->ids || p->ngroups < 1 || p->ngroups > ZOO_EG_MAX_GROUPS || p->world < 2 || p->world > ZOO_EG_MAX_PEERS ||
      p->rank < 0 || p->rank >= p->world || p->B < 0 || p->cap < 1) {
    zoo_set_last_error("eg_route: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  int max_e = 1;
  for (int k = 0; k < p->ngroups; ++k) {
    const ZooEgRouteGroup& g = p->g[k];
    if (!g.counters || g.bag < 1 || (long long)p->B * g.bag > p->cap || (g.rows + p->world - 1) / p->world > INT_MAX) {
      zoo_set_last_error("eg_route: cap must be >= B * bag and a shard must have fewer than 2^31 rows", __FILE__, __LINE__);
      return ZOO_ERR_ARG;
    }
    for (int r = 0; r < p->world; ++r)
      if (!g.pairs[r] || !g.counts[r]) { zoo_set_last_error("eg_route: null mailbox pointer", __FILE__, __LINE__); return ZOO_ERR_ARG; }
    if (p->B * g.bag > max_e) max_e = p->B * g.bag;
  }
  // every group's grid has the same x extent (the ticket count of the last-CTA test is gridDim.x)
  ZOO_CUDA_CHECK(zoo_launch_pdl(eg_route_kernel, dim3((unsigned)((max_e + 255) / 256), (unsigned)p->ngroups), dim3(256), 0, (cudaStream_t)stream, *p));
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_eg_update(const ZooEgParams* p, void* stream) {
  if (!p || p->ngroups < 1 || p->ngroups > ZOO_EG_MAX_GROUPS) {
    zoo_set_last_error("eg_update: bad group count", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  int max_e = 1, max_units = 1;
  for (int k = 0; k < p->ngroups; ++k) {
    const ZooEgGroup& g = p->g[k];
    bool ok = g.rep_of_row && g.dup_flag && g.win_list && g.win_count && g.ntabs >= 1 && g.ntabs <= ZOO_EG_MAX_TABS && g.bag >= 1 &&
              g.world >= 1 && g.world <= ZOO_EG_MAX_PEERS && g.rank >= 0 && g.rank < g.world && g.E_local >= 0 &&
              (g.world == 1 ? g.ids != nullptr : (g.pairs && g.counts && g.cap >= 1)) && g.max_entries >= (g.world == 1 ? g.E_local : 1) && (g.rows + g.world - 1) / g.world <= INT_MAX;
    int units = 0;
    for (int q = 0; q < g.ntabs && ok; ++q) {
      const ZooEgTab& t = g.t[q];
      ok = t.w && t.accum && t.hyper && t.D > 0 && (t.D % 4 == 0) && (t.ldg % 4 == 0) && (((uintptr_t)t.w & 15) == 0) &&
           (((uintptr_t)t.accum & 15) == 0) && (!t.m || ((uintptr_t)t.m & 15) == 0) && (!t.v || ((uintptr_t)t.v & 15) == 0) &&
           (!t.w16 || ((uintptr_t)t.w16 & 7) == 0) && !(t.opt == 1 && (!t.m || !t.v)) && !(t.opt == 2 && !t.m);
      for (int r = 0; r < g.world && ok; ++r) {
        if (!t.dout[r] || ((uintptr_t)t.dout[r] & 7)) ok = false;
        if (t.mul[0] && (!t.mul[r] || ((uintptr_t)t.mul[r] & 7) || (t.ldmul % 4))) ok = false;
      }
      units += t.D / 4;
    }
    if (!ok || units > 128) {
      zoo_set_last_error("eg_update: bad group (rows must be multiples of 4 floats, 16-byte aligned, <= 512 floats per group; "
                         "optimizer state must match the optimizer)", __FILE__, __LINE__);
      return ZOO_ERR_ARG;
    }
    if (g.E_local > max_e) max_e = g.E_local;
    if (units > max_units) max_units = units;
  }
  cudaStream_t s = (cudaStream_t)stream;
  const int sms = eg_sms();
  // entry passes: enough threads for one GPU's share of the entries, striding over whatever the owner really received
  unsigned eb = (unsigned)((max_e + 255) / 256);
  if (eb > (unsigned)sms * 16) eb = (unsigned)sms * 16;
  ZOO_CUDA_CHECK(zoo_launch_pdl(eg_mark_kernel, dim3(eb, (unsigned)p->ngroups), dim3(256), 0, s, *p));
  ZOO_CUDA_CHECK(zoo_launch_pdl(eg_accum_kernel, dim3(eb, (unsigned)p->ngroups), dim3(256), 0, s, *p));
  // apply: blockIdx.y = (group, table); G lanes per row from the table's own width
  int max_d = 4;
  bool any_sharded = false;
  for (int k = 0; k < p->ngroups; ++k) {
    for (int q = 0; q < p->g[k].ntabs; ++q) if (p->g[k].t[q].D > max_d) max_d = p->g[k].t[q].D;
    if (p->g[k].world > 1) any_sharded = true;
  }
  const long long rows_per_cta = 8;                             // 8 warps x >= 1 row
  long long ab = ((long long)max_e + rows_per_cta - 1) / rows_per_cta;
  // sharded: the owner's entry count is only known on the device; a few waves of CTAs stride over the winner list (measured on
  // 2 GPUs: a grid of one pass per warp, 64 CTAs per SM, was 4 % slower per step than 16 CTAs per SM with the loop)
  const bool loop = any_sharded || ab > (long long)sms * 64;
  if (loop) { const long long capb = (long long)sms * 16; if (ab > capb) ab = capb; }
  if (ab < 1) ab = 1;
  const dim3 grid((unsigned)ab, (unsigned)(p->ngroups * ZOO_EG_MAX_TABS));
  if (max_d <= 128) {
    if (loop) ZOO_CUDA_CHECK(zoo_launch_pdl(eg_apply_kernel<1, true>, grid, dim3(256), 0, s, *p));
    else      ZOO_CUDA_CHECK(zoo_launch_pdl(eg_apply_kernel<1, false>, grid, dim3(256), 0, s, *p));
  } else {
    ZOO_CUDA_CHECK(zoo_launch_pdl(eg_apply_kernel<2, true>, grid, dim3(256), 0, s, *p));
  }
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

// sizeof of the parameter blocks, so that a binding in another language can verify its mirror of include/zoo_emb_group.h
ZOO_API int zoo_eg_sizes(int* out6) {
  if (!out6) return ZOO_ERR_ARG;
  out6[0] = (int)sizeof(ZooEgTab); out6[1] = (int)sizeof(ZooEgGroup); out6[2] = (int)sizeof(ZooEgParams);
  out6[3] = (int)sizeof(ZooEgSpec); out6[4] = (int)sizeof(ZooEgGatherParams); out6[5] = (int)sizeof(ZooEgRouteParams);
  return ZOO_OK;
}
