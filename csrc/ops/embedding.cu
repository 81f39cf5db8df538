// K2 / K4: embedding lookup and its sparse backward on sm_100a.
//
// Forward (K2) is a ONE-SIDED gather: table rows are sharded across GPUs (row r lives on
// GPU r % world at local row r / world) and every GPU reads the rows it needs straight from
// the owner's HBM through NVLink peer pointers.  Index dispatch, row gather and the combine
// all-to-all of the classic three-step design collapse into one kernel: the only bytes that
// cross NVLink are the B x D result rows, and no index traffic is sent at all.  Optional
// epilogues write into a column window of a wider activation buffer (fused concat) and
// multiply two gathered rows (the GMF branch of NeuralCF).
//
// Backward (K4): gradient rows are routed to the owner's mailbox with P2P stores, the owner
// de-duplicates rows without sorting (atomicMin election of one representative entry per row,
// fp32 vector reductions only for the duplicates) and applies a fused sparse Adam / Adagrad /
// SGD to the touched rows only.
//
// Replaces LookupTable / LookupTableSparse inside zoo Embedding / SparseEmbedding
// (zoo/.../pipeline/api/keras/layers/Embedding.scala:82-101, SparseEmbedding.scala:76-89),
// the GMF merge of NeuralCF.scala:83-89, and the dense-table all-reduce + dense optimizer
// pass the reference performs for embedding gradients (Topology.scala:1187-1206).
#include "../common/zoo_common.cuh"
#include <limits.h>
#include <stdlib.h>

namespace zoo {

constexpr int kMaxPeers = 8;

struct ShardPtrs {
  const void* p[kMaxPeers];
};
struct ShardPtrsMut {
  void* p[kMaxPeers];
};

template <typename T> struct Vec;  // 16-byte vector traits
template <> struct Vec<float> { static constexpr int N = 4; };
template <> struct Vec<__nv_bfloat16> { static constexpr int N = 8; };

template <typename T>
__device__ __forceinline__ void load_vec(const T* p, float (&v)[Vec<T>::N]);
template <>
__device__ __forceinline__ void load_vec<float>(const float* p, float (&v)[4]) {
  const float4 f = *reinterpret_cast<const float4*>(p);
  v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
}
template <>
__device__ __forceinline__ void load_vec<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y), c = unpack_bf16(u.z), d = unpack_bf16(u.w);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ float to_f(float x) { return x; }
__device__ __forceinline__ float to_f(__nv_bfloat16 x) { return __bfloat162float(x); }

// Select entry `i` of a small pointer table that lives in kernel-parameter space.  A switch keeps the table in the
// constant bank; indexing it with a run-time value would make the compiler copy it to local memory first.
template <typename P>
__device__ __forceinline__ P pick8(P const (&arr)[kMaxPeers], int i) {
  switch (i) {
    case 0: return arr[0]; case 1: return arr[1]; case 2: return arr[2]; case 3: return arr[3];
    case 4: return arr[4]; case 5: return arr[5]; case 6: return arr[6]; default: return arr[7];
  }
}

template <typename T>
__device__ __forceinline__ const T* row_ptr(const ShardPtrs& s, long long id, int world, int D) {
  if (world == 1) return reinterpret_cast<const T*>(s.p[0]) + id * (long long)D;
  const int owner = (int)(id % world);
  const long long local = id / world;
  return reinterpret_cast<const T*>(pick8(s.p, owner)) + local * (long long)D;
}

// One warp per (sample, slot). ids: [B, id_stride] int64, column id_col (and id_col2 for the
// second table when MUL).  Rows whose id is out of [0, total_rows) produce zeros.
// bag > 1: ids[b, id_col .. id_col+bag) are pooled (sum, or mean when mean != 0), negative ids skipped.
// G lanes serve one sample (G = 32 unless the row is narrower than a warp's worth of 16-byte vectors), so narrow rows put
// several samples in one warp and keep every lane loading.
template <typename T, bool MUL, bool VEC>
__global__ void __launch_bounds__(256, MUL ? 4 : 6)
embedding_gather_kernel(const long long* __restrict__ ids, int id_stride, int id_col, int id_col2,
                        int B, int bag, int mean, ShardPtrs ta, long long rows_a, ShardPtrs tb,
                        long long rows_b, int world, int D, __nv_bfloat16* __restrict__ out,
                        long long ldo, __nv_bfloat16* __restrict__ save_a, __nv_bfloat16* __restrict__ save_b, int G) {
  pdl_wait();
  pdl_launch_dependents();
  const int gl = threadIdx.x & 31;
  const int warp = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (32 / G) + gl / G;   // = sample index
  const int lane = gl - (gl / G) * G;
  if (warp >= B) return;
  const long long* idr = ids + (long long)warp * id_stride;
  __nv_bfloat16* o = out + (long long)warp * ldo;
  constexpr int VN = Vec<T>::N;
  if (VEC) {
    for (int c = lane * VN; c < D; c += G * VN) {
      float acc[VN];
#pragma unroll
      for (int t = 0; t < VN; ++t) acc[t] = 0.f;
      int cnt = 0;
      for (int j = 0; j < bag; ++j) {
        const long long ia = idr[id_col + j];
        if (ia < 0 || ia >= rows_a) continue;
        float va[VN];
        load_vec<T>(row_ptr<T>(ta, ia, world, D) + c, va);
        if (MUL) {
          const long long ib = idr[id_col2 + j];
          float vb[VN];
          if (ib >= 0 && ib < rows_b) {
            load_vec<T>(row_ptr<T>(tb, ib, world, D) + c, vb);
          } else {
#pragma unroll
            for (int t = 0; t < VN; ++t) vb[t] = 0.f;
          }
          if (save_a) {  // keep both factors (bf16) for the product rule in backward
            __nv_bfloat16* pa = save_a + (long long)warp * D + c;
            __nv_bfloat16* pb = save_b + (long long)warp * D + c;
#pragma unroll
            for (int t = 0; t < VN; t += 4) {
              uint2 ua, ub;
              ua.x = pack_bf16(va[t], va[t + 1]); ua.y = pack_bf16(va[t + 2], va[t + 3]);
              ub.x = pack_bf16(vb[t], vb[t + 1]); ub.y = pack_bf16(vb[t + 2], vb[t + 3]);
              *reinterpret_cast<uint2*>(pa + t) = ua;
              *reinterpret_cast<uint2*>(pb + t) = ub;
            }
          }
#pragma unroll
          for (int t = 0; t < VN; ++t) va[t] *= vb[t];
        }
#pragma unroll
        for (int t = 0; t < VN; ++t) acc[t] += va[t];
        ++cnt;
      }
      if (mean && cnt > 1) {
        const float s = 1.f / cnt;
#pragma unroll
        for (int t = 0; t < VN; ++t) acc[t] *= s;
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
  } else {
    for (int c = lane; c < D; c += G) {
      float acc = 0.f;
      int cnt = 0;
      for (int j = 0; j < bag; ++j) {
        const long long ia = idr[id_col + j];
        if (ia < 0 || ia >= rows_a) continue;
        float va = to_f(row_ptr<T>(ta, ia, world, D)[c]);
        if (MUL) {
          const long long ib = idr[id_col2 + j];
          const float vb = (ib >= 0 && ib < rows_b) ? to_f(row_ptr<T>(tb, ib, world, D)[c]) : 0.f;
          if (save_a) {
            save_a[(long long)warp * D + c] = __float2bfloat16(va);
            save_b[(long long)warp * D + c] = __float2bfloat16(vb);
          }
          va *= vb;
        }
        acc += va;
        ++cnt;
      }
      if (mean && cnt > 1) acc /= cnt;
      o[c] = __float2bfloat16(acc);
    }
  }
}

// ---------------------------------------------------------------- backward ---
// Entry e of a table's backward = (row id, gradient row).  The gradient row is
//   dOut[b, 0:D] (bf16, leading dim ldg)           optionally times
//   mul[b, 0:D]  (bf16, leading dim ldmul)         (the other factor saved by the GMF forward).
// With nsrc > 1 the entries of `nsrc` GPUs are visited as one list (entry e belongs to source
// e / E_local): ids, gradient rows and multiplier rows of the peers are read in place over NVLink,
// so routing gradient rows to the owner (the backward all-to-all) happens inside these kernels.
struct BwdEntries {
  const long long* ids[kMaxPeers]; int id_stride; int id_col; int bag;  // E_local = B * bag entries per source
  const __nv_bfloat16* dout[kMaxPeers]; long long ldg;
  const __nv_bfloat16* mul[kMaxPeers]; long long ldmul;                 // optional elementwise multiplier rows
  float grad_scale;                                                     // e.g. 1/bag for mean pooling
  int nsrc; int E_local;
};

__device__ __forceinline__ void entry_split(const BwdEntries& en, int e, int& s, int& b, int& j) {
  s = 0;
  int el = e;
  if (en.nsrc > 1) { s = e / en.E_local; el = e - s * en.E_local; }
  b = el / en.bag;
  j = el - b * en.bag;
}

template <typename P>
__device__ __forceinline__ P pick(P const (&arr)[kMaxPeers], int s) {
  // a switch keeps the pointer table in the constant bank (no local-memory copy of the kernel parameters)
  switch (s) {
    case 0: return arr[0]; case 1: return arr[1]; case 2: return arr[2]; case 3: return arr[3];
    case 4: return arr[4]; case 5: return arr[5]; case 6: return arr[6]; default: return arr[7];
  }
}

__device__ __forceinline__ long long entry_row(const BwdEntries& en, int e) {
  int s, b, j;
  entry_split(en, e, s, b, j);
  return pick(en.ids, s)[(long long)b * en.id_stride + en.id_col + j];
}

// Pass 1: elect, per touched row, the smallest entry index as its representative.
__global__ void __launch_bounds__(256)
emb_mark_kernel(BwdEntries en, int E, long long rows, int world, int rank, int* __restrict__ rep_of_row,
                int* __restrict__ win_count, float* __restrict__ step_slot) {
  pdl_wait();
  pdl_launch_dependents();
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e == 0 && win_count) *win_count = 0;     // the winner list is rebuilt by pass 2
  if (e == 0 && step_slot) *step_slot += 1.f;  // optimizer step counter (bias correction) kept on the device
  if (e >= E) return;
  const long long id = entry_row(en, e);
  if (id < 0 || id >= rows) return;
  if (world > 1 && (int)(id % world) != rank) return;
  const long long local = world > 1 ? id / world : id;
  atomicMin(rep_of_row + local, e);
}

template <bool VEC>
__device__ __forceinline__ void entry_grad(const BwdEntries& en, int e, int c, float (&g)[4]) {
  int s, b, j;
  entry_split(en, e, s, b, j);
  const __nv_bfloat16* gp = pick(en.dout, s) + (long long)b * en.ldg + c;
  if (VEC) {
    const uint2 u = *reinterpret_cast<const uint2*>(gp);
    const float2 a = unpack_bf16(u.x), d = unpack_bf16(u.y);
    g[0] = a.x; g[1] = a.y; g[2] = d.x; g[3] = d.y;
  } else {
    g[0] = __bfloat162float(gp[0]); g[1] = g[2] = g[3] = 0.f;
  }
  if (en.mul[0]) {
    const __nv_bfloat16* mp = pick(en.mul, s) + (long long)b * en.ldmul + c;
    if (VEC) {
      const uint2 u = *reinterpret_cast<const uint2*>(mp);
      const float2 a = unpack_bf16(u.x), d = unpack_bf16(u.y);
      g[0] *= a.x; g[1] *= a.y; g[2] *= d.x; g[3] *= d.y;
    } else {
      g[0] *= __bfloat162float(mp[0]);
    }
  }
  g[0] *= en.grad_scale; g[1] *= en.grad_scale; g[2] *= en.grad_scale; g[3] *= en.grad_scale;
}

// Pass 2: entries that lost the election add their gradient into the representative's
// accumulation row (fp32, vector red) and flag it.  A warp scans 32 entries at once (one per
// lane) and then cooperates on each loser, so tables without duplicates cost one coalesced
// pass over ids + rep_of_row and nothing else.
template <bool VEC>
__global__ void __launch_bounds__(256)
emb_accum_kernel(BwdEntries en, int E, long long rows, int world, int rank, const int* __restrict__ rep_of_row,
                 int D, float* __restrict__ accum, int* __restrict__ dup_flag, int2* __restrict__ win_list,
                 int* __restrict__ win_count) {
  pdl_wait();
  pdl_launch_dependents();
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int e_mine = warp_global * 32 + lane;
  int rep_mine = -1;
  bool win = false;
  int local_mine = 0;
  if (e_mine < E) {
    const long long id = entry_row(en, e_mine);
    if (id >= 0 && id < rows && !(world > 1 && (int)(id % world) != rank)) {
      const long long local = world > 1 ? id / world : id;
      const int rep = rep_of_row[local];
      if (rep != e_mine) rep_mine = rep;
      else { win = true; local_mine = (int)local; }
    }
  }
  if (win_list) {
    // Representatives are appended to a compact (entry, local row) list so that pass 3 launches work only for
    // them and never re-reads ids (which may sit on a peer) or the row map.  One atomic per CTA on the list counter:
    // thousands of warps hitting a single address serialise in L2 and used to dominate this kernel.
    __shared__ int s_wcount[8];
    __shared__ int s_base;
    const unsigned wmask = __ballot_sync(0xffffffffu, win);
    const int wib = threadIdx.x >> 5;
    if (lane == 0) s_wcount[wib] = __popc(wmask);
    __syncthreads();
    if (threadIdx.x == 0) {
      int tot = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) { const int c = s_wcount[i]; s_wcount[i] = tot; tot += c; }
      s_base = tot ? atomicAdd(win_count, tot) : 0;
    }
    __syncthreads();
    if (win) win_list[s_base + s_wcount[wib] + __popc(wmask & ((1u << lane) - 1u))] = make_int2(e_mine, local_mine);
  }
  unsigned losers = __ballot_sync(0xffffffffu, rep_mine >= 0);
  while (losers) {
    const int src = __ffs(losers) - 1;
    losers &= losers - 1;
    const int e = warp_global * 32 + src;
    const int rep = __shfl_sync(0xffffffffu, rep_mine, src);
    float* arow = accum + (long long)rep * D;
    if (VEC) {
      for (int c = lane * 4; c < D; c += 128) {
        float g[4];
        entry_grad<true>(en, e, c, g);
        red_add_f32x4(arow + c, g[0], g[1], g[2], g[3]);
      }
    } else {
      for (int c = lane; c < D; c += 32) {
        float g[4];
        entry_grad<false>(en, e, c, g);
        atomicAdd(arow + c, g[0]);
      }
    }
    if (lane == 0) dup_flag[rep] = 1;
  }
}

enum SparseOpt : int { SOPT_SGD = 0, SOPT_ADAM = 1, SOPT_ADAGRAD = 2 };

// hyper[] layout (device memory, fp32): lr, beta1, beta2, eps, weight_decay, step (as float)
struct SparseState {
  float* w;                 // fp32 master rows  [rows_local, D]
  __nv_bfloat16* w16;       // optional bf16 shadow read by the forward gather (may be null)
  float* m;                 // Adam first moment / Adagrad accumulator
  float* v;                 // Adam second moment (null otherwise)
};

// Pass 3: representatives apply the optimizer to their row and reset the scratch they used.
template <bool VEC>
__global__ void __launch_bounds__(256)
emb_apply_kernel(BwdEntries en, int E, long long rows, int world, int rank, int* __restrict__ rep_of_row,
                 int D, float* __restrict__ accum, int* __restrict__ dup_flag, SparseState st, int opt,
                 const float* __restrict__ hyper, const int2* __restrict__ win_list, const int* __restrict__ win_count) {
  pdl_wait();
  pdl_launch_dependents();
  int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (e >= E) return;
  long long local;
  if (win_list) {                                  // compact list of representatives from pass 2
    if (e >= *win_count) return;
    const int2 wl = win_list[e];
    e = wl.x;
    local = wl.y;
  } else {
    const long long id = entry_row(en, e);
    if (id < 0 || id >= rows) return;
    if (world > 1 && (int)(id % world) != rank) return;
    local = world > 1 ? id / world : id;
    if (rep_of_row[local] != e) return;
  }
  const bool dup = dup_flag[e] != 0;
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], step = hyper[5];
  float bc1 = 1.f, bc2 = 1.f;
  if (opt == SOPT_ADAM) {
    bc1 = 1.f - powf(b1, step);
    bc2 = 1.f - powf(b2, step);
  }
  float* wrow = st.w + local * (long long)D;
  float* arow = accum + (long long)e * D;
  constexpr int STEP = VEC ? 4 : 1;
  for (int c = lane * STEP; c < D; c += 32 * STEP) {
    float g[4];
    entry_grad<VEC>(en, e, c, g);
    float w[4], m[4] = {0, 0, 0, 0}, v[4] = {0, 0, 0, 0};
    if (VEC) {
      if (dup) {
        const float4 a = *reinterpret_cast<const float4*>(arow + c);
        g[0] += a.x; g[1] += a.y; g[2] += a.z; g[3] += a.w;
        *reinterpret_cast<float4*>(arow + c) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      const float4 w4 = *reinterpret_cast<const float4*>(wrow + c);
      w[0] = w4.x; w[1] = w4.y; w[2] = w4.z; w[3] = w4.w;
      if (st.m) { const float4 t = *reinterpret_cast<const float4*>(st.m + local * (long long)D + c); m[0] = t.x; m[1] = t.y; m[2] = t.z; m[3] = t.w; }
      if (st.v) { const float4 t = *reinterpret_cast<const float4*>(st.v + local * (long long)D + c); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
    } else {
      if (dup) { g[0] += arow[c]; arow[c] = 0.f; }
      w[0] = wrow[c];
      if (st.m) m[0] = st.m[local * (long long)D + c];
      if (st.v) v[0] = st.v[local * (long long)D + c];
    }
#pragma unroll
    for (int t = 0; t < STEP; ++t) {
      float gg = g[t] + wd * w[t];
      if (opt == SOPT_ADAM) {
        m[t] = b1 * m[t] + (1.f - b1) * gg;
        v[t] = b2 * v[t] + (1.f - b2) * gg * gg;
        w[t] -= lr * (m[t] / bc1) / (sqrtf(v[t] / bc2) + eps);
      } else if (opt == SOPT_ADAGRAD) {
        m[t] += gg * gg;
        w[t] -= lr * gg / (sqrtf(m[t]) + eps);
      } else {
        w[t] -= lr * gg;
      }
    }
    if (VEC) {
      *reinterpret_cast<float4*>(wrow + c) = make_float4(w[0], w[1], w[2], w[3]);
      if (st.m) *reinterpret_cast<float4*>(st.m + local * (long long)D + c) = make_float4(m[0], m[1], m[2], m[3]);
      if (st.v) *reinterpret_cast<float4*>(st.v + local * (long long)D + c) = make_float4(v[0], v[1], v[2], v[3]);
      if (st.w16) {
        uint2 pk; pk.x = pack_bf16(w[0], w[1]); pk.y = pack_bf16(w[2], w[3]);
        *reinterpret_cast<uint2*>(st.w16 + local * (long long)D + c) = pk;
      }
    } else {
      wrow[c] = w[0];
      if (st.m) st.m[local * (long long)D + c] = m[0];
      if (st.v) st.v[local * (long long)D + c] = v[0];
      if (st.w16) st.w16[local * (long long)D + c] = __float2bfloat16(w[0]);
    }
  }
  __syncwarp();
  if (lane == 0) {
    rep_of_row[local] = INT_MAX;
    if (dup) dup_flag[e] = 0;
  }
}

// Pass 3, fast path (winner list, D % 4 == 0, D <= 128): the update is latency-bound -- a row is a handful of 16-byte
// loads (gradient row, possibly from a peer over NVLink, plus w / m / v) followed by a little math -- so the kernel is
// organised for loads in flight rather than for arithmetic: a row uses only as many lanes as it has float4 columns
// (G = 8 / 16 / 32), a warp therefore covers 32/G rows per pass, and every warp issues the loads of R passes before it
// computes anything.
// LOOP == false: one pass per warp over a grid sized for the entries (small entry counts: no register cost of the loop, full
// occupancy); LOOP == true: a grid of a few waves strides over the winner list (entry counts far above the winner count).
template <int R, bool LOOP>
__global__ void __launch_bounds__(256, R == 1 ? (LOOP ? 5 : 6) : 1)
emb_apply_fast_kernel(BwdEntries en, int D, int G, float* __restrict__ accum, int* __restrict__ dup_flag, int* __restrict__ rep_of_row,
                      SparseState st, int opt, const float* __restrict__ hyper, const int2* __restrict__ win_list,
                      const int* __restrict__ win_count) {
  pdl_wait();
  pdl_launch_dependents();
  const int count = *win_count;
  const int lane = threadIdx.x & 31;
  const int P = 32 / G;                                   // rows per warp pass
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int slot = lane / G, c = (lane - slot * G) * 4;
  const bool col_ok = c < D;
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], step = hyper[5];
  float bc1 = 1.f, bc2 = 1.f;
  if (opt == SOPT_ADAM) { bc1 = 1.f - powf(b1, step); bc2 = 1.f - powf(b2, step); }
  // The grid is sized for full occupancy, not for the entry count: the winner count is only known on the device, and a grid
  // of one warp per POSSIBLE entry (world x batch of them for a sharded table) spent most of its time launching empty CTAs.
  int base = warp * P * R;
  if (base >= count) return;
  do {                                                     // LOOP == false: exactly one pass (the grid covers every possible entry)
  int e[R]; long long local[R]; bool ok[R], dup[R];
  float g[R][4]; float4 w4[R], m4[R], v4[R], a4[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {                            // ---- issue every load first
    const int idx = base + r * P + slot;
    ok[r] = idx < count && col_ok;
    e[r] = 0; local[r] = 0; dup[r] = false;
    if (idx < count) { const int2 wl = win_list[idx]; e[r] = wl.x; local[r] = wl.y; }
    if (ok[r]) {
      dup[r] = dup_flag[e[r]] != 0;
      entry_grad<true>(en, e[r], c, g[r]);
      const long long o = local[r] * (long long)D + c;
      w4[r] = *reinterpret_cast<const float4*>(st.w + o);
      if (st.m) m4[r] = *reinterpret_cast<const float4*>(st.m + o);
      if (st.v) v4[r] = *reinterpret_cast<const float4*>(st.v + o);
      if (dup[r]) a4[r] = *reinterpret_cast<const float4*>(accum + (long long)e[r] * D + c);
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {                            // ---- then compute and store
    if (!ok[r]) continue;
    float w[4] = {w4[r].x, w4[r].y, w4[r].z, w4[r].w};
    float m[4] = {0, 0, 0, 0}, v[4] = {0, 0, 0, 0};
    if (st.m) { m[0] = m4[r].x; m[1] = m4[r].y; m[2] = m4[r].z; m[3] = m4[r].w; }
    if (st.v) { v[0] = v4[r].x; v[1] = v4[r].y; v[2] = v4[r].z; v[3] = v4[r].w; }
    if (dup[r]) {
      g[r][0] += a4[r].x; g[r][1] += a4[r].y; g[r][2] += a4[r].z; g[r][3] += a4[r].w;
      *reinterpret_cast<float4*>(accum + (long long)e[r] * D + c) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float gg = g[r][t] + wd * w[t];
      if (opt == SOPT_ADAM) {
        m[t] = b1 * m[t] + (1.f - b1) * gg;
        v[t] = b2 * v[t] + (1.f - b2) * gg * gg;
        w[t] -= lr * (m[t] / bc1) / (sqrtf(v[t] / bc2) + eps);
      } else if (opt == SOPT_ADAGRAD) {
        m[t] += gg * gg;
        w[t] -= lr * gg / (sqrtf(m[t]) + eps);
      } else {
        w[t] -= lr * gg;
      }
    }
    const long long o = local[r] * (long long)D + c;
    *reinterpret_cast<float4*>(st.w + o) = make_float4(w[0], w[1], w[2], w[3]);
    if (st.m) *reinterpret_cast<float4*>(st.m + o) = make_float4(m[0], m[1], m[2], m[3]);
    if (st.v) *reinterpret_cast<float4*>(st.v + o) = make_float4(v[0], v[1], v[2], v[3]);
    if (st.w16) {
      uint2 pk; pk.x = pack_bf16(w[0], w[1]); pk.y = pack_bf16(w[2], w[3]);
      *reinterpret_cast<uint2*>(st.w16 + o) = pk;
    }
    if (c == 0) {                                          // one lane per row restores the scratch
      rep_of_row[local[r]] = INT_MAX;
      if (dup[r]) dup_flag[e[r]] = 0;
    }
  }
  base += nwarps * P * R;
  } while (LOOP && base < count);
}

// Dense-gradient mode (exact reference semantics for small tables): dTable[id] += dOut rows.
template <bool VEC>
__global__ void __launch_bounds__(256)
emb_dense_grad_kernel(BwdEntries en, int E, long long rows, int D, float* __restrict__ dtable) {
  const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (e >= E) return;
  const long long id = entry_row(en, e);
  if (id < 0 || id >= rows) return;
  float* drow = dtable + id * (long long)D;
  if (VEC) {
    for (int c = lane * 4; c < D; c += 128) {
      float g[4];
      entry_grad<true>(en, e, c, g);
      red_add_f32x4(drow + c, g[0], g[1], g[2], g[3]);
    }
  } else {
    for (int c = lane; c < D; c += 32) {
      float g[4];
      entry_grad<false>(en, e, c, g);
      atomicAdd(drow + c, g[0]);
    }
  }
}

__global__ void fill_i32_kernel(int* p, long long n, int v) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = v;
}

static BwdEntries single_source(const long long* ids, int id_stride, int id_col, int bag, const void* dout, long long ldg,
                                const void* mul, long long ldmul, float grad_scale, int E) {
  BwdEntries en{};
  en.ids[0] = ids; en.id_stride = id_stride; en.id_col = id_col; en.bag = bag;
  en.dout[0] = (const __nv_bfloat16*)dout; en.ldg = ldg;
  en.mul[0] = (const __nv_bfloat16*)mul; en.ldmul = ldmul;
  en.grad_scale = grad_scale; en.nsrc = 1; en.E_local = E;
  return en;
}

static ShardPtrs make_ptrs(const void* const* tabs, int world) {
  ShardPtrs s{};
  for (int i = 0; i < kMaxPeers; ++i) s.p[i] = (tabs && i < world) ? tabs[i] : nullptr;
  return s;
}

// The three passes.  win_list (optional, int32 [2*E + 4]): [0..2E) = (entry, local row) pairs, [2E] = count.
static int launch_sparse_update(const BwdEntries& en, int E, long long rows, int world, int rank, int D, int* rep_of_row,
                                float* accum, int* dup_flag, const SparseState& st, int opt, const float* hyper, bool vec,
                                int* win_list, cudaStream_t s, bool bump_step = false) {
  const int threads = 256;
  int2* wl = reinterpret_cast<int2*>(win_list);
  int* wc = win_list ? win_list + 2 * (size_t)E : nullptr;
  ZOO_CUDA_CHECK(zoo_launch_pdl(emb_mark_kernel, dim3((E + threads - 1) / threads), dim3(threads), 0, s, en, E, rows, world, rank, rep_of_row, wc,
                                bump_step ? const_cast<float*>(hyper) + 5 : (float*)nullptr));
  const unsigned wb = (unsigned)(((long long)E * 32 + threads - 1) / threads);
  const unsigned ab = (unsigned)((E + threads - 1) / threads);  // accum: one lane per entry
  if (vec && win_list && D <= 128) {
    ZOO_CUDA_CHECK(zoo_launch_pdl(emb_accum_kernel<true>, dim3(ab), dim3(threads), 0, s, en, E, rows, world, rank, (const int*)rep_of_row, D, accum, dup_flag, wl, wc));
    const int G = D <= 32 ? 8 : (D <= 64 ? 16 : 32);
    static int R = -1;
    if (R < 0) { const char* ev = getenv("ZOO_APPLY_R"); R = ev ? atoi(ev) : 1; if (R != 1 && R != 2 && R != 4) R = 1; }
    const int rows_per_warp = (32 / G) * R;
    const long long warps = ((long long)E + rows_per_warp - 1) / rows_per_warp;
    long long grid_ll = (warps * 32 + threads - 1) / threads;
    static int sm_count = 0;
    if (!sm_count) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev); if (sm_count <= 0) sm_count = 148; }
    // up to 64 CTAs per SM worth of entries (one GPU's own batch): one pass, no loop; beyond that (world x batch possible entries
    // of a sharded table, most of them not owned here) a few waves stride over the winner list
    const bool loop = grid_ll > (long long)sm_count * 64;
    if (loop) grid_ll = (long long)sm_count * 16;
    const unsigned grid = (unsigned)grid_ll;
    if (R == 1 && !loop) ZOO_CUDA_CHECK(zoo_launch_pdl(emb_apply_fast_kernel<1, false>, dim3(grid), dim3(threads), 0, s, en, D, G, accum, dup_flag, rep_of_row, st, opt, hyper, (const int2*)wl, (const int*)wc));
    else if (R == 1) ZOO_CUDA_CHECK(zoo_launch_pdl(emb_apply_fast_kernel<1, true>, dim3(grid), dim3(threads), 0, s, en, D, G, accum, dup_flag, rep_of_row, st, opt, hyper, (const int2*)wl, (const int*)wc));
    else if (R == 2) emb_apply_fast_kernel<2, true><<<grid, threads, 0, s>>>(en, D, G, accum, dup_flag, rep_of_row, st, opt, hyper, wl, wc);
    else emb_apply_fast_kernel<4, true><<<grid, threads, 0, s>>>(en, D, G, accum, dup_flag, rep_of_row, st, opt, hyper, wl, wc);
  } else if (vec) {
    emb_accum_kernel<true><<<ab, threads, 0, s>>>(en, E, rows, world, rank, rep_of_row, D, accum, dup_flag, wl, wc);
    emb_apply_kernel<true><<<wb, threads, 0, s>>>(en, E, rows, world, rank, rep_of_row, D, accum, dup_flag, st, opt, hyper, wl, wc);
  } else {
    emb_accum_kernel<false><<<ab, threads, 0, s>>>(en, E, rows, world, rank, rep_of_row, D, accum, dup_flag, wl, wc);
    emb_apply_kernel<false><<<wb, threads, 0, s>>>(en, E, rows, world, rank, rep_of_row, D, accum, dup_flag, st, opt, hyper, wl, wc);
  }
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

}  // namespace zoo

using namespace zoo;

// out[b, 0:D] (bf16, leading dim ldo) = pool_j table[ids[b, id_col + j]]  (x table2[ids[b, id_col2 + j]] if tables2)
//   tables / tables2 : host arrays of `world` device pointers (one shard per GPU; peers via P2P)
//   table_dtype      : 0 = fp32 rows, 1 = bf16 rows
//   rows_a / rows_b  : global (unsharded) row counts
//   save_a / save_b  : optional [B, D] bf16 buffers receiving the two factors of a product (bag must be 1)
ZOO_API int zoo_embedding_gather(const long long* ids, int id_stride, int id_col, int id_col2, int B, int bag,
                                 int mean, const void* const* tables, long long rows_a,
                                 const void* const* tables2, long long rows_b, int world, int D,
                                 int table_dtype, void* out, long long ldo, void* save_a, void* save_b,
                                 void* stream) {
  if (!ids || !tables || !out || B < 0 || D <= 0 || world < 1 || world > kMaxPeers || bag < 1 ||
      ((save_a != nullptr) != (save_b != nullptr)) || (save_a && (!tables2 || bag != 1))) {
    zoo_set_last_error("embedding_gather: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (B == 0) return ZOO_OK;
  const ShardPtrs ta = make_ptrs(tables, world);
  const ShardPtrs tb = make_ptrs(tables2, world);
  const bool mul = tables2 != nullptr;
  const int vn = table_dtype == 0 ? 4 : 8;
  bool vec = (D % vn == 0) && (ldo % vn == 0) && (((uintptr_t)out & (vn * 2 - 1)) == 0);
  for (int i = 0; i < world && vec; ++i) {
    if ((uintptr_t)ta.p[i] & 15) vec = false;
    if (mul && ((uintptr_t)tb.p[i] & 15)) vec = false;
  }
  const int threads = 256;
  // lanes per sample: enough 16-byte vectors (or scalars) to cover one row, rounded up to a power of two
  const int units = vec ? (D + vn - 1) / vn : D;
  int G = 32;
  while (G > 1 && G / 2 >= units) G /= 2;
  const long long warps = ((long long)B + (32 / G) - 1) / (32 / G);
  const long long blocks = (warps * 32 + threads - 1) / threads;
  cudaStream_t s = (cudaStream_t)stream;
  __nv_bfloat16* o = (__nv_bfloat16*)out;
  __nv_bfloat16* sa = (__nv_bfloat16*)save_a;
  __nv_bfloat16* sb = (__nv_bfloat16*)save_b;
#define ZOO_GATHER(T, MUL, VEC)                                                                     \
  ZOO_CUDA_CHECK(zoo_launch_pdl(embedding_gather_kernel<T, MUL, VEC>, dim3((unsigned)blocks), dim3(threads), 0, s, \
      ids, id_stride, id_col, id_col2, B, bag, mean, ta, rows_a, tb, rows_b, world, D, o, ldo, sa, sb, G))
  if (table_dtype == 0) {
    if (mul) { if (vec) ZOO_GATHER(float, true, true); else ZOO_GATHER(float, true, false); }
    else     { if (vec) ZOO_GATHER(float, false, true); else ZOO_GATHER(float, false, false); }
  } else {
    if (mul) { if (vec) ZOO_GATHER(__nv_bfloat16, true, true); else ZOO_GATHER(__nv_bfloat16, true, false); }
    else     { if (vec) ZOO_GATHER(__nv_bfloat16, false, true); else ZOO_GATHER(__nv_bfloat16, false, false); }
  }
#undef ZOO_GATHER
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

// Scratch owned by the caller (see zoo/ops/embedding.py):
//   rep_of_row : int32 [rows_local]  -- must be INT_MAX everywhere on entry; restored on exit
//   accum      : fp32  [E, D]        -- must be zero on entry; restored on exit
//   dup_flag   : int32 [E]           -- must be zero on entry; restored on exit
ZOO_API int zoo_embedding_scratch_init(int* rep_of_row, long long rows_local, void* stream) {
  if (!rep_of_row || rows_local <= 0) return ZOO_ERR_ARG;
  fill_i32_kernel<<<1024, 256, 0, (cudaStream_t)stream>>>(rep_of_row, rows_local, INT_MAX);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

// Fused sparse optimizer step for the table shard owned by `rank`.
//   ids/dout describe ALL entries visible to this GPU (world == 1: its own batch; world > 1: the
//   mailbox the peers filled); entries owned by other ranks are skipped.
//   mul (optional, bf16 [B, ldmul]): gradient rows are multiplied elementwise by mul rows
//   (the other factor of a GMF product, saved by the forward).
//   opt: 0 = SGD, 1 = Adam (lazy / touched rows only), 2 = Adagrad.   hyper: device fp32[6] =
//   {lr, beta1, beta2, eps, weight_decay, step}.
ZOO_API int zoo_embedding_sparse_update(const long long* ids, int id_stride, int id_col, int B, int bag,
                                        const void* dout, long long ldg, const void* mul, long long ldmul,
                                        float grad_scale, long long rows, int world, int rank, int D,
                                        float* w, void* w16, float* m, float* v, int opt, const float* hyper,
                                        int* rep_of_row, float* accum, int* dup_flag, void* stream) {
  if (!ids || !dout || !w || !hyper || !rep_of_row || !accum || !dup_flag || D <= 0 || bag < 1 || world < 1 ||
      world > kMaxPeers || rank < 0 || rank >= world) {
    zoo_set_last_error("embedding_sparse_update: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if ((opt == SOPT_ADAM && (!m || !v)) || (opt == SOPT_ADAGRAD && !m)) {
    zoo_set_last_error("embedding_sparse_update: optimizer state missing", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const int E = B * bag;
  if (E == 0) return ZOO_OK;
  BwdEntries en = single_source(ids, id_stride, id_col, bag, dout, ldg, mul, ldmul, grad_scale, E);
  SparseState st{w, (__nv_bfloat16*)w16, opt == SOPT_SGD ? nullptr : m, opt == SOPT_ADAM ? v : nullptr};
  const bool vec = (D % 4 == 0) && (ldg % 4 == 0) && (((uintptr_t)dout & 7) == 0) && (((uintptr_t)w & 15) == 0) &&
                   (((uintptr_t)accum & 15) == 0) && (!m || ((uintptr_t)m & 15) == 0) &&
                   (!v || ((uintptr_t)v & 15) == 0) && (!w16 || ((uintptr_t)w16 & 7) == 0) &&
                   (!mul || ((ldmul % 4 == 0) && ((uintptr_t)mul & 7) == 0));
  cudaStream_t s = (cudaStream_t)stream;
  return launch_sparse_update(en, E, rows, world, rank, D, rep_of_row, accum, dup_flag, st, opt, hyper, vec, nullptr, s);
}

// zoo_embedding_sparse_update with one more caller-owned scratch buffer, `win_list` (int32 [2*E + 4], any
// contents): pass 2 compacts the representative entries into it and pass 3 runs only over that list.
ZOO_API int zoo_embedding_sparse_update_w(const long long* ids, int id_stride, int id_col, int B, int bag,
                                          const void* dout, long long ldg, const void* mul, long long ldmul,
                                          float grad_scale, long long rows, int world, int rank, int D,
                                          float* w, void* w16, float* m, float* v, int opt, const float* hyper,
                                          int* rep_of_row, float* accum, int* dup_flag, int* win_list, int bump_step, void* stream) {
  if (!ids || !dout || !w || !hyper || !rep_of_row || !accum || !dup_flag || !win_list || D <= 0 || bag < 1 || world < 1 ||
      world > kMaxPeers || rank < 0 || rank >= world || (rows + world - 1) / world > INT_MAX) {
    zoo_set_last_error("embedding_sparse_update_w: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if ((opt == SOPT_ADAM && (!m || !v)) || (opt == SOPT_ADAGRAD && !m)) {
    zoo_set_last_error("embedding_sparse_update_w: optimizer state missing", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const int E = B * bag;
  if (E == 0) return ZOO_OK;
  BwdEntries en = single_source(ids, id_stride, id_col, bag, dout, ldg, mul, ldmul, grad_scale, E);
  SparseState st{w, (__nv_bfloat16*)w16, opt == SOPT_SGD ? nullptr : m, opt == SOPT_ADAM ? v : nullptr};
  const bool vec = (D % 4 == 0) && (ldg % 4 == 0) && (((uintptr_t)dout & 7) == 0) && (((uintptr_t)w & 15) == 0) &&
                   (((uintptr_t)accum & 15) == 0) && (!m || ((uintptr_t)m & 15) == 0) &&
                   (!v || ((uintptr_t)v & 15) == 0) && (!w16 || ((uintptr_t)w16 & 7) == 0) &&
                   (!mul || ((ldmul % 4 == 0) && ((uintptr_t)mul & 7) == 0));
  return launch_sparse_update(en, E, rows, world, rank, D, rep_of_row, accum, dup_flag, st, opt, hyper, vec, win_list, (cudaStream_t)stream, bump_step != 0);
}

// Multi-GPU form of zoo_embedding_sparse_update: the backward of a row-sharded table.
//   ids_src / dout_src / mul_src : HOST arrays of `world` device pointers -- every rank's id matrix,
//   gradient-row window and (optional) multiplier rows, all living in the symmetric heap so this GPU
//   can read them in place.  B, id_stride, ldg, ldmul are the same on every rank.
// The owner PULLS what it needs: all world*B*bag entries are scanned (ids are tiny), and only the
// gradient rows of entries whose row lives here cross NVLink -- the all-to-all of a classic
// dispatch/combine design without a separate collective or a staging copy on the receiver.
// Callers bracket this with zoo_p2p_barrier (peers' buffers complete before / untouched until after).
//   accum : fp32 [world*B*bag, D], dup_flag : int32 [world*B*bag] (zero on entry, restored on exit).
//   bump_step : 1 = hyper[5] (the optimizer step) is incremented on the device before it is used.
ZOO_API int zoo_embedding_sparse_update_multi(const long long* const* ids_src, int id_stride, int id_col, int B, int bag,
                                              const void* const* dout_src, long long ldg, const void* const* mul_src,
                                              long long ldmul, float grad_scale, long long rows, int world, int rank,
                                              int D, float* w, void* w16, float* m, float* v, int opt, const float* hyper,
                                              int* rep_of_row, float* accum, int* dup_flag, int* win_list, int bump_step, void* stream) {
  if ((rows + world - 1) / (world > 0 ? world : 1) > INT_MAX) { zoo_set_last_error("embedding_sparse_update_multi: shard too large", __FILE__, __LINE__); return ZOO_ERR_TOO_LARGE; }
  if (!ids_src || !dout_src || !w || !hyper || !rep_of_row || !accum || !dup_flag || D <= 0 || bag < 1 || world < 1 ||
      world > kMaxPeers || rank < 0 || rank >= world) {
    zoo_set_last_error("embedding_sparse_update_multi: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if ((opt == SOPT_ADAM && (!m || !v)) || (opt == SOPT_ADAGRAD && !m)) {
    zoo_set_last_error("embedding_sparse_update_multi: optimizer state missing", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const long long E64 = (long long)world * B * bag;
  if (E64 == 0) return ZOO_OK;
  if (E64 > INT_MAX / 32) { zoo_set_last_error("embedding_sparse_update_multi: too many entries", __FILE__, __LINE__); return ZOO_ERR_TOO_LARGE; }
  const int E = (int)E64;
  BwdEntries en{};
  bool vec = (D % 4 == 0) && (ldg % 4 == 0) && (((uintptr_t)w & 15) == 0) && (((uintptr_t)accum & 15) == 0) &&
             (!m || ((uintptr_t)m & 15) == 0) && (!v || ((uintptr_t)v & 15) == 0) && (!w16 || ((uintptr_t)w16 & 7) == 0);
  for (int s = 0; s < world; ++s) {
    if (!ids_src[s] || !dout_src[s]) { zoo_set_last_error("embedding_sparse_update_multi: null source pointer", __FILE__, __LINE__); return ZOO_ERR_ARG; }
    en.ids[s] = ids_src[s];
    en.dout[s] = (const __nv_bfloat16*)dout_src[s];
    en.mul[s] = mul_src ? (const __nv_bfloat16*)mul_src[s] : nullptr;
    vec = vec && (((uintptr_t)dout_src[s] & 7) == 0) && (!en.mul[s] || ((ldmul % 4 == 0) && ((uintptr_t)en.mul[s] & 7) == 0));
  }
  en.id_stride = id_stride; en.id_col = id_col; en.bag = bag; en.ldg = ldg; en.ldmul = ldmul;
  en.grad_scale = grad_scale; en.nsrc = world; en.E_local = B * bag;
  SparseState st{w, (__nv_bfloat16*)w16, opt == SOPT_SGD ? nullptr : m, opt == SOPT_ADAM ? v : nullptr};
  cudaStream_t s = (cudaStream_t)stream;
  return launch_sparse_update(en, E, rows, world, rank, D, rep_of_row, accum, dup_flag, st, opt, hyper, vec, win_list, s, bump_step != 0);
}

// K4b: backward DISPATCH of a row-sharded table.  Every source GPU pushes the gradient row of each of its entries into the
// receive buffer of the GPU that owns the row (row r lives on rank r % world): posted NVLink writes of whole 16-byte
// vectors, issued by the side that already has the data, instead of latency-bound remote reads by the owner.  Slot
// `rank * cap + pos` of the owner's buffers receives (local row index, gradient row); `pos` comes from a per-destination
// counter on the source (one global atomic per destination per CTA).  After a barrier the owner runs the ordinary
// single-GPU sparse update over its receive buffer, whose unused slots hold the id -1.
//   recv_ids  : HOST array of `world` device pointers, int64 [world * cap]  (peer memory, pre-filled with -1)
//   recv_grad : HOST array of `world` device pointers, bf16  [world * cap, D]
//   counters  : int32 [world] on this GPU (zeroed here)
// The multiplier rows (GMF factor) and grad_scale are applied at the source, so only the final gradient row travels.
namespace zoo {
struct DispatchArgs {
  long long* recv_ids[kMaxPeers];
  __nv_bfloat16* recv_grad[kMaxPeers];
};

template <bool VEC>
__global__ void __launch_bounds__(256)
emb_dispatch_kernel(const long long* __restrict__ ids, int id_stride, int id_col, int E, int bag,
                    const __nv_bfloat16* __restrict__ dout, long long ldg, const __nv_bfloat16* __restrict__ mul, long long ldmul,
                    float grad_scale, long long rows, int world, int rank, int D, int G, DispatchArgs a,
                    int* __restrict__ counters, int cap) {
  __shared__ int s_cnt[kMaxPeers];     // entries of this CTA per destination, then the CTA's base position there
  __shared__ int s_dest[256];          // destination rank of entry t of this CTA (-1: skipped)
  __shared__ int s_slot[256];          // slot in the destination's receive buffers
  const int t = threadIdx.x;
  if (t < kMaxPeers) s_cnt[t] = 0;
  __syncthreads();
  // ---- phase 1: one thread per entry picks destination and position
  const int e = blockIdx.x * 256 + t;
  int o = -1, mypos = 0;
  long long local = 0;
  if (e < E) {
    const int b = e / bag, j = e - b * bag;
    const long long id = ids[(long long)b * id_stride + id_col + j];
    if (id >= 0 && id < rows) {
      o = (int)(id % world);
      local = id / world;
      mypos = atomicAdd(&s_cnt[o], 1);
    }
  }
  __syncthreads();
  if (t < world) {
    const int c = s_cnt[t];
    s_cnt[t] = c ? atomicAdd(counters + t, c) : 0;
  }
  __syncthreads();
  int slot = 0;
  if (o >= 0) {
    const int pos = s_cnt[o] + mypos;
    if (pos >= cap) { printf("zoo emb_dispatch: receive buffer overflow (rank %d -> %d)\n", rank, o); __trap(); }
    slot = rank * cap + pos;
    pick8(a.recv_ids, o)[slot] = local;
  }
  s_dest[t] = o;
  s_slot[t] = slot;
  __syncthreads();
  // ---- phase 2: G lanes move one gradient row (16-byte vectors), 32 / G rows per warp pass
  const int lane = t & 31, warp = t >> 5, P = 32 / G, sub = lane / G, l = lane - sub * G;
  for (int base = warp * P; base < 256; base += 8 * P) {
    const int k = base + sub;
    const int dst = s_dest[k];
    if (dst < 0) continue;
    const int ek = blockIdx.x * 256 + k;
    const int b = ek / bag;
    const __nv_bfloat16* src = dout + (long long)b * ldg;
    const __nv_bfloat16* mrow = mul ? mul + (long long)b * ldmul : nullptr;
    __nv_bfloat16* out = pick8(a.recv_grad, dst) + (long long)s_slot[k] * D;
    if (VEC) {
      for (int c = l * 8; c < D; c += G * 8) {
        uint4 v = *reinterpret_cast<const uint4*>(src + c);
        if (mrow || grad_scale != 1.f) {
          uint32_t w[4] = {v.x, v.y, v.z, v.w};
          uint4 m4 = make_uint4(0, 0, 0, 0);
          if (mrow) m4 = *reinterpret_cast<const uint4*>(mrow + c);
          const uint32_t mw[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float2 f = unpack_bf16(w[q]);
            if (mrow) { const float2 g = unpack_bf16(mw[q]); f.x *= g.x; f.y *= g.y; }
            w[q] = pack_bf16(f.x * grad_scale, f.y * grad_scale);
          }
          v = make_uint4(w[0], w[1], w[2], w[3]);
        }
        *reinterpret_cast<uint4*>(out + c) = v;
      }
    } else {
      for (int c = l; c < D; c += G) {
        float f = __bfloat162float(src[c]);
        if (mrow) f *= __bfloat162float(mrow[c]);
        out[c] = __float2bfloat16(f * grad_scale);
      }
    }
  }
  __threadfence_system();     // pushed rows are ordered before the flag stores of the barrier kernel that follows
}

}  // namespace zoo
using namespace zoo;

ZOO_API int zoo_embedding_dispatch(const long long* ids, int id_stride, int id_col, int B, int bag, const void* dout, long long ldg,
                                   const void* mul, long long ldmul, float grad_scale, long long rows, int world, int rank, int D,
                                   void* const* recv_ids, void* const* recv_grad, int* counters, int cap, void* stream) {
  if (!ids || !dout || !recv_ids || !recv_grad || !counters || B < 0 || bag < 1 || D <= 0 || world < 1 || world > kMaxPeers ||
      rank < 0 || rank >= world || cap < 1) {
    zoo_set_last_error("embedding_dispatch: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const long long E64 = (long long)B * bag;
  if (E64 > INT_MAX) { zoo_set_last_error("embedding_dispatch: too many entries", __FILE__, __LINE__); return ZOO_ERR_TOO_LARGE; }
  const int E = (int)E64;
  cudaStream_t s = (cudaStream_t)stream;
  ZOO_CUDA_CHECK(cudaMemsetAsync(counters, 0, sizeof(int) * world, s));
  if (E == 0) return ZOO_OK;
  DispatchArgs a{};
  bool vec = (D % 8 == 0) && (ldg % 8 == 0) && (((uintptr_t)dout & 15) == 0) && (!mul || ((ldmul % 8 == 0) && ((uintptr_t)mul & 15) == 0));
  for (int p = 0; p < world; ++p) {
    if (!recv_ids[p] || !recv_grad[p]) { zoo_set_last_error("embedding_dispatch: null receive buffer", __FILE__, __LINE__); return ZOO_ERR_ARG; }
    a.recv_ids[p] = (long long*)recv_ids[p];
    a.recv_grad[p] = (__nv_bfloat16*)recv_grad[p];
    vec = vec && (((uintptr_t)recv_grad[p] & 15) == 0);
  }
  const int units = vec ? (D + 7) / 8 : D;
  int G = 32;
  while (G > 1 && G / 2 >= units) G /= 2;
  const unsigned blocks = (unsigned)((E + 255) / 256);
  if (vec) emb_dispatch_kernel<true><<<blocks, 256, 0, s>>>(ids, id_stride, id_col, E, bag, (const __nv_bfloat16*)dout, ldg, (const __nv_bfloat16*)mul,
                                                           ldmul, grad_scale, rows, world, rank, D, G, a, counters, cap);
  else     emb_dispatch_kernel<false><<<blocks, 256, 0, s>>>(ids, id_stride, id_col, E, bag, (const __nv_bfloat16*)dout, ldg, (const __nv_bfloat16*)mul,
                                                            ldmul, grad_scale, rows, world, rank, D, G, a, counters, cap);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

// Dense gradient: dtable[ids] += dout rows (fp32 atomics); used when the table is trained as a
// dense parameter exactly like the reference does.
ZOO_API int zoo_embedding_dense_grad(const long long* ids, int id_stride, int id_col, int B, int bag,
                                     const void* dout, long long ldg, const void* mul, long long ldmul,
                                     float grad_scale, long long rows, int D, float* dtable, void* stream) {
  if (!ids || !dout || !dtable || D <= 0 || bag < 1) {
    zoo_set_last_error("embedding_dense_grad: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const int E = B * bag;
  if (E == 0) return ZOO_OK;
  BwdEntries en = single_source(ids, id_stride, id_col, bag, dout, ldg, mul, ldmul, grad_scale, E);
  const bool vec = (D % 4 == 0) && (ldg % 4 == 0) && (((uintptr_t)dout & 7) == 0) && (((uintptr_t)dtable & 15) == 0) &&
                   (!mul || ((ldmul % 4 == 0) && ((uintptr_t)mul & 7) == 0));
  const int threads = 256;
  const unsigned wb = (unsigned)(((long long)E * 32 + threads - 1) / threads);
  if (vec) emb_dense_grad_kernel<true><<<wb, threads, 0, (cudaStream_t)stream>>>(en, E, rows, D, dtable);
  else     emb_dense_grad_kernel<false><<<wb, threads, 0, (cudaStream_t)stream>>>(en, E, rows, D, dtable);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

// ------------------------------------------------------------------------------------------
// Allocation-free variant of zoo_embedding_sparse_update.
//
// Who needs it: code that must apply an embedding update but is not allowed to obtain device
// memory at that point -- a latency-critical serving/online-learning thread (cudaMalloc and
// cudaFree synchronise the whole device and take the driver's allocator lock), a host callback
// running inside another framework's stream where allocating is forbidden, or a process that
// handed its entire HBM budget to a pool at start-up so a late cudaMalloc would simply fail.
// Such callers still own the table, its optimizer state, the ids and the gradient rows; what
// they cannot provide is the three scratch buffers of the regular entry point.  Here that
// scratch lives in ONE statically allocated device array (below), so the limits are fixed:
//
//     local table rows  (rows / world, rounded up)  <= ZOO_NOALLOC_MAX_ROWS     (1,048,576)
//     entries per call  (B * bag)                   <= ZOO_NOALLOC_MAX_ENTRIES  (8,192)
//     embedding width   D                           <= ZOO_NOALLOC_MAX_DIM      (128)
//
// Larger problems are refused with ZOO_ERR_TOO_LARGE (use the regular entry point and pass your
// own scratch).  The array costs about 8.03 MiB of device memory on EVERY GPU whose context
// loads this library, from load until the process exits, whether or not the entry point is used.
//
// Sharing rules.  All calls on a device use the same array and it keeps whatever the previous
// call left behind, so every call first re-initialises the part it is going to use (row map to
// INT_MAX, accumulation rows and duplicate flags to zero) on the caller's stream.  Two calls
// whose kernels can overlap on the GPU (different streams without an ordering between them, or
// several host threads) would corrupt each other's scratch: serialise them.
//
// First call on a device.  The device address of the array is looked up once per device (with
// that device current) on the first call made there; do that first call outside of any CUDA
// stream capture.
#define ZOO_NOALLOC_MAX_ROWS (1 << 20)
#define ZOO_NOALLOC_MAX_ENTRIES 8192
#define ZOO_NOALLOC_MAX_DIM 128

namespace zoo {
constexpr size_t kNaRepBytes = sizeof(int) * (size_t)ZOO_NOALLOC_MAX_ROWS;                              // row -> entry map
constexpr size_t kNaAccumBytes = sizeof(float) * (size_t)ZOO_NOALLOC_MAX_ENTRIES * ZOO_NOALLOC_MAX_DIM;  // duplicate sums
constexpr size_t kNaDupBytes = sizeof(int) * (size_t)ZOO_NOALLOC_MAX_ENTRIES;                            // duplicate flags
constexpr size_t kNaTotalBytes = kNaRepBytes + kNaAccumBytes + kNaDupBytes;
}  // namespace zoo

// File-scope device storage.  In HOST code this name is only a handle for the runtime's symbol
// table -- it is not a device address and must never be handed to a kernel; the address comes
// from cudaGetSymbolAddress (see noalloc_scratch_base below).
static __device__ __align__(256) unsigned char g_zoo_noalloc_scratch[zoo::kNaTotalBytes];

namespace zoo {
constexpr int kNaMaxDevices = 64;
// Cache of the array's device address per device ordinal.  Several host threads may fill an entry at the same time without a
// lock: every writer stores the SAME value (cudaGetSymbolAddress is a pure lookup for a given device), a pointer-sized store
// is not torn on the supported hosts, and a reader that still sees null simply repeats the lookup.
static void* g_na_base[kNaMaxDevices] = {nullptr};

// Per-device set-up: resolve the device address of the static array for the CURRENT device.
static int noalloc_scratch_base(unsigned char** base) {
  int dev = -1;
  ZOO_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNaMaxDevices) {
    zoo_set_last_error("embedding_sparse_update_noalloc: device ordinal out of range", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (!g_na_base[dev]) {
    void* p = nullptr;
    ZOO_CUDA_CHECK(cudaGetSymbolAddress(&p, g_zoo_noalloc_scratch));
    if (!p) {
      zoo_set_last_error("embedding_sparse_update_noalloc: cudaGetSymbolAddress returned null", __FILE__, __LINE__);
      return ZOO_ERR_CUDA;
    }
    g_na_base[dev] = p;
  }
  *base = static_cast<unsigned char*>(g_na_base[dev]);
  return ZOO_OK;
}
}  // namespace zoo

// Same arguments and results as zoo_embedding_sparse_update, minus the three scratch pointers.
// Calls no cudaMalloc / cudaFree and no allocating library routine; see the block comment above
// for the limits, the sharing rules and the first-call note.
ZOO_API int zoo_embedding_sparse_update_noalloc(const long long* ids, int id_stride, int id_col, int B, int bag,
                                                const void* dout, long long ldg, const void* mul, long long ldmul,
                                                float grad_scale, long long rows, int world, int rank, int D,
                                                float* w, void* w16, float* m, float* v, int opt,
                                                const float* hyper, void* stream) {
  if (world < 1 || world > kMaxPeers || B < 0 || bag < 1 || D <= 0 || rows <= 0) {
    zoo_set_last_error("embedding_sparse_update_noalloc: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const long long rows_local = (rows + world - 1) / world;
  const long long E = (long long)B * bag;
  if (rows_local > ZOO_NOALLOC_MAX_ROWS || E > ZOO_NOALLOC_MAX_ENTRIES || D > ZOO_NOALLOC_MAX_DIM) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "embedding_sparse_update_noalloc: problem exceeds the static scratch (local rows %lld > %d, or entries "
             "%lld > %d, or dim %d > %d); use zoo_embedding_sparse_update with caller-owned scratch",
             rows_local, ZOO_NOALLOC_MAX_ROWS, E, ZOO_NOALLOC_MAX_ENTRIES, D, ZOO_NOALLOC_MAX_DIM);
    zoo_set_last_error(msg, __FILE__, __LINE__);
    return ZOO_ERR_TOO_LARGE;
  }
  if (E == 0) return ZOO_OK;
  unsigned char* base = nullptr;
  const int rc = noalloc_scratch_base(&base);
  if (rc != ZOO_OK) return rc;
  int* rep_of_row = reinterpret_cast<int*>(base);
  float* accum = reinterpret_cast<float*>(base + kNaRepBytes);
  int* dup_flag = reinterpret_cast<int*>(base + kNaRepBytes + kNaAccumBytes);
  cudaStream_t s = (cudaStream_t)stream;
  // The array still holds the previous call's state: bring the used part to its start values.
  fill_i32_kernel<<<256, 256, 0, s>>>(rep_of_row, rows_local, INT_MAX);
  ZOO_CUDA_CHECK(cudaGetLastError());
  ZOO_CUDA_CHECK(cudaMemsetAsync(accum, 0, sizeof(float) * (size_t)E * D, s));
  ZOO_CUDA_CHECK(cudaMemsetAsync(dup_flag, 0, sizeof(int) * (size_t)E, s));
  return zoo_embedding_sparse_update(ids, id_stride, id_col, B, bag, dout, ldg, mul, ldmul, grad_scale, rows, world,
                                     rank, D, w, w16, m, v, opt, hyper, rep_of_row, accum, dup_flag, stream);
}
