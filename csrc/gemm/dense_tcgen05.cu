// K3: Dense-layer GEMM family on 5th-gen tensor cores (tcgen05 + TMEM + TMA).
//
// One warp-specialised, persistent kernel computes
//     D[Mo, No] = sum_r A(Mo, r) * B(No, r)            (bf16 x bf16 -> fp32 in TMEM)
// with either operand K-major (row-major [rows, R]) or MN-major (row-major [R, rows]),
// so the three GEMMs of a Dense layer need no transposes:
//     forward  Y  = act(X W^T + b)        A = X  (K-major)   B = W  (K-major)
//     dgrad    dX = (dY W) * relu'(X)     A = dY (K-major)   B = W  (MN-major)
//     wgrad    dW += dY^T X               A = dY (MN-major)  B = X  (MN-major), split over batch
// Replaces the reference's Dense stacks over MKL sgemm
// (zoo/.../models/recommendation/NeuralCF.scala:73-96, WideAndDeep.scala:146-157) and the
// separate bias/activation passes of BigDL's Linear+ReLU modules.
//
// Roles per CTA (384 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane),
// warp 2 = TMEM allocator, warps 4-11 = epilogue (TMEM -> registers -> global; two warps per
// TMEM lane quarter so every SM sub-partition has two epilogue warps to interleave).  The epilogue
// stages each 32-row x 128-byte piece of the tile in swizzled shared memory and hands it to the TMA
// engine (bulk tensor store; reduce-add for the split-K fp32 partial sums), so global writes are full
// 128-byte lines issued asynchronously instead of 32 scattered 16-byte stores per instruction.
// Pipelines: 4-stage smem ring (TMA <-> MMA) and a 2-stage TMEM accumulator ring
// (MMA <-> epilogue) so the epilogue of tile i overlaps the main loop of tile i+1.
#include "../common/zoo_common.cuh"
#include <stdlib.h>

namespace zoo {

constexpr int kBlockM = 128;       // UMMA M
constexpr int kBlockK = 64;        // 64 bf16 = 128 B = one swizzle-128B row
constexpr int kUmmaK = 16;
constexpr int kMaxN = 256;         // UMMA N upper bound
constexpr int kStages = 4;          // ring depth when a stage carries A and B
constexpr int kMaxStages = 12;      // ring depth limit when B is resident and the ring carries only A
constexpr int kAccStages = 2;
constexpr int kABytes = kBlockM * kBlockK * 2;   // 16 KiB
constexpr int kBBytes = kMaxN * kBlockK * 2;     // 32 KiB
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kEpiSlotBytes = 4096;               // per epilogue warp: 32 rows x 128 B staging box for the TMA store
constexpr int kEpiStageOff = kStages * kStageBytes;
constexpr int kSmemBytes = kStages * kStageBytes + 8 * kEpiSlotBytes + 1024;  // + alignment slack
constexpr int kEpiWarps = 8;
constexpr int kThreads = 128 + kEpiWarps * 32;  // 4 control warps + 8 epilogue warps
constexpr int kMnBoxBytes = kBlockK * 128;       // one MN-major TMA box: 64 k-rows x 128 B

enum Epilogue : int {
  EPI_BIAS_ACT = 0,   // out(bf16) = act(acc + bias[n])
  EPI_MASK = 1,       // out(bf16) = acc * (mask[m,n] > 0)          (mask may be null)
  EPI_RED_F32 = 2     // out(fp32) += acc                           (split-K partial sums)
};
enum Act : int { ACT_NONE = 0, ACT_RELU = 1, ACT_GELU = 2, ACT_SIGMOID = 3, ACT_TANH = 4 };

struct GemmArgs {
  int Mo, No, R;            // output rows, output cols, reduction length
  int block_n;              // UMMA N for this launch (multiple of 16, <= 256)
  int a_mn, b_mn;           // 1 = MN-major operand
  int m_blocks, n_blocks, k_splits, kb_per_split;
  int epilogue, act;
  int b_resident;           // 1: the whole B operand (<= 4 k-blocks) is loaded once per CTA and stays in shared memory
  int stages;               // ring depth of this launch
  long long ldd;            // leading dimension of out (elements)
  long long ldmask;         // leading dimension of mask (elements)
  int mask_cols;            // the mask applies to output columns [0, mask_cols) only
  void* out;
  const float* bias;                  // [No] fp32 or null
  const __nv_bfloat16* mask;          // [Mo, ldmask] or null
  int tma_epi;                        // 1: epilogue through shared memory + TMA store / reduce-add
  uint32_t* relu_bits;                // forward, optional: bit (n & 31) of word [m, n / 32] = (out[m, n] > 0)
  const uint32_t* mask_bits;          // dgrad, optional: same layout, replaces `mask`
  int bits_wpr;                       // words per row of relu_bits / mask_bits
  float* colsum;                      // split-K mode, optional: colsum[m] += sum_r A(m, r)  (the bias gradient of a wgrad)
  // Low-precision inference (forward only): A and B are FP8 E4M3 (one byte per element, K-major), products accumulate in fp32
  // on the tensor core (tcgen05.mma kind::f8f6f4), alpha = scale_a * scale_b brings the accumulator back to real units before
  // bias / activation, and the output is bf16 or -- quantised for the next FP8 layer -- E4M3 of (value * out_qscale).
  int fp8;
  float alpha;
  int out_fp8;
  float out_qscale;
};

__device__ __forceinline__ uint32_t pack_e4m3x4(float a, float b, float c, float d) {
  uint16_t lo, hi;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(lo) : "f"(a), "f"(b));      // (upper, lower) operand order
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(hi) : "f"(c), "f"(d));
  return (uint32_t)lo | ((uint32_t)hi << 16);
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; FP8 (E4M3 / E5M2 per the instruction descriptor) -> fp32, K = 32 per instruction.
__device__ __forceinline__ void umma_f8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n"
      "}\n"
      :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// kind::f8f6f4 instruction descriptor: fp32 accumulator, A and B E4M3 (format 0), both K-major.
__host__ __device__ constexpr uint32_t make_idesc_e4m3(uint32_t m, uint32_t n) {
  return (1u << 4) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

template <int ACT>
__device__ __forceinline__ float act_fn(float v) {
  if (ACT == ACT_RELU) return fmaxf(v, 0.f);
  if (ACT == ACT_GELU) return 0.5f * v * (1.f + erff(v * 0.70710678118654752f));  // exact erf GELU
  if (ACT == ACT_SIGMOID) return 1.f / (1.f + __expf(-v));
  if (ACT == ACT_TANH) return tanhf(v);
  return v;
}

__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(nthreads) : "memory");
}

// One 32-column chunk of the accumulator tile -> global memory.  FULL == true: the whole chunk is
// inside the matrix (no per-element guards, 16-byte accesses only).
template <int EPI, int ACT, bool FULL>
__device__ __forceinline__ void epilogue_chunk(const GemmArgs& g, const uint32_t (&r)[32], long long row, int col,
                                               const float* __restrict__ s_bias /* of this chunk */,
                                               const uint4 (&mk)[4]) {
  if (EPI == EPI_RED_F32) {
    float* o = reinterpret_cast<float*>(g.out) + row * g.ldd + col;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      if (FULL || col + j + 4 <= g.No) {
        red_add_f32x4(o + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                      __uint_as_float(r[j + 3]));
      } else {
#pragma unroll
        for (int t = 0; t < 4; ++t)
          if (col + j + t < g.No) atomicAdd(o + j + t, __uint_as_float(r[j + t]));
      }
    }
    return;
  }
  __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(g.out) + row * g.ldd + col;
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    float v[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) v[t] = __uint_as_float(r[j + t]);
    if (EPI == EPI_BIAS_ACT) {
      if (g.fp8) {
#pragma unroll
        for (int t = 0; t < 8; ++t) v[t] *= g.alpha;
      }
#pragma unroll
      for (int t = 0; t < 8; ++t) v[t] = act_fn<ACT>(v[t] + s_bias[j + t]);
      if (g.out_fp8) {       // quantised output for the next FP8 layer: one byte per element
        uint8_t* o8 = reinterpret_cast<uint8_t*>(g.out) + row * g.ldd + col + j;
        const float q = g.out_qscale;
        if (FULL || col + j + 8 <= g.No) {
          uint2 pk8;
          pk8.x = pack_e4m3x4(v[0] * q, v[1] * q, v[2] * q, v[3] * q);
          pk8.y = pack_e4m3x4(v[4] * q, v[5] * q, v[6] * q, v[7] * q);
          *reinterpret_cast<uint2*>(o8) = pk8;
        } else {
#pragma unroll
          for (int t = 0; t < 8; ++t)
            if (col + j + t < g.No) o8[t] = (uint8_t)(pack_e4m3x4(v[t] * q, 0.f, 0.f, 0.f) & 0xffu);
        }
        continue;
      }
    } else {  // EPI_MASK
      if (g.mask && col + j < g.mask_cols) {
        const uint4 m4 = mk[j >> 3];
        const uint32_t mw[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const float2 f = unpack_bf16(mw[t]);
          if (!(f.x > 0.f)) v[2 * t] = 0.f;
          if (!(f.y > 0.f)) v[2 * t + 1] = 0.f;
        }
      }
    }
    if (FULL || col + j + 8 <= g.No) {
      uint4 pk;
      pk.x = pack_bf16(v[0], v[1]); pk.y = pack_bf16(v[2], v[3]);
      pk.z = pack_bf16(v[4], v[5]); pk.w = pack_bf16(v[6], v[7]);
      *reinterpret_cast<uint4*>(o + j) = pk;
    } else {
#pragma unroll
      for (int t = 0; t < 8; ++t)
        if (col + j + t < g.No) o[j + t] = __float2bfloat16(v[t]);
    }
  }
}

template <int EPI, int ACT>
__global__ void __launch_bounds__(kThreads, 1)
dense_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a,
                     const __grid_constant__ CUtensorMap tmap_b,
                     const __grid_constant__ CUtensorMap tmap_out,
                     const GemmArgs g) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[kAccStages];
  __shared__ __align__(8) uint64_t tmem_empty_bar[kAccStages];
  __shared__ __align__(8) uint64_t b_full_bar;
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float s_bias[kMaxN];

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = threadIdx.x & 31u;
  // 128B-swizzled tiles must start on a 1024 B boundary.
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    if (g.tma_epi) tma_prefetch_desc(&tmap_out);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < g.stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < kAccStages; ++i) { mbar_init(&tmem_full_bar[i], 1); mbar_init(&tmem_empty_bar[i], kEpiWarps); }
    mbar_init(&b_full_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(&tmem_base_slot, 512);
  if (EPI == EPI_RED_F32 && g.colsum && warp == 3) {
    // B operand of the bias-gradient MMA: a 16 x 64 tile of ones (2 KiB; any swizzle of all-ones is all-ones).  It borrows
    // the first epilogue staging slot, which is not written before every MMA of this CTA's only work item has retired.
    uint32_t* ones = reinterpret_cast<uint32_t*>(smem_al + kEpiStageOff);
    for (int i = lane; i < 512; i += 32) ones[i] = 0x3F803F80u;
    fence_proxy_async_smem();
  }
  pdl_wait();        // everything above overlapped the previous kernel's tail; global memory is touched from here on
  if (EPI == EPI_BIAS_ACT && g.n_blocks == 1 && threadIdx.x < kMaxN)   // bias of the only N block, once
    s_bias[threadIdx.x] = (g.bias && (int)threadIdx.x < g.No) ? g.bias[threadIdx.x] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_launch_dependents();
  const uint32_t tmem_base = tmem_base_slot;

  const int total_work = g.m_blocks * g.n_blocks * g.k_splits;
  const int kbe = g.fp8 ? 2 * kBlockK : kBlockK;      // elements per k-block: always one 128-byte swizzle row
  const int total_kb = (g.R + kbe - 1) / kbe;
  const uint32_t a_stage_bytes = kABytes;
  const uint32_t b_boxes = g.b_mn ? (uint32_t)((g.block_n + 63) / 64) : 1u;
  const uint32_t b_stage_bytes = g.b_mn ? b_boxes * kMnBoxBytes : (uint32_t)g.block_n * 128u;
  // Weight-stationary mode: the layers here are skinny (R <= 256), so B (the weights) is at most 4 k-blocks = 128 KiB.
  // Streaming it through the ring again for every M tile makes the main loop L2-bandwidth bound (B is 2/3 of the
  // bytes of a stage); instead it is loaded once and the ring carries only A.
  const bool resident = g.b_resident != 0;
  const uint32_t a_ring_stride = resident ? (uint32_t)kABytes : (uint32_t)kStageBytes;
  const uint32_t b_res_base = (uint32_t)g.stages * kABytes;   // resident B lives behind the A ring (1024-byte aligned)
  const int nstages = g.stages;

  if (warp == 0) {
    // ===================== TMA producer (one elected lane) =====================
    if (elect_one()) {
      uint32_t stage = 0, phase = 0;
      if (resident && (int)blockIdx.x < total_work) {
        mbar_arrive_expect_tx(&b_full_bar, (uint32_t)total_kb * b_stage_bytes);
        for (int kb = 0; kb < total_kb; ++kb) {
          uint8_t* sb = smem_al + b_res_base + kb * b_stage_bytes;
          if (!g.b_mn) {
            tma_load_2d(sb, &tmap_b, &b_full_bar, kb * kbe, 0);
          } else {
            for (uint32_t bx = 0; bx < b_boxes; ++bx) tma_load_2d(sb + bx * kMnBoxBytes, &tmap_b, &b_full_bar, (int)bx * 64, kb * kBlockK);
          }
        }
      }
      for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
        const int ks = w % g.k_splits;
        const int n_blk = (w / g.k_splits) % g.n_blocks;
        const int m_blk = w / (g.k_splits * g.n_blocks);
        const int kb0 = ks * g.kb_per_split;
        const int kb1 = min(kb0 + g.kb_per_split, total_kb);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = smem_al + stage * a_ring_stride;
          uint8_t* sb = sa + kABytes;
          mbar_arrive_expect_tx(&full_bar[stage], resident ? a_stage_bytes : a_stage_bytes + b_stage_bytes);
          const int r0 = kb * kbe;
          if (!g.a_mn) {
            tma_load_2d(sa, &tmap_a, &full_bar[stage], r0, m_blk * kBlockM);
          } else {
            tma_load_2d(sa, &tmap_a, &full_bar[stage], m_blk * kBlockM, r0);
            tma_load_2d(sa + kMnBoxBytes, &tmap_a, &full_bar[stage], m_blk * kBlockM + 64, r0);
          }
          if (resident) {
            // B is already in shared memory
          } else if (!g.b_mn) {
            tma_load_2d(sb, &tmap_b, &full_bar[stage], r0, n_blk * g.block_n);
          } else {
            for (uint32_t bx = 0; bx < b_boxes; ++bx)
              tma_load_2d(sb + bx * kMnBoxBytes, &tmap_b, &full_bar[stage],
                          n_blk * g.block_n + (int)bx * 64, r0);
          }
          if (++stage == (uint32_t)nstages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one elected lane) =======================
    if (elect_one()) {
      const uint32_t idesc = g.fp8 ? make_idesc_e4m3(kBlockM, (uint32_t)g.block_n)
                                   : make_idesc_bf16(kBlockM, (uint32_t)g.block_n, (uint32_t)g.a_mn, (uint32_t)g.b_mn);
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
      if (resident && (int)blockIdx.x < total_work) {
        mbar_wait(&b_full_bar, 0);
        tc_fence_after();
      }
      const uint32_t idesc_ones = make_idesc_bf16(kBlockM, 16u, (uint32_t)g.a_mn, 0u);
      const uint64_t d_ones = make_smem_desc_sw128(smem_base + kEpiStageOff, 16, 1024);
      for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
        const int ks = w % g.k_splits;
        const int kb0 = ks * g.kb_per_split;
        const int kb1 = min(kb0 + g.kb_per_split, total_kb);
        const bool do_colsum = EPI == EPI_RED_F32 && g.colsum && ((w / g.k_splits) % g.n_blocks) == 0;
        mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * kMaxN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * a_ring_stride;
          const uint32_t sb = resident ? smem_base + b_res_base + (uint32_t)kb * b_stage_bytes : sa + kABytes;
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            // K-major: step 16 elements (32 B) inside the swizzled 128 B row.
            // MN-major: step 16 k-rows (16 x 128 B); LBO = next 64 MN elements (next TMA box).
            const uint64_t da = g.a_mn ? make_smem_desc_sw128(sa + k * 2048, kMnBoxBytes, 1024)
                                       : make_smem_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t db = g.b_mn ? make_smem_desc_sw128(sb + k * 2048, kMnBoxBytes, 1024)
                                       : make_smem_desc_sw128(sb + k * 32, 16, 1024);
            if (g.fp8) umma_f8(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);      // 32 one-byte elements = the same 32-byte step
            else umma_bf16(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            // row sums of A ride along as a 16-column MMA against ones (accumulator: the unused second TMEM stage)
            if (do_colsum) umma_bf16(tmem_base + kMaxN, da, d_ones, idesc_ones, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);             // frees the smem slot when the MMAs retire
          if (kb == kb1 - 1) umma_commit(&tmem_full_bar[acc]);  // accumulator complete
          if (++stage == (uint32_t)nstages) { stage = 0; phase ^= 1u; }
        }
        if (kb1 <= kb0) umma_commit(&tmem_full_bar[acc]);  // empty split: still hand the stage over
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else if (warp >= 4) {
    // ============ Epilogue: 8 warps (TMEM -> registers -> global) ==============
    // Warp w may only touch TMEM lanes 32*(w%4)..+31; warps 4-7 take the first half of the
    // 32-column chunks of a tile, warps 8-11 the second half.
    const uint32_t q = warp & 3u;
    const uint32_t half = (warp - 4u) >> 2;
    const int nchunks = (g.block_n + 31) / 32;
    const int c_begin = half == 0 ? 0 : (nchunks + 1) / 2;
    const int c_end = half == 0 ? (nchunks + 1) / 2 : nchunks;
    uint32_t acc = 0, acc_phase = 0;
    int bias_blk = g.n_blocks == 1 ? 0 : -1;
    if (g.tma_epi) {
      // ---- staged epilogue: registers -> swizzled smem box (32 rows x 128 B) -> TMA store / reduce-add ----
      const uint32_t slot = smem_base + kEpiStageOff + (warp - 4u) * kEpiSlotBytes;   // 1024-byte aligned
      const uint32_t my_row = slot + lane * 128u;
      const uint32_t swz = lane & 7u;
      constexpr int UNIT = (EPI == EPI_RED_F32) ? 32 : 64;     // output columns per box (128 bytes)
      const int nunits = (g.block_n + UNIT - 1) / UNIT;
      const int u_begin = half == 0 ? 0 : (nunits + 1) / 2;
      const int u_end = half == 0 ? (nunits + 1) / 2 : nunits;
      for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
        const int ks = w % g.k_splits;
        const int n_blk = (w / g.k_splits) % g.n_blocks;
        const int m_blk = w / (g.k_splits * g.n_blocks);
        const bool has_work = (ks * g.kb_per_split) < total_kb;
        if (EPI == EPI_BIAS_ACT && n_blk != bias_blk) {   // N > 256: reload this block's bias slice
          named_bar_sync(1, kEpiWarps * 32);
          const int t = threadIdx.x - 128;
          const int n = n_blk * g.block_n + t;
          if (t < kMaxN) s_bias[t] = (g.bias && t < g.block_n && n < g.No) ? g.bias[n] : 0.f;
          named_bar_sync(1, kEpiWarps * 32);
          bias_blk = n_blk;
        }
        mbar_wait(&tmem_full_bar[acc], acc_phase);
        tc_fence_after();
        const int row0 = m_blk * kBlockM + (int)q * 32;
        const long long row = (long long)row0 + lane;
        const bool row_ok = row < g.Mo;
        const int col0 = n_blk * g.block_n;
        const uint32_t taddr = tmem_base + acc * kMaxN + ((q * 32u) << 16);
        for (int u = u_begin; u < u_end; ++u) {
          const int col = col0 + u * UNIT;
          if (EPI == EPI_RED_F32) {
            uint32_t r[32];
            tmem_ld_32x32(taddr + u * 32, r);
            tmem_ld_wait();
            if (lane == 0) bulk_wait_read0();              // the previous box has left shared memory
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j) st_shared_u4(my_row + ((j ^ swz) << 4), r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && has_work && col < g.No && row0 < g.Mo) { tma_reduce_add_2d(&tmap_out, slot, col, row0); bulk_commit(); }
            if (g.colsum && n_blk == 0 && half == 0 && u == u_begin) {       // bias gradient: TMEM column 256 of this lane's row
              uint32_t cs[32];
              tmem_ld_32x32(tmem_base + kMaxN + ((q * 32u) << 16), cs);
              tmem_ld_wait();
              if (has_work && row_ok) atomicAdd(g.colsum + row, __uint_as_float(cs[0]));
            }
          } else {
            uint32_t r[2][32];
            const bool second = u * 64 + 32 < g.block_n;
            tmem_ld_32x32(taddr + u * 64, r[0]);
            if (second) tmem_ld_32x32(taddr + u * 64 + 32, r[1]);
            uint32_t mw[2] = {0xffffffffu, 0xffffffffu};
            uint4 mk[2][4];
            bool use_mk[2] = {false, false};
            if (EPI == EPI_MASK) {                          // issue the mask loads under the TMEM latency
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                const int ch = col + h * 32;
                if (ch >= g.mask_cols || ch >= g.No || !row_ok) continue;
                if (g.mask_bits) {
                  mw[h] = __ldg(g.mask_bits + row * g.bits_wpr + (ch >> 5));
                } else if (g.mask) {
                  use_mk[h] = true;
                  if (ch + 32 <= g.No) {
                    const uint4* mp = reinterpret_cast<const uint4*>(g.mask + row * g.ldmask + ch);
#pragma unroll
                    for (int t = 0; t < 4; ++t) mk[h][t] = ld_nc_u4(mp + t);
                  } else {
                    __nv_bfloat16 tmp[32];
                    for (int t = 0; t < 32; ++t) tmp[t] = (ch + t < g.No) ? g.mask[row * g.ldmask + ch + t] : __float2bfloat16(0.f);
#pragma unroll
                    for (int t = 0; t < 4; ++t) mk[h][t] = reinterpret_cast<const uint4*>(tmp)[t];
                  }
                }
              }
            }
            tmem_ld_wait();
            uint32_t pk[2][16];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              uint32_t bits = 0;
#pragma unroll
              for (int j = 0; j < 32; j += 2) {
                float v0 = __uint_as_float(r[h][j]), v1 = __uint_as_float(r[h][j + 1]);
                if (EPI == EPI_BIAS_ACT) {
                  v0 = act_fn<ACT>(v0 + s_bias[u * 64 + h * 32 + j]);
                  v1 = act_fn<ACT>(v1 + s_bias[u * 64 + h * 32 + j + 1]);
                  bits |= (v0 > 0.f ? 1u : 0u) << j;
                  bits |= (v1 > 0.f ? 1u : 0u) << (j + 1);
                } else {
                  if (use_mk[h]) {
                    const uint4 m4 = mk[h][j >> 3];
                    const uint32_t m32 = ((j >> 1) & 3) == 0 ? m4.x : ((j >> 1) & 3) == 1 ? m4.y : ((j >> 1) & 3) == 2 ? m4.z : m4.w;
                    const float2 f = unpack_bf16(m32);
                    if (!(f.x > 0.f)) v0 = 0.f;
                    if (!(f.y > 0.f)) v1 = 0.f;
                  } else {
                    if (!((mw[h] >> j) & 1u)) v0 = 0.f;
                    if (!((mw[h] >> (j + 1)) & 1u)) v1 = 0.f;
                  }
                }
                pk[h][j >> 1] = pack_bf16(v0, v1);
              }
              if (EPI == EPI_BIAS_ACT && g.relu_bits && row_ok && (h == 0 || second) && col + h * 32 < g.No)
                g.relu_bits[row * g.bits_wpr + ((col + h * 32) >> 5)] = bits;
            }
            if (lane == 0) bulk_wait_read0();
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int h = j >> 2, o = (j & 3) * 4;
              st_shared_u4(my_row + ((j ^ swz) << 4), pk[h][o], pk[h][o + 1], pk[h][o + 2], pk[h][o + 3]);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && has_work && col < g.No && row0 < g.Mo) { tma_store_2d(&tmap_out, slot, col, row0); bulk_commit(); }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
      }
      if (lane == 0) bulk_wait_all();                      // all stores of this warp are complete before the CTA exits
    } else
    for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
      const int ks = w % g.k_splits;
      const int n_blk = (w / g.k_splits) % g.n_blocks;
      const int m_blk = w / (g.k_splits * g.n_blocks);
      const bool has_work = (ks * g.kb_per_split) < total_kb;
      if (EPI == EPI_BIAS_ACT && n_blk != bias_blk) {   // N > 256: reload this block's bias slice
        named_bar_sync(1, kEpiWarps * 32);
        const int t = threadIdx.x - 128;
        const int n = n_blk * g.block_n + t;
        if (t < kMaxN) s_bias[t] = (g.bias && t < g.block_n && n < g.No) ? g.bias[n] : 0.f;
        named_bar_sync(1, kEpiWarps * 32);
        bias_blk = n_blk;
      }
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tc_fence_after();
      const long long row = (long long)m_blk * kBlockM + q * 32 + lane;
      const int col0 = n_blk * g.block_n;
      const uint32_t taddr = tmem_base + acc * kMaxN + ((q * 32u) << 16);
      const bool row_ok = row < g.Mo && has_work;
      for (int c = c_begin; c < c_end; ++c) {
        const int col = col0 + c * 32;
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        uint4 mk[4];
        const bool full = col + 32 <= g.No;
        if (EPI == EPI_MASK && g.mask && row_ok && full && col < g.mask_cols) {   // issue the mask loads under the TMEM latency
          const uint4* mp = reinterpret_cast<const uint4*>(g.mask + row * g.ldmask + col);
#pragma unroll
          for (int t = 0; t < 4; ++t) mk[t] = ld_nc_u4(mp + t);
        }
        tmem_ld_wait();
        if (row_ok) {
          if (full) {
            epilogue_chunk<EPI, ACT, true>(g, r, row, col, s_bias + c * 32, mk);
          } else {
            if (EPI == EPI_MASK && g.mask && col < g.mask_cols) {
              const __nv_bfloat16* mp = g.mask + row * g.ldmask + col;
              __nv_bfloat16 tmp[32];
#pragma unroll
              for (int t = 0; t < 32; ++t) tmp[t] = (col + t < g.No) ? mp[t] : __float2bfloat16(0.f);
#pragma unroll
              for (int t = 0; t < 4; ++t) mk[t] = reinterpret_cast<const uint4*>(tmp)[t];
            }
            epilogue_chunk<EPI, ACT, false>(g, r, row, col, s_bias + c * 32, mk);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
      if (++acc == kAccStages) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ----------------------------------------------------------------- host side --
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

// rows x cols bf16 matrix, row-major with leading dimension ld (elements).
static int make_tmap(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld,
                     int box_cols, int box_rows, bool f32 = false, bool u8 = false) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) { zoo_set_last_error("cuTensorMapEncodeTiled unavailable", __FILE__, __LINE__); return ZOO_ERR_CUDA; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * (f32 ? 4 : (u8 ? 1 : 2))};
  const CUtensorMapDataType dt = f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : (u8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    // A thread that has issued no CUDA runtime call yet (e.g. an autograd worker whose first backward node is
    // one of our kernels) has no context bound; the runtime binds the primary context of its current device
    // on the first real call.  Do that and encode again.
    cudaFree(0);
    r = enc(m, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) {
    char msg[256];
    snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled failed (CUresult %d): base %p rows %lld cols %lld ld %lld box %dx%d -- "
             "base must be 16-byte aligned and ld a multiple of 8 elements", (int)r, base, rows, cols, ld, box_rows, box_cols);
    zoo_set_last_error(msg, __FILE__, __LINE__);
    return ZOO_ERR_CUDA;
  }
  return ZOO_OK;
}

static int g_num_sms = 0;
static int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

static int launch(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn,
                  GemmArgs g, int max_ctas, cudaStream_t stream) {
  if (!A || !B || !g.out || g.Mo <= 0 || g.No <= 0 || g.R <= 0) {
    zoo_set_last_error("dense_tcgen05: null pointer or empty shape", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  const int ld_mult = g.fp8 ? 16 : 8;
  if ((lda % ld_mult) || (ldb % ld_mult) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15)) {
    zoo_set_last_error("dense_tcgen05: operands need 16-byte aligned rows (ld % 8 == 0 for bf16, % 16 for fp8)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (g.fp8 && (a_mn || b_mn || g.epilogue != EPI_BIAS_ACT)) {
    zoo_set_last_error("dense_tcgen05: the FP8 mode is forward only (K-major operands, bias/activation epilogue)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  g.a_mn = a_mn; g.b_mn = b_mn;
  // UMMA N: whole output width when it fits, rounded up to a multiple of 16.
  int bn = g.No >= kMaxN ? kMaxN : ((g.No + 15) / 16) * 16;
  g.block_n = bn;
  g.m_blocks = (g.Mo + kBlockM - 1) / kBlockM;
  g.n_blocks = (g.No + bn - 1) / bn;
  const int kbe = g.fp8 ? 2 * kBlockK : kBlockK;
  const int total_kb = (g.R + kbe - 1) / kbe;
  if (g.epilogue == EPI_RED_F32) {
    int want = max_ctas / (g.m_blocks * g.n_blocks);
    if (want < 1) want = 1;
    if (want > total_kb) want = total_kb;
    g.kb_per_split = (total_kb + want - 1) / want;
    g.k_splits = (total_kb + g.kb_per_split - 1) / g.kb_per_split;
  } else {
    g.k_splits = 1;
    g.kb_per_split = total_kb;
  }
  // weight-stationary B: one N block, no split, and all of B's k-blocks fit behind the A ring
  static int allow_resident = -1;
  if (allow_resident < 0) { const char* ev = getenv("ZOO_DENSE_RESIDENT"); allow_resident = ev ? atoi(ev) : 1; }
  const int b_stage = b_mn ? ((bn + 63) / 64) * kMnBoxBytes : bn * 128;
  g.b_resident = (allow_resident && g.k_splits == 1 && g.n_blocks == 1 && g.m_blocks > 1 &&
                  kStages * kABytes + total_kb * b_stage <= kEpiStageOff) ? 1 : 0;
  g.stages = kStages;
  if (g.b_resident) {
    static int max_stages = -1;
    if (max_stages < 0) { const char* ev = getenv("ZOO_DENSE_ASTAGES"); max_stages = ev ? atoi(ev) : kMaxStages; if (max_stages < 2 || max_stages > kMaxStages) max_stages = kMaxStages; }
    g.stages = (kEpiStageOff - total_kb * b_stage) / kABytes;
    if (g.stages > max_stages) g.stages = max_stages;
  }
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn) rc = make_tmap(&ta, A, g.Mo, g.R, lda, kbe, kBlockM, false, g.fp8 != 0);
  else       rc = make_tmap(&ta, A, g.R, g.Mo, lda, 64, kBlockK);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap(&tb, B, g.No, g.R, ldb, kbe, bn, false, g.fp8 != 0);
  else       rc = make_tmap(&tb, B, g.R, g.No, ldb, 64, kBlockK);
  if (rc) return rc;
  static int allow_tma_epi = -1;
  if (allow_tma_epi < 0) { const char* ev = getenv("ZOO_DENSE_TMA_EPI"); allow_tma_epi = ev ? atoi(ev) : 1; }
  g.tma_epi = (allow_tma_epi || g.relu_bits || g.mask_bits || g.colsum) ? 1 : 0;   // the bit-mask forms exist only in the staged epilogue
  if (g.fp8) g.tma_epi = 0;            // scaled / one-byte outputs leave through the direct epilogue
  CUtensorMap to = ta;
  if (g.tma_epi) {
    if (g.epilogue == EPI_RED_F32) rc = make_tmap(&to, g.out, g.Mo, g.No, g.ldd, 32, 32, true);
    else                           rc = make_tmap(&to, g.out, g.Mo, g.No, g.ldd, 64, 32);
    if (rc) return rc;
  }
  const int total_work = g.m_blocks * g.n_blocks * g.k_splits;
  const int grid = total_work < max_ctas ? total_work : max_ctas;
#define ZOO_LAUNCH(EPI, ACT)                                                                                   \
  do {                                                                                                         \
    static bool attr_set = false;                                                                              \
    if (!attr_set) {                                                                                           \
      ZOO_CUDA_CHECK(cudaFuncSetAttribute(dense_tcgen05_kernel<EPI, ACT>,                                      \
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));          \
      attr_set = true;                                                                                         \
    }                                                                                                          \
    ZOO_CUDA_CHECK(zoo_launch_pdl(dense_tcgen05_kernel<EPI, ACT>, dim3(grid), dim3(kThreads), kSmemBytes, stream, ta, tb, to, g));                        \
  } while (0)
  if (g.epilogue == EPI_RED_F32) ZOO_LAUNCH(EPI_RED_F32, ACT_NONE);
  else if (g.epilogue == EPI_MASK) ZOO_LAUNCH(EPI_MASK, ACT_NONE);
  else if (g.act == ACT_RELU) ZOO_LAUNCH(EPI_BIAS_ACT, ACT_RELU);
  else if (g.act == ACT_GELU) ZOO_LAUNCH(EPI_BIAS_ACT, ACT_GELU);
  else if (g.act == ACT_SIGMOID) ZOO_LAUNCH(EPI_BIAS_ACT, ACT_SIGMOID);
  else if (g.act == ACT_TANH) ZOO_LAUNCH(EPI_BIAS_ACT, ACT_TANH);
  else ZOO_LAUNCH(EPI_BIAS_ACT, ACT_NONE);
#undef ZOO_LAUNCH
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

}  // namespace zoo

using namespace zoo;

// Y[M,N] (bf16, ldy) = act(X[M,K] (bf16, ldx) * W[N,K]^T (bf16, ldw) + bias[N] (fp32 or null))
ZOO_API int zoo_dense_fwd_bf16(const void* x, long long ldx, const void* w, long long ldw, const float* bias,
                               void* y, long long ldy, int M, int N, int K, int act, void* stream) {
  GemmArgs g{};
  g.Mo = M; g.No = N; g.R = K; g.epilogue = EPI_BIAS_ACT; g.act = act;
  g.out = y; g.ldd = ldy; g.bias = bias; g.mask = nullptr; g.ldmask = 0;
  if (ldy % 8 || ((uintptr_t)y & 15)) { zoo_set_last_error("dense_fwd: ldy % 8 != 0", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  return launch(x, ldx, 0, w, ldw, 0, g, num_sms(), (cudaStream_t)stream);
}

// Forward that also emits the sign bits of its output: bit (n & 31) of relu_bits[m * bits_wpr + n / 32] = (Y[m, n] > 0).
// The next layer's dgrad reads them instead of re-reading Y as its ReLU mask (1/16 of the bytes).
ZOO_API int zoo_dense_fwd_bf16_bits(const void* x, long long ldx, const void* w, long long ldw, const float* bias,
                                    void* y, long long ldy, int M, int N, int K, int act, unsigned int* relu_bits,
                                    int bits_wpr, void* stream) {
  GemmArgs g{};
  g.Mo = M; g.No = N; g.R = K; g.epilogue = EPI_BIAS_ACT; g.act = act;
  g.out = y; g.ldd = ldy; g.bias = bias; g.mask = nullptr; g.ldmask = 0;
  g.relu_bits = relu_bits; g.bits_wpr = bits_wpr;
  if (ldy % 8 || ((uintptr_t)y & 15)) { zoo_set_last_error("dense_fwd: ldy % 8 != 0", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (relu_bits && bits_wpr * 32 < N) { zoo_set_last_error("dense_fwd: bits_wpr too small", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  return launch(x, ldx, 0, w, ldw, 0, g, num_sms(), (cudaStream_t)stream);
}

// dgrad whose ReLU mask is the bit matrix written by zoo_dense_fwd_bf16_bits (mask_cols must be a multiple of 32 or K).
ZOO_API int zoo_dense_dgrad_bf16_bits(const void* dy, long long lddy, const void* w, long long ldw, const unsigned int* mask_bits,
                                      int bits_wpr, int mask_cols, void* dx, long long lddx, int M, int N, int K, void* stream) {
  GemmArgs g{};
  g.Mo = M; g.No = K; g.R = N; g.epilogue = EPI_MASK; g.act = ACT_NONE;
  g.out = dx; g.ldd = lddx; g.bias = nullptr; g.mask = nullptr; g.ldmask = 0;
  g.mask_bits = mask_bits; g.bits_wpr = bits_wpr;
  g.mask_cols = (mask_cols <= 0 || mask_cols > K) ? K : mask_cols;
  if (!mask_bits || (g.mask_cols % 32 && g.mask_cols != K) || bits_wpr * 32 < g.mask_cols) {
    zoo_set_last_error("dense_dgrad_bits: need mask_bits, mask_cols % 32 == 0 and bits_wpr * 32 >= mask_cols", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (lddx % 8 || ((uintptr_t)dx & 15)) { zoo_set_last_error("dense_dgrad: output rows must be 16-byte aligned", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  return launch(dy, lddy, 0, w, ldw, 1, g, num_sms(), (cudaStream_t)stream);
}

// dX[M,K] (bf16, lddx) = (dY[M,N] (bf16, lddy) * W[N,K] (bf16, ldw)) * (mask[M,K] > 0)   (mask may be null)
// mask_cols: the ReLU mask covers dX columns [0, mask_cols) (multiple of 8; <= 0 means all K columns).
ZOO_API int zoo_dense_dgrad_bf16(const void* dy, long long lddy, const void* w, long long ldw, const void* mask,
                                 long long ldmask, int mask_cols, void* dx, long long lddx, int M, int N, int K,
                                 void* stream) {
  GemmArgs g{};
  g.Mo = M; g.No = K; g.R = N; g.epilogue = EPI_MASK; g.act = ACT_NONE;
  g.out = dx; g.ldd = lddx; g.bias = nullptr; g.mask = (const __nv_bfloat16*)mask; g.ldmask = ldmask;
  g.mask_cols = (mask_cols <= 0 || mask_cols > K) ? K : mask_cols;
  if (g.mask_cols % 8 && g.mask_cols != K) { zoo_set_last_error("dense_dgrad: mask_cols must be a multiple of 8", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (lddx % 8 || ((uintptr_t)dx & 15) || (mask && (ldmask % 8 || ((uintptr_t)mask & 15)))) {
    zoo_set_last_error("dense_dgrad: output/mask rows must be 16-byte aligned", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  return launch(dy, lddy, 0, w, ldw, 1, g, num_sms(), (cudaStream_t)stream);
}

extern "C" int zoo_colsum_bf16(const void* x, long long ldx, int rows, int cols, float* out, void* stream);

// zoo_dense_wgrad_bf16 that also accumulates the bias gradient db[N] += column sums of dY.  When every CTA owns a single
// split (always the case for layers up to 256 x 256) the sums come out of the same kernel as one extra 16-column MMA per
// k-step; otherwise a separate column-sum kernel runs.
ZOO_API int zoo_dense_wgrad_bias_bf16(const void* dy, long long lddy, const void* x, long long ldx, float* dw,
                                      long long lddw, float* db, int M, int N, int K, void* stream) {
  GemmArgs g{};
  g.Mo = N; g.No = K; g.R = M; g.epilogue = EPI_RED_F32; g.act = ACT_NONE;
  g.out = dw; g.ldd = lddw; g.bias = nullptr; g.mask = nullptr; g.ldmask = 0;
  if (lddw % 4 || ((uintptr_t)dw & 15)) { zoo_set_last_error("dense_wgrad: lddw % 4 != 0", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  const int m_blocks = (N + kBlockM - 1) / kBlockM, n_blocks = (K + kMaxN - 1) / kMaxN;
  static int allow_fused = -1;
  if (allow_fused < 0) { const char* ev = getenv("ZOO_DENSE_FUSED_COLSUM"); allow_fused = ev ? atoi(ev) : 1; }
  const bool fused = db && allow_fused && m_blocks * n_blocks <= num_sms();
  if (db && !fused) {
    int rc = zoo_colsum_bf16(dy, lddy, M, N, db, stream);
    if (rc) return rc;
  }
  g.colsum = fused ? db : nullptr;
  return launch(dy, lddy, 1, x, ldx, 1, g, num_sms(), (cudaStream_t)stream);
}

// dW[N,K] (fp32, lddw) += dY[M,N]^T (bf16, lddy) * X[M,K] (bf16, ldx); split over the batch dimension.
ZOO_API int zoo_dense_wgrad_bf16(const void* dy, long long lddy, const void* x, long long ldx, float* dw,
                                 long long lddw, int M, int N, int K, void* stream) {
  GemmArgs g{};
  g.Mo = N; g.No = K; g.R = M; g.epilogue = EPI_RED_F32; g.act = ACT_NONE;
  g.out = dw; g.ldd = lddw; g.bias = nullptr; g.mask = nullptr; g.ldmask = 0;
  if (lddw % 4 || ((uintptr_t)dw & 15)) { zoo_set_last_error("dense_wgrad: lddw % 4 != 0", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  return launch(dy, lddy, 1, x, ldx, 1, g, num_sms(), (cudaStream_t)stream);
}

// FP8 inference GEMM: Y[M,N] = act(alpha * (Xq[M,K] (E4M3, ldx) * Wq[N,K]^T (E4M3, ldw)) + bias[N]); alpha = scale_x * scale_w.
// out_fp8 == 0: Y is bf16 (ldy elements).  out_fp8 != 0: Y is E4M3 of (value * out_qscale) -- the quantised input of the next
// FP8 layer, ldy in bytes.  Stands where the reference runs OpenVINO / BigDL int8 inference
// (zoo/.../pipeline/inference/OpenVinoInferenceSupportive.scala:35-57; examples/vnni): 8-bit operands, 32-bit accumulation.
ZOO_API int zoo_dense_fwd_fp8(const void* xq, long long ldx, const void* wq, long long ldw, const float* bias, void* y, long long ldy,
                              int M, int N, int K, int act, float alpha, int out_fp8, float out_qscale, void* stream) {
  GemmArgs g{};
  g.Mo = M; g.No = N; g.R = K; g.epilogue = EPI_BIAS_ACT; g.act = act;
  g.out = y; g.ldd = ldy; g.bias = bias; g.mask = nullptr; g.ldmask = 0;
  g.fp8 = 1; g.alpha = alpha; g.out_fp8 = out_fp8 ? 1 : 0; g.out_qscale = out_qscale;
  if ((out_fp8 ? (ldy % 16) : (ldy % 8)) || ((uintptr_t)y & 15)) { zoo_set_last_error("dense_fwd_fp8: output rows must be 16-byte aligned", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  return launch(xq, ldx, 0, wq, ldw, 0, g, num_sms(), (cudaStream_t)stream);
}
