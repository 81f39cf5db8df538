// K3, whole-tower form: three Dense + ReLU layers (K0 -> N1 -> N2 -> N3) in ONE persistent tcgen05 kernel.
//
// The recommendation towers are three skinny layers (NeuralCF 256 -> 256 -> 128 -> 64, zoo/.../models/recommendation/
// NeuralCF.scala:73-89; Wide&Deep's deep part, WideAndDeep.scala:146-157).  Layer by layer (csrc/gemm/dense_tcgen05.cu) every layer
// writes its activation to HBM and the next one reads it back, and every launch pays its prologue / fill / drain at 3.5 tiles per
// CTA: 40 us for 151 MB of traffic.  Here a CTA walks 128-row tiles of the input and keeps the intermediate activations on chip:
//
//   warp 0  TMA producer : per tile, the k-blocks of X with W1 (layer 1) and then the k-blocks of W2 (layer 2) through one ring;
//                          W3 is loaded once and stays resident
//   warp 1  MMA issuer   : acc1 = X W1^T (TMEM columns [0, 256)), acc2 = h1 W2^T ([256, 384)), acc3 = h2 W3^T ([384, 448));
//                          the A operands of layers 2 and 3 are the shared-memory tiles the epilogue warps wrote
//   warps 4-11 epilogue  : TMEM -> registers -> bias + ReLU -> bf16 -> the 128B-swizzled K-major tile H in shared memory, which is
//                          BOTH the source of the TMA store of the activation (backward needs h1, h2, h3) and the A operand of
//                          the next layer's MMA; ReLU sign bits go to global memory for the dgrad epilogues
//
// HBM traffic: X read once, h1 / h2 / h3 written once (92 MB instead of 151 MB), one launch instead of three.
#include "../common/zoo_common.cuh"

namespace zoo {

constexpr int kT3Stages = 3;
constexpr int kT3ABytes = 128 * 64 * 2;                 // 16 KiB: one k-block of a 128-row A tile
constexpr int kT3BBytes = 256 * 64 * 2;                 // 32 KiB: one k-block of up to 256 weight rows
constexpr int kT3Stage = kT3ABytes + kT3BBytes;         // 48 KiB
constexpr int kT3OffH = kT3Stages * kT3Stage;           // H: 4 k-blocks x 16 KiB (h1; h2 reuses blocks 0-1, h3 staging block 2)
constexpr int kT3OffW3 = kT3OffH + 4 * kT3ABytes;       // W3: up to 2 k-blocks x (64 rows x 128 B)
constexpr int kT3OffBar = kT3OffW3 + 2 * 8192;          // barriers, TMEM slot, biases
constexpr int kT3Smem = kT3OffBar + 256 + (256 + 128 + 64) * 4;
constexpr int kT3Threads = 384;                         // 4 control warps + 8 epilogue warps

struct Mlp3Args {
  int M, K0, N1, N2, N3;
  const float* b1; const float* b2; const float* b3;
  uint32_t* bits1; uint32_t* bits2; uint32_t* bits3;    // optional ReLU sign bits, [M, N / 32] words per row
  int m_blocks;
};

__device__ __forceinline__ void t3_named_bar(uint32_t id, uint32_t n) { asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(n) : "memory"); }

// One 64-column unit of an accumulator: bias + ReLU, sign bits, bf16 into rows [32q, 32q+32) of k-block `unit` of the H tile,
// then one TMA store of that 32 x 64 box.
__device__ __forceinline__ void t3_unit(uint32_t tmem_acc, uint32_t lane_addr, int unit, const float* s_bias, uint32_t h_block /* smem addr of k-block */,
                                        uint32_t q, uint32_t lane, const CUtensorMap* tmap_out, int row0, long long row, bool row_ok,
                                        uint32_t* bits, int bits_wpr, int m_rows) {
  uint32_t r0[32], r1[32];
  tmem_ld_32x32(tmem_acc + lane_addr + unit * 64, r0);
  tmem_ld_32x32(tmem_acc + lane_addr + unit * 64 + 32, r1);
  tmem_ld_wait();
  uint32_t pk[32];
  uint32_t bw[2] = {0u, 0u};
#pragma unroll
  for (int j = 0; j < 32; j += 2) {
    const float a0 = fmaxf(__uint_as_float(r0[j]) + s_bias[unit * 64 + j], 0.f), a1 = fmaxf(__uint_as_float(r0[j + 1]) + s_bias[unit * 64 + j + 1], 0.f);
    const float c0 = fmaxf(__uint_as_float(r1[j]) + s_bias[unit * 64 + 32 + j], 0.f), c1 = fmaxf(__uint_as_float(r1[j + 1]) + s_bias[unit * 64 + 32 + j + 1], 0.f);
    bw[0] |= (a0 > 0.f ? 1u : 0u) << j; bw[0] |= (a1 > 0.f ? 1u : 0u) << (j + 1);
    bw[1] |= (c0 > 0.f ? 1u : 0u) << j; bw[1] |= (c1 > 0.f ? 1u : 0u) << (j + 1);
    pk[j >> 1] = pack_bf16(a0, a1);
    pk[16 + (j >> 1)] = pack_bf16(c0, c1);
  }
  if (bits && row_ok) {
    bits[row * bits_wpr + unit * 2] = bw[0];
    bits[row * bits_wpr + unit * 2 + 1] = bw[1];
  }
  const uint32_t my_row = h_block + (q * 32u + lane) * 128u;
  const uint32_t swz = lane & 7u;                      // (32 q + lane) & 7
#pragma unroll
  for (int p = 0; p < 8; ++p) st_shared_u4(my_row + (((uint32_t)p ^ swz) << 4), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
  fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0 && row0 + (int)q * 32 < m_rows) { tma_store_2d(tmap_out, h_block + q * 4096u, unit * 64, row0 + (int)q * 32); bulk_commit(); }
}

__global__ void __launch_bounds__(kT3Threads, 1)
mlp3_fwd_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w1,
                const __grid_constant__ CUtensorMap tmap_w2, const __grid_constant__ CUtensorMap tmap_w3,
                const __grid_constant__ CUtensorMap tmap_h1, const __grid_constant__ CUtensorMap tmap_h2,
                const __grid_constant__ CUtensorMap tmap_h3, const Mlp3Args a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = smem_u32(smem_raw);
  if (smem_base & 1023u) { if (threadIdx.x == 0) printf("zoo mlp3: shared memory base is not 1024-byte aligned\n"); __trap(); }
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + kT3OffBar);
  uint64_t* full_bar = bars;                 // [3]
  uint64_t* empty_bar = bars + 3;            // [3]
  uint64_t* acc_full = bars + 6;             // [3] accumulator of layer l complete
  uint64_t* acc_empty = bars + 9;            // [3] accumulator of layer l read back by all epilogue warps
  uint64_t* h_ready = bars + 12;             // [2] h1 / h2 tile written to shared memory
  uint64_t& w3_bar = bars[14];
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(bars + 15);
  float* s_b1 = reinterpret_cast<float*>(smem_raw + kT3OffBar + 256);
  float* s_b2 = s_b1 + 256;
  float* s_b3 = s_b2 + 128;
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const int kb0n = a.K0 / 64, kb1n = a.N1 / 64, kb2n = a.N2 / 64;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x); tma_prefetch_desc(&tmap_w1); tma_prefetch_desc(&tmap_w2); tma_prefetch_desc(&tmap_w3);
    tma_prefetch_desc(&tmap_h1); tma_prefetch_desc(&tmap_h2); tma_prefetch_desc(&tmap_h3);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kT3Stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 3; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 8); }
    mbar_init(&h_ready[0], 8); mbar_init(&h_ready[1], 8);
    mbar_init(&w3_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(&tmem_slot, 512);
  pdl_wait();
  for (int i = threadIdx.x; i < 256; i += kT3Threads) s_b1[i] = (a.b1 && i < a.N1) ? a.b1[i] : 0.f;
  for (int i = threadIdx.x; i < 128; i += kT3Threads) s_b2[i] = (a.b2 && i < a.N2) ? a.b2[i] : 0.f;
  for (int i = threadIdx.x; i < 64; i += kT3Threads) s_b3[i] = (a.b3 && i < a.N3) ? a.b3[i] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_launch_dependents();
  const uint32_t tmem = tmem_slot;
  const uint32_t acc1 = tmem, acc2 = tmem + 256, acc3 = tmem + 384;
  const uint32_t sH = smem_base + kT3OffH, sW3 = smem_base + kT3OffW3;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      if ((int)blockIdx.x < a.m_blocks) {
        mbar_arrive_expect_tx(&w3_bar, (uint32_t)kb2n * (uint32_t)a.N3 * 128u);
        for (int kb = 0; kb < kb2n; ++kb) tma_load_2d(smem_raw + kT3OffW3 + kb * 8192, &tmap_w3, &w3_bar, kb * 64, 0);
      }
      uint32_t stage = 0, phase = 0;
      for (int tile = blockIdx.x; tile < a.m_blocks; tile += gridDim.x) {
        for (int kb = 0; kb < kb0n; ++kb) {                       // layer 1: X k-block + W1 k-block
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = smem_raw + stage * kT3Stage;
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)kT3ABytes + (uint32_t)a.N1 * 128u);
          tma_load_2d(sa, &tmap_x, &full_bar[stage], kb * 64, tile * 128);
          tma_load_2d(sa + kT3ABytes, &tmap_w1, &full_bar[stage], kb * 64, 0);
          if (++stage == kT3Stages) { stage = 0; phase ^= 1u; }
        }
        for (int kb = 0; kb < kb1n; ++kb) {                       // layer 2: W2 k-block only (A is the on-chip h1 tile)
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = smem_raw + stage * kT3Stage;
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)a.N2 * 128u);
          tma_load_2d(sa + kT3ABytes, &tmap_w2, &full_bar[stage], kb * 64, 0);
          if (++stage == kT3Stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (elect_one() && (int)blockIdx.x < a.m_blocks) {
      const uint32_t id1 = make_idesc_bf16(128u, (uint32_t)a.N1, 0u, 0u), id2 = make_idesc_bf16(128u, (uint32_t)a.N2, 0u, 0u),
                     id3 = make_idesc_bf16(128u, (uint32_t)a.N3, 0u, 0u);
      uint32_t stage = 0, phase = 0;
      mbar_wait(&w3_bar, 0);
      tc_fence_after();
      // Layer 1 of the NEXT tile is issued between layers 2 and 3 of the current one (its accumulator is free once h1 is out), so its
      // loads and MMAs run under the epilogues of layers 2 and 3; the ring is consumed in the producer's order L1(0) L2(0) L1(1) L2(1) ...
      auto layer1 = [&](int it_) {
        mbar_wait(&acc_empty[0], (uint32_t)(it_ & 1) ^ 1u);
        tc_fence_after();
        for (int kb = 0; kb < kb0n; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * kT3Stage, sb = sa + kT3ABytes;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(acc1, make_smem_desc_sw128(sa + k * 32, 16, 1024), make_smem_desc_sw128(sb + k * 32, 16, 1024), id1, (kb | k) ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if (++stage == kT3Stages) { stage = 0; phase ^= 1u; }
        }
        umma_commit(&acc_full[0]);
      };
      int it = 0;
      layer1(0);
      for (int tile = blockIdx.x; tile < a.m_blocks; tile += gridDim.x, ++it) {
        const uint32_t par = (uint32_t)(it & 1);
        // ---- layer 2: A = h1 tile in shared memory
        mbar_wait(&h_ready[0], par);
        mbar_wait(&acc_empty[1], par ^ 1u);
        tc_fence_after();
        for (int kb = 0; kb < kb1n; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = sH + kb * kT3ABytes, sb = smem_base + stage * kT3Stage + kT3ABytes;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(acc2, make_smem_desc_sw128(sa + k * 32, 16, 1024), make_smem_desc_sw128(sb + k * 32, 16, 1024), id2, (kb | k) ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if (++stage == kT3Stages) { stage = 0; phase ^= 1u; }
        }
        umma_commit(&acc_full[1]);
        if (tile + (int)gridDim.x < a.m_blocks) layer1(it + 1);
        // ---- layer 3: A = h2 tile (H k-blocks 0..), B = resident W3
        mbar_wait(&h_ready[1], par);
        mbar_wait(&acc_empty[2], par ^ 1u);
        tc_fence_after();
        for (int kb = 0; kb < kb2n; ++kb) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(acc3, make_smem_desc_sw128(sH + kb * kT3ABytes + k * 32, 16, 1024), make_smem_desc_sw128(sW3 + kb * 8192 + k * 32, 16, 1024),
                      id3, (kb | k) ? 1u : 0u);
        }
        umma_commit(&acc_full[2]);
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ epilogue (8 warps)
    const uint32_t q = warp & 3u, half = (warp - 4u) >> 2;
    const uint32_t lane_addr = (q * 32u) << 16;
    int it = 0;
    for (int tile = blockIdx.x; tile < a.m_blocks; tile += gridDim.x, ++it) {
      const uint32_t par = (uint32_t)(it & 1);
      const int row0 = tile * 128;
      const long long row = (long long)row0 + q * 32 + lane;
      const bool row_ok = row < a.M;
      // ---- h1: every k-block of H is rewritten; the previous tile's stores out of H (h2, h3 staging) must have been read
      mbar_wait(&acc_full[0], par);
      tc_fence_after();
      if (lane == 0) bulk_wait_read0();
      t3_named_bar(1, 256);
      {
        const int ub = half == 0 ? 0 : (kb1n + 1) / 2, ue = half == 0 ? (kb1n + 1) / 2 : kb1n;
        for (int u = ub; u < ue; ++u)
          t3_unit(acc1, lane_addr, u, s_b1, sH + u * kT3ABytes, q, lane, &tmap_h1, row0, row, row_ok, a.bits1, a.N1 / 32, a.M);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&acc_empty[0]); mbar_arrive(&h_ready[0]); }
      // ---- h2 into H k-blocks 0 .. N2/64-1 (layer 2's MMAs have read h1: acc_full[1]; h1's stores must have been read too)
      mbar_wait(&acc_full[1], par);
      tc_fence_after();
      if (lane == 0) bulk_wait_read0();
      t3_named_bar(1, 256);
      {
        const int ub = half == 0 ? 0 : (kb2n + 1) / 2, ue = half == 0 ? (kb2n + 1) / 2 : kb2n;
        for (int u = ub; u < ue; ++u)
          t3_unit(acc2, lane_addr, u, s_b2, sH + u * kT3ABytes, q, lane, &tmap_h2, row0, row, row_ok, a.bits2, a.N2 / 32, a.M);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&acc_empty[1]); mbar_arrive(&h_ready[1]); }
      // ---- h3: staged in H k-block 2 (h1's copy of it was read before the barrier above; h2 lives in k-blocks 0-1)
      mbar_wait(&acc_full[2], par);
      tc_fence_after();
      if (half == 0) {
        // the unit index selects accumulator columns and the output column; the staging block is fixed
        uint32_t r0[32], r1[32];
        tmem_ld_32x32(acc3 + lane_addr, r0);
        tmem_ld_32x32(acc3 + lane_addr + 32, r1);
        tmem_ld_wait();
        uint32_t pk[32], bw[2] = {0u, 0u};
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float a0 = fmaxf(__uint_as_float(r0[j]) + s_b3[j], 0.f), a1 = fmaxf(__uint_as_float(r0[j + 1]) + s_b3[j + 1], 0.f);
          const float c0 = fmaxf(__uint_as_float(r1[j]) + s_b3[32 + j], 0.f), c1 = fmaxf(__uint_as_float(r1[j + 1]) + s_b3[32 + j + 1], 0.f);
          bw[0] |= (a0 > 0.f ? 1u : 0u) << j; bw[0] |= (a1 > 0.f ? 1u : 0u) << (j + 1);
          bw[1] |= (c0 > 0.f ? 1u : 0u) << j; bw[1] |= (c1 > 0.f ? 1u : 0u) << (j + 1);
          pk[j >> 1] = pack_bf16(a0, a1);
          pk[16 + (j >> 1)] = pack_bf16(c0, c1);
        }
        if (a.bits3 && row_ok) { a.bits3[row * (a.N3 / 32)] = bw[0]; a.bits3[row * (a.N3 / 32) + 1] = bw[1]; }
        const uint32_t blk = sH + 2 * kT3ABytes;
        const uint32_t my_row = blk + (q * 32u + lane) * 128u;
        const uint32_t swz = lane & 7u;
#pragma unroll
        for (int p = 0; p < 8; ++p) st_shared_u4(my_row + (((uint32_t)p ^ swz) << 4), pk[4 * p], pk[4 * p + 1], pk[4 * p + 2], pk[4 * p + 3]);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && row0 + (int)q * 32 < a.M) { tma_store_2d(&tmap_h3, blk + q * 4096u, 0, row0 + (int)q * 32); bulk_commit(); }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[2]);
    }
    if (lane == 0) bulk_wait_all();                        // every store of this warp is complete before the CTA exits
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

}  // namespace zoo

using namespace zoo;

typedef CUresult (*PFN_t3_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                       const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                       CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int t3_tmap(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, int box_cols, int box_rows) {
  static PFN_t3_encodeTiled enc = nullptr;
  if (!enc) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
      zoo_set_last_error("cuTensorMapEncodeTiled unavailable", __FILE__, __LINE__);
      return ZOO_ERR_CUDA;
    }
    enc = reinterpret_cast<PFN_t3_encodeTiled>(p);
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = CUDA_SUCCESS;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT) break;
    cudaFree(0);          // a thread without a bound context (e.g. an autograd worker): bind the primary context and retry
  }
  if (r != CUDA_SUCCESS) {
    char msg[200];
    snprintf(msg, sizeof(msg), "mlp3: cuTensorMapEncodeTiled failed (CUresult %d): rows %lld cols %lld ld %lld box %dx%d", (int)r, rows, cols, ld, box_rows, box_cols);
    zoo_set_last_error(msg, __FILE__, __LINE__);
    return ZOO_ERR_CUDA;
  }
  return ZOO_OK;
}

// 1 when zoo_mlp3_fwd_bf16 covers these layer widths.
ZOO_API int zoo_mlp3_supported(int K0, int N1, int N2, int N3) {
  return K0 >= 64 && K0 <= 256 && K0 % 64 == 0 && N1 >= 64 && N1 <= 256 && N1 % 64 == 0 && N2 >= 64 && N2 <= 128 && N2 % 64 == 0 && N3 == 64;
}

// h1 = relu(x W1^T + b1), h2 = relu(h1 W2^T + b2), h3 = relu(h2 W3^T + b3) in one launch.  x [M, K0], W_l [N_l, K_l] (K-major, bf16),
// h_l [M, N_l] bf16 with leading dimensions ld* (multiples of 8); bits_l (optional): ReLU sign bits, N_l / 32 words per row.
ZOO_API int zoo_mlp3_fwd_bf16(const void* x, long long ldx, const void* w1, long long ldw1, const float* b1, const void* w2, long long ldw2,
                              const float* b2, const void* w3, long long ldw3, const float* b3, void* h1, long long ldh1, void* h2,
                              long long ldh2, void* h3, long long ldh3, int M, int K0, int N1, int N2, int N3, unsigned int* bits1,
                              unsigned int* bits2, unsigned int* bits3, void* stream) {
  if (!zoo_mlp3_supported(K0, N1, N2, N3)) { zoo_set_last_error("mlp3: unsupported layer widths", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (!x || !w1 || !w2 || !w3 || !h1 || !h2 || !h3 || M <= 0) { zoo_set_last_error("mlp3: null pointer or empty batch", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  const long long lds[7] = {ldx, ldw1, ldw2, ldw3, ldh1, ldh2, ldh3};
  const void* ptrs[7] = {x, w1, w2, w3, h1, h2, h3};
  for (int i = 0; i < 7; ++i)
    if ((lds[i] & 7) || ((uintptr_t)ptrs[i] & 15)) { zoo_set_last_error("mlp3: operands need 16-byte aligned rows", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  CUtensorMap tx, tw1, tw2, tw3, th1, th2, th3;
  int rc;
  if ((rc = t3_tmap(&tx, x, M, K0, ldx, 64, 128))) return rc;
  if ((rc = t3_tmap(&tw1, w1, N1, K0, ldw1, 64, N1))) return rc;
  if ((rc = t3_tmap(&tw2, w2, N2, N1, ldw2, 64, N2))) return rc;
  if ((rc = t3_tmap(&tw3, w3, N3, N2, ldw3, 64, N3))) return rc;
  if ((rc = t3_tmap(&th1, h1, M, N1, ldh1, 64, 32))) return rc;
  if ((rc = t3_tmap(&th2, h2, M, N2, ldh2, 64, 32))) return rc;
  if ((rc = t3_tmap(&th3, h3, M, N3, ldh3, 64, 32))) return rc;
  Mlp3Args a;
  a.M = M; a.K0 = K0; a.N1 = N1; a.N2 = N2; a.N3 = N3; a.b1 = b1; a.b2 = b2; a.b3 = b3; a.bits1 = bits1; a.bits2 = bits2; a.bits3 = bits3;
  a.m_blocks = (M + 127) / 128;
  static bool attr_set = false;
  if (!attr_set) {
    ZOO_CUDA_CHECK(cudaFuncSetAttribute(mlp3_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kT3Smem));
    attr_set = true;
  }
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  const int grid = a.m_blocks < sms ? a.m_blocks : sms;
  ZOO_CUDA_CHECK(zoo_launch_pdl(mlp3_fwd_kernel, dim3(grid), dim3(kT3Threads), (size_t)kT3Smem, (cudaStream_t)stream, tx, tw1, tw2, tw3, th1, th2, th3, a));
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
