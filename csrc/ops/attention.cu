// K8: blockwise (flash-style) attention forward on tcgen05 / TMEM / TMA, head dimension 64, bf16 in / bf16 out.
//
// Replaces the dense O(L^2) attention of the reference's TransformerLayer / BERT
// (zoo/.../pipeline/api/keras/layers/TransformerLayer.scala:163-181: mm(q,k) -> mask -> softmax -> mm(w,v), scores
// materialised as [batch, heads, L, L]).  One CTA owns 128 query rows of one (batch, head) and walks the keys in blocks of 128:
//
//   warp 5  TMA producer : Q tile once; K_j [128 keys x 64] (K-major) and V_j (MN-major, two 64-key boxes), double-buffered
//   warp 4  MMA issuer   : S = Q K_j^T  (128 x 128, fp32 in TMEM columns [0,128)), then O_j = P_j V_j (128 x 64, columns [128,192))
//   warps 0-3 softmax    : thread t owns query row t (= TMEM lane t): two sweeps over S with tcgen05.ld (row max, then
//                          p = exp2((s - m) * log2e)), P_j written as bf16 into a swizzled K-major shared-memory tile that is the
//                          A operand of the second MMA; the running output row lives in 64 fp32 registers and is rescaled by
//                          alpha = exp2(m_old - m_new) when the maximum moves (online softmax), O_j is added from TMEM.
//
// The L x L score matrix never leaves the SM.  Key padding masks are an additive fp32 vector per (batch, key); `causal` skips
// the key blocks above the diagonal.  STATUS: written at the end of round 1 after the GPU budget was spent -- compiled for
// sm_100a, not yet run on hardware; `zoo.ops.attention` only takes this path when ZOO_ATTN_KERNEL=1 (tests/test_gpu_attention.py
// is skipped otherwise) and torch SDPA remains the default.
#include "../common/zoo_common.cuh"

namespace zoo {

constexpr int kAttD = 64;                 // head dimension (one 128-byte swizzle row of bf16)
constexpr int kAttBM = 128;               // query rows per CTA (UMMA M)
constexpr int kAttBN = 128;               // keys per block (UMMA N of the score MMA)
constexpr int kAttQBytes = kAttBM * kAttD * 2;          // 16 KiB
constexpr int kAttKBytes = kAttBN * kAttD * 2;          // 16 KiB
constexpr int kAttVBytes = kAttBN * kAttD * 2;          // 16 KiB (two MN-major boxes of 64 keys)
constexpr int kAttPBytes = kAttBM * kAttBN * 2;         // 32 KiB (two k-blocks of 64 keys)
constexpr int kAttTiles = kAttQBytes + 2 * (kAttKBytes + kAttVBytes) + kAttPBytes;      // 112 KiB of operand tiles
// Barriers live behind the tiles in the SAME dynamic allocation and the kernel has no static shared memory, so the 1024-byte
// aligned tiles start at the CTA's shared-memory base and two CTAs (2 x 112.1 KiB + the per-CTA reserve) fit one SM: while one
// CTA's softmax warps work, the other CTA's MMAs run.  (With a padded allocation only one CTA fit: ncu showed 9 % active warps.)
constexpr int kAttSmem = kAttTiles + 128;
constexpr int kAttThreads = 192;
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct AttArgs {
  int B, H, L;
  float scale;                 // multiplies q.k (1/sqrt(d) or 1)
  int causal;
  const float* mask;           // additive, [B, L] per key, or null
  __nv_bfloat16* out;          // [B, H, L, 64]
  float* lse;                  // optional [B, H, L]: log2-domain log-sum-exp of every row's scaled, masked scores (for backward)
};

__global__ void __launch_bounds__(kAttThreads, 2)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                     const __grid_constant__ CUtensorMap tmap_v, const AttArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + kAttTiles);
  uint64_t& q_bar = bars[0];
  uint64_t* kv_full = bars + 1;
  uint64_t* kv_empty = bars + 3;
  uint64_t* s_bar = bars + 10;                              // one per score buffer (the issuer runs one block ahead)
  uint64_t& p_bar = bars[6];
  uint64_t& o_bar = bars[7];
  uint64_t& oread_bar = bars[8];
  uint32_t& tmem_slot = *reinterpret_cast<uint32_t*>(bars + 12);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t smem_base = smem_u32(smem_raw);
  uint8_t* smem_al = smem_raw;
  if (smem_base & 1023u) { if (threadIdx.x == 0) printf("zoo attention: shared memory base is not 1024-byte aligned\n"); __trap(); }
  const uint32_t off_q = 0, off_kv = kAttQBytes, off_p = kAttQBytes + 2 * (kAttKBytes + kAttVBytes);

  const int q_tiles = (a.L + kAttBM - 1) / kAttBM;
  const int bh = blockIdx.x / q_tiles, qt = blockIdx.x - bh * q_tiles;
  const int b = bh / a.H;
  const int q0 = qt * kAttBM;
  const long long row_base = (long long)bh * a.L;          // first row of this (batch, head) in the [B*H*L, 64] views
  int n_blocks = (a.L + kAttBN - 1) / kAttBN;
  if (a.causal) n_blocks = min(n_blocks, (q0 + kAttBM + kAttBN - 1) / kAttBN);

  if (warp == 5 && lane == 0) { tma_prefetch_desc(&tmap_q); tma_prefetch_desc(&tmap_k); tma_prefetch_desc(&tmap_v); }
  if (warp == 4 && lane == 0) {
    mbar_init(&q_bar, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1); }
    mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); mbar_init(&o_bar, 1);
    mbar_init(&p_bar, 128);                                  // every softmax thread arrives once per key block
    mbar_init(&oread_bar, 128);                              // ... and once per output read-back
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(&tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  // Two score buffers of 128 columns: S_{j+1} = Q K_{j+1}^T is issued while the softmax warps still work on S_j.  O_j = P_j V_j is
  // written over the first 64 columns of S_j's own buffer once S_j has been consumed (no third region: 256 columns per CTA, two
  // CTAs per SM).
  const uint32_t tmem_sbuf[2] = {tmem, tmem + kAttBN};

  if (warp == 5) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      mbar_arrive_expect_tx(&q_bar, kAttQBytes);
      tma_load_2d(smem_al + off_q, &tmap_q, &q_bar, 0, (int)(row_base + q0));
      for (int j = 0; j < n_blocks; ++j) {
        const int s = j & 1;
        mbar_wait(&kv_empty[s], ((j >> 1) & 1) ^ 1u);
        uint8_t* sk = smem_al + off_kv + s * (kAttKBytes + kAttVBytes);
        uint8_t* sv = sk + kAttKBytes;
        mbar_arrive_expect_tx(&kv_full[s], kAttKBytes + kAttVBytes);
        const int k0 = (int)(row_base + (long long)j * kAttBN);
        tma_load_2d(sk, &tmap_k, &kv_full[s], 0, k0);
        tma_load_2d(sv, &tmap_v, &kv_full[s], 0, k0);                 // keys k0 .. k0+63   (MN-major box: 64 d x 64 keys)
        tma_load_2d(sv + 8192, &tmap_v, &kv_full[s], 0, k0 + 64);     // keys k0+64 .. k0+127
      }
    }
  } else if (warp == 4) {
    // ------------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      const uint32_t idesc_s = make_idesc_bf16(kAttBM, kAttBN, 0u, 0u);      // A = Q (K-major), B = K_j (K-major)
      const uint32_t idesc_o = make_idesc_bf16(kAttBM, kAttD, 0u, 1u);       // A = P (K-major), B = V_j (MN-major)
      mbar_wait(&q_bar, 0);
      tc_fence_after();
      const uint32_t sq = smem_base + off_q, sp = smem_base + off_p;
      auto issue_scores = [&](int j) {
        const int s = j & 1;
        const uint32_t sk = smem_base + off_kv + s * (kAttKBytes + kAttVBytes);
        mbar_wait(&kv_full[s], (j >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < kAttD / 16; ++k)
          umma_bf16(tmem_sbuf[s], make_smem_desc_sw128(sq + k * 32, 16, 1024), make_smem_desc_sw128(sk + k * 32, 16, 1024), idesc_s, k ? 1u : 0u);
        umma_commit(&s_bar[s]);                                // scores of block j are in TMEM
      };
      issue_scores(0);
      for (int j = 0; j < n_blocks; ++j) {
        const int s = j & 1;
        const uint32_t sv = smem_base + off_kv + s * (kAttKBytes + kAttVBytes) + kAttKBytes;
        if (j + 1 < n_blocks) {
          // buffer (j+1)&1 held S_{j-1}, then O_{j-1}: free once every softmax thread has read O_{j-1} back
          if (j > 0) { mbar_wait(&oread_bar, (j - 1) & 1); tc_fence_after(); }
          issue_scores(j + 1);
        }
        mbar_wait(&p_bar, j & 1);                              // P_j is in shared memory (and S_j has been read)
        tc_fence_after();
#pragma unroll
        for (int kb = 0; kb < 2; ++kb)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem_sbuf[s], make_smem_desc_sw128(sp + kb * 16384 + k * 32, 16, 1024),
                      make_smem_desc_sw128(sv + kb * 8192 + k * 2048, 8192, 1024), idesc_o, (kb | k) ? 1u : 0u);
        umma_commit(&kv_empty[s]);                             // K_j / V_j may be overwritten
        umma_commit(&o_bar);                                   // O_j is in TMEM (first 64 columns of buffer s)
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + output (128 threads, thread = query row)
    const int t = threadIdx.x;                                 // 0..127 = TMEM lane
    const int q = q0 + t;
    const uint32_t lane_addr = (warp * 32u) << 16;
    const uint32_t p_row = smem_base + off_p + (uint32_t)t * 128u;
    const uint32_t swz = (uint32_t)t & 7u;
    const float sc = a.scale * kLog2e;
    const float* mrow = a.mask ? a.mask + (long long)b * a.L : nullptr;
    float o[kAttD];
#pragma unroll
    for (int i = 0; i < kAttD; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    // O_j = P_j V_j sits in the first 64 columns of score buffer j & 1 once o_bar completes for block j
    auto add_output = [&](int j) {
      mbar_wait(&o_bar, j & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < kAttD / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_sbuf[j & 1] + lane_addr + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c * 32 + i] += __uint_as_float(r[i]);
      }
      tc_fence_before();
      mbar_arrive(&oread_bar);                                 // the issuer may overwrite this buffer with S_{j+2}
    };
    for (int j = 0; j < n_blocks; ++j) {
      const int key0 = j * kAttBN;
      const uint32_t tmem_s = tmem_sbuf[j & 1];
      mbar_wait(&s_bar[j & 1], (j >> 1) & 1);
      tc_fence_after();
      // A key block is FULL when every key is in range, no key is above the diagonal for any row of this CTA and there is no
      // additive mask: then no per-element predicate is needed.  (ncu on the general form: 2,760 warp instructions per tile,
      // most of them ISETP / FSEL / VIADD of the masking logic; the full form needs ~700.)
      const bool full = !mrow && key0 + kAttBN <= a.L && (!a.causal || key0 + kAttBN - 1 <= q0) && sc > 0.f;
      float lsum = 0.f;
      float m_new, alpha, m_use;
      // ---- sweep 1: row maximum (log2 domain)
      if (full) {                                              // raw scores (scale > 0 keeps the order), two loads in flight
        float raw = -INFINITY;
#pragma unroll 1
        for (int c = 0; c < kAttBN / 64; ++c) {
          uint32_t r0[32], r1[32];
          tmem_ld_32x32(tmem_s + lane_addr + c * 64, r0);
          tmem_ld_32x32(tmem_s + lane_addr + c * 64 + 32, r1);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) raw = fmaxf(raw, fmaxf(__uint_as_float(r0[i]), __uint_as_float(r1[i])));
        }
        m_new = fmaxf(m, raw * sc);
        alpha = (m == -INFINITY) ? 0.f : ex2_approx(m - m_new);
        m_use = m_new;                                         // finite: a full block has live keys
      } else {
        float mx = m;
#pragma unroll 1
        for (int c = 0; c < kAttBN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_s + lane_addr + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int key = key0 + c * 32 + i;
            float sv = __uint_as_float(r[i]) * sc;
            if (mrow && key < a.L) sv += mrow[key] * kLog2e;
            if (key >= a.L || (a.causal && key > q)) sv = -INFINITY;
            mx = fmaxf(mx, sv);
          }
        }
        m_new = mx;
        alpha = (m == -INFINITY) ? 0.f : exp2f(m - m_new);
        m_use = (m_new == -INFINITY) ? 0.f : m_new;            // a fully masked row keeps p = 0 without NaNs
      }
      // ---- the previous block's output: its MMA ran under sweep 1; P may be rewritten once it has completed
      if (j > 0) add_output(j - 1);
      // ---- sweep 2: probabilities -> bf16 -> swizzled K-major tile (A operand of the second MMA)
      if (full) {
        const float neg_m = -m_use;
#pragma unroll 1
        for (int c = 0; c < kAttBN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_s + lane_addr + c * 32, r);
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float p0 = ex2_approx(fmaf(__uint_as_float(r[i]), sc, neg_m)), p1 = ex2_approx(fmaf(__uint_as_float(r[i + 1]), sc, neg_m));
            lsum += p0 + p1;
            pk[i >> 1] = pack_bf16(p0, p1);
          }
          const uint32_t kb_off = (uint32_t)(c >> 1) * 16384u;
#pragma unroll
          for (int piece = 0; piece < 4; ++piece) {
            const uint32_t pi = (uint32_t)((c & 1) * 4 + piece);
            st_shared_u4(p_row + kb_off + ((pi ^ swz) << 4), pk[4 * piece], pk[4 * piece + 1], pk[4 * piece + 2], pk[4 * piece + 3]);
          }
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < kAttBN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_s + lane_addr + c * 32, r);
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            float p2[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
              const int key = key0 + c * 32 + i + u;
              float sv = __uint_as_float(r[i + u]) * sc;
              if (mrow && key < a.L) sv += mrow[key] * kLog2e;
              const bool dead = key >= a.L || (a.causal && key > q);
              p2[u] = dead ? 0.f : exp2f(sv - m_use);
            }
            lsum += p2[0] + p2[1];
            pk[i >> 1] = pack_bf16(p2[0], p2[1]);
          }
          // chunk c covers keys [32c, 32c+32) = 64 bytes of row t inside k-block (c >> 1): 16-byte pieces 4*(c&1) .. +3
          const uint32_t kb_off = (uint32_t)(c >> 1) * 16384u;
#pragma unroll
          for (int piece = 0; piece < 4; ++piece) {
            const uint32_t pi = (uint32_t)((c & 1) * 4 + piece);
            st_shared_u4(p_row + kb_off + ((pi ^ swz) << 4), pk[4 * piece], pk[4 * piece + 1], pk[4 * piece + 2], pk[4 * piece + 3]);
          }
        }
      }
      l = l * alpha + lsum;
      m = m_new;
      if (alpha != 1.f) {                                      // the maximum moved (rare after the first blocks); o includes O_{j-1}
#pragma unroll
        for (int i = 0; i < kAttD; ++i) o[i] *= alpha;
      }
      tc_fence_before();                                       // our tcgen05.ld of S_j are done before the issuer overwrites S
      fence_proxy_async_smem();                                // P_j visible to the tensor core (async proxy)
      mbar_arrive(&p_bar);
    }
    add_output(n_blocks - 1);
    if (q < a.L) {
      if (a.lse) a.lse[row_base + q] = l > 0.f ? m + log2f(l) : INFINITY;      // a row with no live key: p = exp2(s - inf) = 0 in backward
      const float inv = l > 0.f ? 1.f / l : 0.f;
      __nv_bfloat16* orow = a.out + (row_base + q) * kAttD;
#pragma unroll
      for (int i = 0; i < kAttD; i += 8) {
        uint4 v;
        v.x = pack_bf16(o[i] * inv, o[i + 1] * inv); v.y = pack_bf16(o[i + 2] * inv, o[i + 3] * inv);
        v.z = pack_bf16(o[i + 4] * inv, o[i + 5] * inv); v.w = pack_bf16(o[i + 6] * inv, o[i + 7] * inv);
        *reinterpret_cast<uint4*>(orow + i) = v;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

}  // namespace zoo

using namespace zoo;

typedef CUresult (*PFN_att_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int att_tmap(CUtensorMap* m, const void* base, long long rows, int box_rows) {
  static PFN_att_encodeTiled enc = nullptr;
  if (!enc) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
      zoo_set_last_error("cuTensorMapEncodeTiled unavailable", __FILE__, __LINE__);
      return ZOO_ERR_CUDA;
    }
    enc = reinterpret_cast<PFN_att_encodeTiled>(p);
  }
  cuuint64_t dims[2] = {(cuuint64_t)kAttD, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)kAttD * 2};
  cuuint32_t box[2] = {(cuuint32_t)kAttD, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = CUDA_SUCCESS;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT) break;
    cudaFree(0);       // a thread without a bound context (autograd worker): bind the primary context and encode again
  }
  if (r != CUDA_SUCCESS) {
    zoo_set_last_error("attention: cuTensorMapEncodeTiled failed (q/k/v must be contiguous [B,H,L,64] bf16, 16-byte aligned)", __FILE__, __LINE__);
    return ZOO_ERR_CUDA;
  }
  return ZOO_OK;
}

// out[B,H,L,64] = softmax(scale * q k^T + mask[b, key] (+ causal)) v ;  q, k, v, out: contiguous bf16 [B,H,L,64];
// mask: additive fp32 [B, L] or null.
static int attention_fwd_launch(const void* q, const void* k, const void* v, void* out, float* lse, const float* mask, int B, int H, int L,
                                int D, float scale, int causal, void* stream) {
  if (!q || !k || !v || !out || B <= 0 || H <= 0 || L <= 0) { zoo_set_last_error("attention: bad arguments", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  if (D != kAttD) { zoo_set_last_error("attention: the kernel is specialised for head dimension 64", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  const long long rows = (long long)B * H * L;
  if (rows > 0x7fffffffLL - 256) { zoo_set_last_error("attention: B*H*L too large", __FILE__, __LINE__); return ZOO_ERR_TOO_LARGE; }
  CUtensorMap tq, tk, tv;
  int rc;
  if ((rc = att_tmap(&tq, q, rows, kAttBM))) return rc;
  if ((rc = att_tmap(&tk, k, rows, kAttBN))) return rc;
  if ((rc = att_tmap(&tv, v, rows, 64))) return rc;
  AttArgs a{};
  a.B = B; a.H = H; a.L = L; a.scale = scale; a.causal = causal; a.mask = mask; a.out = (__nv_bfloat16*)out; a.lse = lse;
  static bool attr_set = false;
  if (!attr_set) {
    ZOO_CUDA_CHECK(cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttSmem));
    attr_set = true;
  }
  const int q_tiles = (L + kAttBM - 1) / kAttBM;
  attention_fwd_kernel<<<(unsigned)((long long)B * H * q_tiles), kAttThreads, kAttSmem, (cudaStream_t)stream>>>(tq, tk, tv, a);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_attention_fwd_bf16(const void* q, const void* k, const void* v, void* out, const float* mask, int B, int H, int L, int D,
                                   float scale, int causal, void* stream) {
  return attention_fwd_launch(q, k, v, out, nullptr, mask, B, H, L, D, scale, causal, stream);
}

// Forward that also writes lse[B,H,L] (fp32, log2 domain), which zoo_attention_bwd_bf16 needs to recompute the probabilities.
ZOO_API int zoo_attention_fwd_lse_bf16(const void* q, const void* k, const void* v, void* out, float* lse, const float* mask, int B, int H,
                                       int L, int D, float scale, int causal, void* stream) {
  if (!lse) { zoo_set_last_error("attention: lse buffer missing", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  return attention_fwd_launch(q, k, v, out, lse, mask, B, H, L, D, scale, causal, stream);
}
