// K8 backward: dQ / dK / dV of blockwise attention on tcgen05 / TMEM / TMA, head dimension 64, probabilities recomputed.
//
// The reference differentiates the dense [batch, heads, L, L] softmax of TransformerLayer.scala:163-181 through BigDL's
// autograd (mm -> softmax -> mm), keeping the L x L probabilities of every layer alive between forward and backward.  Here
// nothing L x L is stored: forward keeps one fp32 log-sum-exp per row (csrc/ops/attention.cu), and backward rebuilds each
// 128 x 128 block of probabilities in TMEM.
//
// One CTA owns a block of 128 KEYS of one (batch, head) and walks the query blocks i that can see it:
//   warp 5  TMA producer : K_j, V_j once; Q_i and dO_i double-buffered
//   warp 4  MMA issuer   : S  = Q_i K_j^T          (TMEM cols [0,128))      dP = dO_i V_j^T   (cols [128,256))
//                          dV += P^T dO_i           (cols [256,320))         dK += dS^T Q_i    (cols [320,384))
//                          dQ_i = dS K_j            (cols [384,448))
//   warps 0-3            : thread t = query row t: p = exp2(s * scale * log2e + mask - lse), ds = p * (dp - delta) * scale,
//                          P and dS written as bf16 into swizzled shared-memory tiles that are MMA operands in BOTH
//                          orientations (K-major A for dQ, MN-major A for dV / dK -- the same bytes, another descriptor);
//                          then dQ_i is read back from TMEM and added to the fp32 dQ buffer (16-byte vector reductions);
//                          at the end thread t = key row t stores dK_j / dV_j.
// delta[row] = sum_d dO * O comes from a small pre-pass.  Every operand tile serves two MMAs: [rows][64] bf16 is K-major for a
// product over d and MN-major for a product over rows.
#include "../common/zoo_common.cuh"

namespace zoo {

constexpr int kAbD = 64, kAbB = 128;                     // head dimension, block edge (queries and keys)
constexpr int kAbTile = kAbB * kAbD * 2;                 // 16 KiB: one [128][64] bf16 tile
constexpr int kAbPS = kAbB * kAbB * 2;                   // 32 KiB: one [128 q][128 keys] bf16 tile (two 64-key halves)
constexpr int kAbSmem = 2 * kAbTile /*K,V*/ + 4 * kAbTile /*Q,dO x2*/ + 2 * kAbPS /*P,dS*/ + 1024;
constexpr int kAbThreads = 192;
constexpr float kAbLog2e = 1.4426950408889634f;

__device__ __forceinline__ float ab_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct AttBwdArgs {
  int B, H, L;
  float scale;
  int causal;
  const float* mask;           // additive [B, L] per key or null
  const float* lse;            // [B, H, L] log2 domain
  const float* delta;          // [B, H, L]
  float* dq;                   // [B, H, L, 64] fp32, zero on entry
  __nv_bfloat16* dk;           // [B, H, L, 64]
  __nv_bfloat16* dv;
};

// delta[row] = sum_d dO[row, d] * O[row, d]   (one warp per row, 2 elements per lane)
__global__ void __launch_bounds__(256)
attention_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o, float* __restrict__ delta, long long rows) {
  const long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float2 a = unpack_bf16(reinterpret_cast<const uint32_t*>(o + row * kAbD)[lane]);
  const float2 b = unpack_bf16(reinterpret_cast<const uint32_t*>(d_o + row * kAbD)[lane]);
  const float s = warp_sum(a.x * b.x + a.y * b.y);
  if (lane == 0) delta[row] = s;
}

__global__ void __launch_bounds__(kAbThreads, 1)
attention_bwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                     const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_do, const AttBwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t kv_bar, qdo_full[2], qdo_empty[2], s_bar, pds_bar, dq_bar;
  __shared__ uint32_t tmem_slot;
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t off_k = 0, off_v = kAbTile, off_qdo = 2 * kAbTile, off_p = 6 * kAbTile, off_ds = 6 * kAbTile + kAbPS;

  const int blocks = (a.L + kAbB - 1) / kAbB;
  const int bh = blockIdx.x / blocks, jb = blockIdx.x - bh * blocks;
  const int b = bh / a.H;
  const int k0 = jb * kAbB;
  const long long row_base = (long long)bh * a.L;
  const int i_begin = a.causal ? jb : 0;                   // query blocks below the diagonal block see none of these keys
  const int n_iter = blocks - i_begin;

  if (warp == 5 && lane == 0) { tma_prefetch_desc(&tmap_q); tma_prefetch_desc(&tmap_k); tma_prefetch_desc(&tmap_v); tma_prefetch_desc(&tmap_do); }
  if (warp == 4 && lane == 0) {
    mbar_init(&kv_bar, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&qdo_full[i], 1); mbar_init(&qdo_empty[i], 1); }
    mbar_init(&s_bar, 1); mbar_init(&dq_bar, 1);
    mbar_init(&pds_bar, 128);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(&tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t t_s = tmem, t_dp = tmem + 128, t_dv = tmem + 256, t_dk = tmem + 320, t_dq = tmem + 384;

  if (warp == 5) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      mbar_arrive_expect_tx(&kv_bar, 2 * kAbTile);
      tma_load_2d(smem_al + off_k, &tmap_k, &kv_bar, 0, (int)(row_base + k0));
      tma_load_2d(smem_al + off_v, &tmap_v, &kv_bar, 0, (int)(row_base + k0));
      for (int it = 0; it < n_iter; ++it) {
        const int s = it & 1;
        mbar_wait(&qdo_empty[s], ((it >> 1) & 1) ^ 1u);
        uint8_t* sq = smem_al + off_qdo + s * 2 * kAbTile;
        mbar_arrive_expect_tx(&qdo_full[s], 2 * kAbTile);
        const int q0 = (i_begin + it) * kAbB;
        tma_load_2d(sq, &tmap_q, &qdo_full[s], 0, (int)(row_base + q0));
        tma_load_2d(sq + kAbTile, &tmap_do, &qdo_full[s], 0, (int)(row_base + q0));
      }
    }
  } else if (warp == 4) {
    // ------------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      const uint32_t id_kk = make_idesc_bf16(128, 128, 0u, 0u);     // S, dP: A and B K-major (product over d)
      const uint32_t id_mm = make_idesc_bf16(128, 64, 1u, 1u);      // dV, dK: A and B MN-major (product over queries)
      const uint32_t id_km = make_idesc_bf16(128, 64, 0u, 1u);      // dQ: A K-major (dS), B MN-major (K_j)
      const uint32_t sk = smem_base + off_k, sv = smem_base + off_v, sp = smem_base + off_p, sds = smem_base + off_ds;
      mbar_wait(&kv_bar, 0);
      tc_fence_after();
      auto issue_scores = [&](int it) {
        const int s = it & 1;
        mbar_wait(&qdo_full[s], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t sq = smem_base + off_qdo + s * 2 * kAbTile, sdo = sq + kAbTile;
#pragma unroll
        for (int k = 0; k < kAbD / 16; ++k)
          umma_bf16(t_s, make_smem_desc_sw128(sq + k * 32, 16, 1024), make_smem_desc_sw128(sk + k * 32, 16, 1024), id_kk, k ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < kAbD / 16; ++k)
          umma_bf16(t_dp, make_smem_desc_sw128(sdo + k * 32, 16, 1024), make_smem_desc_sw128(sv + k * 32, 16, 1024), id_kk, k ? 1u : 0u);
        umma_commit(&s_bar);
      };
      if (n_iter > 0) issue_scores(0);
      for (int it = 0; it < n_iter; ++it) {
        const int s = it & 1;
        const uint32_t sq = smem_base + off_qdo + s * 2 * kAbTile, sdo = sq + kAbTile;
        mbar_wait(&pds_bar, it & 1);                               // P_i, dS_i are in shared memory; S / dP / dQ of TMEM are free
        tc_fence_after();
        // dV += P^T dO_i, dK += dS^T Q_i : reduction over the 128 queries = 8 steps of 16 rows (2048 B); the two 64-key halves of
        // P / dS are 16 KiB apart (leading byte offset)
#pragma unroll
        for (int k = 0; k < kAbB / 16; ++k) {
          umma_bf16(t_dv, make_smem_desc_sw128(sp + k * 2048, 16384, 1024), make_smem_desc_sw128(sdo + k * 2048, 8192, 1024), id_mm, (it | k) ? 1u : 0u);
          umma_bf16(t_dk, make_smem_desc_sw128(sds + k * 2048, 16384, 1024), make_smem_desc_sw128(sq + k * 2048, 8192, 1024), id_mm, (it | k) ? 1u : 0u);
        }
        // dQ_i = dS K_j : reduction over the 128 keys = 2 halves x 4 steps of 16 keys (32 B inside the swizzled row of dS)
#pragma unroll
        for (int kb = 0; kb < 2; ++kb)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(t_dq, make_smem_desc_sw128(sds + kb * 16384 + k * 32, 16, 1024),
                      make_smem_desc_sw128(sk + kb * 8192 + k * 2048, 8192, 1024), id_km, (kb | k) ? 1u : 0u);
        umma_commit(&qdo_empty[s]);                                // Q_i / dO_i may be overwritten
        umma_commit(&dq_bar);                                      // dQ_i (and everything before it) is complete
        if (it + 1 < n_iter) issue_scores(it + 1);                 // overlaps the dQ read-back of block i
      }
    }
  } else {
    // ------------------------------------------------------------------ probabilities, dS, dQ read-back (thread = query row)
    const int t = threadIdx.x;
    const uint32_t lane_addr = (warp * 32u) << 16;
    const uint32_t p_row = smem_base + off_p + (uint32_t)t * 128u, ds_row = smem_base + off_ds + (uint32_t)t * 128u;
    const uint32_t swz = (uint32_t)t & 7u;
    const float sc = a.scale * kAbLog2e;
    const float* mrow = a.mask ? a.mask + (long long)b * a.L : nullptr;
    for (int it = 0; it < n_iter; ++it) {
      const int q = (i_begin + it) * kAbB + t;
      const bool q_ok = q < a.L;
      const float lse = q_ok ? a.lse[row_base + q] : INFINITY;
      const float dl = q_ok ? a.delta[row_base + q] : 0.f;
      const int qb0 = (i_begin + it) * kAbB;                 // a block pair is FULL when nothing in it needs masking
      const bool full = !mrow && k0 + kAbB <= a.L && qb0 + kAbB <= a.L && (!a.causal || k0 + kAbB - 1 <= qb0) && sc > 0.f;
      mbar_wait(&s_bar, it & 1);
      tc_fence_after();
      // (P / dS of the previous block were MMA operands until its dQ commit, which was waited for before the previous read-back)
#pragma unroll 1
      for (int c = 0; c < kAbB / 32; ++c) {
        uint32_t rs[32], rd[32];
        tmem_ld_32x32(t_s + lane_addr + c * 32, rs);
        tmem_ld_32x32(t_dp + lane_addr + c * 32, rd);
        tmem_ld_wait();
        uint32_t pk[16], dk[16];
        if (full) {                                          // no per-element predicate (see the forward kernel)
          const float neg_lse = -lse, ndl = -dl * a.scale;
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float p0 = ab_ex2(fmaf(__uint_as_float(rs[i]), sc, neg_lse)), p1 = ab_ex2(fmaf(__uint_as_float(rs[i + 1]), sc, neg_lse));
            const float d0 = p0 * fmaf(__uint_as_float(rd[i]), a.scale, ndl), d1 = p1 * fmaf(__uint_as_float(rd[i + 1]), a.scale, ndl);
            pk[i >> 1] = pack_bf16(p0, p1);
            dk[i >> 1] = pack_bf16(d0, d1);
          }
        } else {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p2[2], d2[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int key = k0 + c * 32 + i + u;
            float s = __uint_as_float(rs[i + u]) * sc;
            if (mrow && key < a.L) s += mrow[key] * kAbLog2e;
            const bool dead = !q_ok || key >= a.L || (a.causal && key > q);
            p2[u] = dead ? 0.f : exp2f(s - lse);
            d2[u] = p2[u] * (__uint_as_float(rd[i + u]) - dl) * a.scale;
          }
          pk[i >> 1] = pack_bf16(p2[0], p2[1]);
          dk[i >> 1] = pack_bf16(d2[0], d2[1]);
        }
        }
        const uint32_t kb_off = (uint32_t)(c >> 1) * 16384u;
#pragma unroll
        for (int piece = 0; piece < 4; ++piece) {
          const uint32_t pi = (uint32_t)((c & 1) * 4 + piece);
          st_shared_u4(p_row + kb_off + ((pi ^ swz) << 4), pk[4 * piece], pk[4 * piece + 1], pk[4 * piece + 2], pk[4 * piece + 3]);
          st_shared_u4(ds_row + kb_off + ((pi ^ swz) << 4), dk[4 * piece], dk[4 * piece + 1], dk[4 * piece + 2], dk[4 * piece + 3]);
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(&pds_bar);
      // dQ_i back from TMEM into the fp32 accumulation buffer (other key blocks add to the same rows)
      mbar_wait(&dq_bar, it & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < kAbD / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(t_dq + lane_addr + c * 32, r);
        tmem_ld_wait();
        if (q_ok) {
          float* dst = a.dq + (row_base + q) * kAbD + c * 32;
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            red_add_f32x4(dst + i, __uint_as_float(r[i]), __uint_as_float(r[i + 1]), __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
        }
      }
      tc_fence_before();
    }
    // dK_j, dV_j: thread t = key row t.  The last dq_bar wait covered every MMA of this CTA.  (tcgen05.ld is warp-collective:
    // every lane of the warp executes it, only the rows inside the sequence are stored.)
    const int key = k0 + t;
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      __nv_bfloat16* dst = (which ? a.dk : a.dv) + (row_base + key) * kAbD;
#pragma unroll
      for (int c = 0; c < kAbD / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32((which ? t_dk : t_dv) + lane_addr + c * 32, r);
        tmem_ld_wait();
        if (key < a.L) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 v;
            v.x = pack_bf16(__uint_as_float(r[i]), __uint_as_float(r[i + 1])); v.y = pack_bf16(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
            v.z = pack_bf16(__uint_as_float(r[i + 4]), __uint_as_float(r[i + 5])); v.w = pack_bf16(__uint_as_float(r[i + 6]), __uint_as_float(r[i + 7]));
            *reinterpret_cast<uint4*>(dst + c * 32 + i) = v;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

}  // namespace zoo

using namespace zoo;

typedef CUresult (*PFN_ab_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                       const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                       CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int ab_tmap(CUtensorMap* m, const void* base, long long rows) {
  static PFN_ab_encodeTiled enc = nullptr;
  if (!enc) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
      zoo_set_last_error("cuTensorMapEncodeTiled unavailable", __FILE__, __LINE__);
      return ZOO_ERR_CUDA;
    }
    enc = reinterpret_cast<PFN_ab_encodeTiled>(p);
  }
  cuuint64_t dims[2] = {(cuuint64_t)kAbD, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)kAbD * 2};
  cuuint32_t box[2] = {(cuuint32_t)kAbD, (cuuint32_t)kAbB};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = CUDA_SUCCESS;
  for (int attempt = 0; attempt < 2; ++attempt) {
    r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT) break;
    cudaFree(0);
  }
  if (r != CUDA_SUCCESS) {
    zoo_set_last_error("attention_bwd: cuTensorMapEncodeTiled failed (operands must be contiguous [B,H,L,64] bf16, 16-byte aligned)", __FILE__, __LINE__);
    return ZOO_ERR_CUDA;
  }
  return ZOO_OK;
}

// Gradients of out = softmax(scale * q k^T + mask (+ causal)) v.
//   q, k, v, o, d_o : contiguous bf16 [B,H,L,64] (o, lse from zoo_attention_fwd_lse_bf16)
//   delta           : fp32 [B,H,L] scratch (written here)
//   dq              : fp32 [B,H,L,64], ZERO on entry (key blocks accumulate into it);  dk, dv : bf16 [B,H,L,64]
ZOO_API int zoo_attention_bwd_bf16(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                                   float* delta, const float* mask, float* dq, void* dk, void* dv, int B, int H, int L, int D,
                                   float scale, int causal, void* stream) {
  if (!q || !k || !v || !o || !d_o || !lse || !delta || !dq || !dk || !dv || B <= 0 || H <= 0 || L <= 0) {
    zoo_set_last_error("attention_bwd: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (D != kAbD) { zoo_set_last_error("attention_bwd: the kernel is specialised for head dimension 64", __FILE__, __LINE__); return ZOO_ERR_ARG; }
  const long long rows = (long long)B * H * L;
  if (rows > 0x7fffffffLL - 256) { zoo_set_last_error("attention_bwd: B*H*L too large", __FILE__, __LINE__); return ZOO_ERR_TOO_LARGE; }
  cudaStream_t s = (cudaStream_t)stream;
  attention_delta_kernel<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, s>>>((const __nv_bfloat16*)o, (const __nv_bfloat16*)d_o, delta, rows);
  ZOO_CUDA_CHECK(cudaGetLastError());
  CUtensorMap tq, tk, tv, tdo;
  int rc;
  if ((rc = ab_tmap(&tq, q, rows))) return rc;
  if ((rc = ab_tmap(&tk, k, rows))) return rc;
  if ((rc = ab_tmap(&tv, v, rows))) return rc;
  if ((rc = ab_tmap(&tdo, d_o, rows))) return rc;
  AttBwdArgs a{};
  a.B = B; a.H = H; a.L = L; a.scale = scale; a.causal = causal; a.mask = mask; a.lse = lse; a.delta = delta;
  a.dq = dq; a.dk = (__nv_bfloat16*)dk; a.dv = (__nv_bfloat16*)dv;
  static bool attr_set = false;
  if (!attr_set) {
    ZOO_CUDA_CHECK(cudaFuncSetAttribute(attention_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAbSmem));
    attr_set = true;
  }
  const int blocks = (L + kAbB - 1) / kAbB;
  attention_bwd_kernel<<<(unsigned)((long long)B * H * blocks), kAbThreads, kAbSmem, s>>>(tq, tk, tv, tdo, a);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
