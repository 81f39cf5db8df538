// K1 + C1: gradient exchange fused with the optimizer over NVLink peer memory, the symmetric
// heap it runs on, and the flag barrier.
//
// The reference's distributed step is BigDL `AllReduceParameter` over the Spark BlockManager:
// gradient slices are put as blocks, task n fetches slice n from every executor, sums it, runs
// OptimMethod.optimize on that slice and publishes the weight slice, and every task fetches all
// weight slices at the start of the next iteration -- two Spark jobs per iteration
// (zoo/.../pipeline/api/keras/models/Topology.scala:1128-1206, docs/docs/wp-bigdl.md:140-158).
// The same owner-computes schedule is ONE kernel here:
//
//   phase A  flag handshake: "my gradients for this range are complete" (st.release.sys into every
//            peer's flag array, ld.acquire.sys spin on the local one) -- replaces job boundary 1
//   reduce   rank r owns shard r of the range; it LOADS that shard of the gradient arena from all
//            `world` GPUs through peer pointers (NVLink, 16-byte loads, all peers in flight at
//            once) and sums in fp32                                   -- the reduce-scatter
//   update   optimizer math on the fp32 master shard and its local m / v        -- zoo_opt_update
//   publish  STORES the new fp32 weights (and the bf16 shadow) into every GPU's parameter arena
//            through peer pointers                                    -- the all-gather
//   phase B  second handshake ("my stores have landed, I am done reading your gradients"), then
//            every rank zeroes its own gradient range for the next iteration.
//
// No NCCL call, no staging buffer, no separate optimizer pass.  Each CTA only ever waits for the
// CTA with the same index on the other GPUs, so the kernel needs no grid-wide synchronisation.
//
// Forward-progress assumption (a plain launch, not a cooperative one): CTA b of this GPU waits for CTA b of every peer, so every
// GPU must eventually RUN its CTA b.  That holds when the grid is co-resident (the launcher caps it at 592 = 4 CTAs x 148 SMs,
// the kernel's occupancy at 256 threads and no shared memory) or when CTAs are dispatched in index order on every GPU and earlier
// CTAs finish without waiting on later ones -- both true for the step as it runs today, where K1 is alone on its stream position
// (it starts after backward).  Running it BESIDE other kernels that hold SMs for long would need the grid clamped to what is
// co-resident next to them; until then a wait that exceeds its clock budget (~10 s) traps with the rank / CTA in
// zoo_last_error instead of hanging.
//
// `emulate` != 0 runs all ranks inside one cooperative launch on ONE GPU (blockIdx.y = rank) --
// used by the single-GPU tests; the real path is one process per GPU with IPC-mapped arenas.
#include "../common/zoo_common.cuh"
#include "../ops/optim_math.cuh"
#include <string.h>

namespace zoo {

constexpr int kMaxRanks = 8;
constexpr int kRsThreads = 256;
constexpr int kFlagStride = 1024; // flag slots per (phase, rank): fixed so launches with different grids never alias; also the CTA cap

struct RsArgs {
  float* g[kMaxRanks];               // gradient arenas (index = rank; peers via IPC)
  float* p[kMaxRanks];               // parameter arenas
  __nv_bfloat16* p16[kMaxRanks];     // bf16 shadows (all null or all set)
  float* m[kMaxRanks];               // optimizer state; only [rank] is touched (emulate: all)
  float* v[kMaxRanks];
  uint32_t* flags[kMaxRanks];        // per-rank flag array: [2 phases][world][kFlagStride] uint32
  uint32_t* epoch[kMaxRanks];        // per-rank local epoch counters [grid]
  const float* hyper[kMaxRanks];     // device hyper-parameter vectors (same values on every rank)
  long long start, end;              // arena range of this launch; (end - start) % (4 * world) == 0
  int world, rank, emulate, opt, zero_grad;
  int replicated;                    // 1: every rank reduces and updates the WHOLE range locally (no publish stores)
  // NVLS: multicast addresses of the local gradient / parameter / shadow arenas (null = peer loads and stores).  One
  // multimem.ld_reduce returns the sum of every GPU's gradients (reduced inside the NVSwitch), one multimem.st writes all copies.
  const float* g_mc;
  float* p_mc;
  __nv_bfloat16* p16_mc;
  int clip;                          // 1: clamp the summed gradient to [clip_lo, clip_hi] (bounds already divided by the gradient scale)
  float clip_lo, clip_hi;
};

__device__ __forceinline__ float4 mm_ld_reduce_f32x4(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ void mm_st_f32x4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void mm_st_bf16x4(__nv_bfloat16* mc, const uint2& v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1,%2};" :: "l"(mc), "r"(v.x), "r"(v.y) : "memory");
}

__device__ __forceinline__ float4 ld_f4(const float* p) { return *reinterpret_cast<const float4*>(p); }

__global__ void __launch_bounds__(kRsThreads)
rs_opt_ag_kernel(const RsArgs a) {
  __shared__ uint32_t s_ep;
  const int rank = a.emulate ? (int)blockIdx.y : a.rank;
  const int W = a.world, G = (int)gridDim.x, b = (int)blockIdx.x;
  if (threadIdx.x == 0) s_ep = a.epoch[rank][b] + 1u;
  __syncthreads();
  const uint32_t ep = s_ep;

  // ---- phase A: every rank's gradients for [start, end) are final ------------------------
  if ((int)threadIdx.x < W) {
    const int peer = threadIdx.x;
    st_release_sys(a.flags[peer] + ((0 * W + rank) * kFlagStride + b), ep);
    const uint32_t* mine = a.flags[rank] + ((0 * W + peer) * kFlagStride + b);
    const long long t0 = clock64();
    while (ld_acquire_sys(mine) < ep) {
      if (clock64() - t0 > 20000000000LL) { printf("zoo rs_opt_ag: phase A timeout rank %d cta %d peer %d\n", rank, b, peer); __trap(); }
    }
  }
  __syncthreads();

  // ---- reduce-scatter + optimizer + all-gather on the owned shard --------------------------
  const long long units = (a.end - a.start) >> 2;       // float4 units in the range
  const OptHyper h = load_hyper(a.hyper[rank], a.opt);
  float* m_loc = a.m[rank];
  float* v_loc = a.v[rank];
  // Small ranges (the dense part of a recommender is a few hundred KB): the two extra NVLink round trips of the
  // owner-computes schedule cost more than the math, so every rank reads all `world` gradient copies of the whole
  // range, sums them in the same order and applies the same update to its own replica -- bitwise identical
  // parameters everywhere, no publish stores, and phase B only has to say "done reading".
  const long long shard_units = a.replicated ? units : units / W;
  const long long u0 = a.replicated ? 0 : (long long)rank * shard_units;    // first unit of my shard
  for (long long u = u0 + b * (long long)kRsThreads + threadIdx.x; u < u0 + shard_units; u += (long long)G * kRsThreads) {
    const long long off = a.start + (u << 2);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.g_mc) {
      acc = mm_ld_reduce_f32x4(a.g_mc + off);                 // the switch adds all `world` copies
    } else {
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r) {
        if (r < W) {
          const float4 t = ld_f4(a.g[r] + off);
          acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
        }
      }
    }
    if (a.clip) {
      acc.x = fminf(fmaxf(acc.x, a.clip_lo), a.clip_hi); acc.y = fminf(fmaxf(acc.y, a.clip_lo), a.clip_hi);
      acc.z = fminf(fmaxf(acc.z, a.clip_lo), a.clip_hi); acc.w = fminf(fmaxf(acc.w, a.clip_lo), a.clip_hi);
    }
    float4 pp = ld_f4(a.p[rank] + off);
    float4 mm = m_loc ? ld_f4(m_loc + off) : make_float4(0, 0, 0, 0);
    float4 vv = v_loc ? ld_f4(v_loc + off) : make_float4(0, 0, 0, 0);
    zoo_opt_update(pp.x, acc.x, mm.x, vv.x, h);
    zoo_opt_update(pp.y, acc.y, mm.y, vv.y, h);
    zoo_opt_update(pp.z, acc.z, mm.z, vv.z, h);
    zoo_opt_update(pp.w, acc.w, mm.w, vv.w, h);
    if (m_loc) *reinterpret_cast<float4*>(m_loc + off) = mm;
    if (v_loc) *reinterpret_cast<float4*>(v_loc + off) = vv;
    uint2 pk;
    pk.x = pack_bf16(pp.x, pp.y); pk.y = pack_bf16(pp.z, pp.w);
    if (a.replicated) {
      *reinterpret_cast<float4*>(a.p[rank] + off) = pp;
      if (a.p16[rank]) *reinterpret_cast<uint2*>(a.p16[rank] + off) = pk;
    } else if (a.p_mc) {
      mm_st_f32x4(a.p_mc + off, pp);                          // one store, replicated by the switch into every copy
      if (a.p16_mc) mm_st_bf16x4(a.p16_mc + off, pk);
      else if (a.p16[rank]) {
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r) if (r < W) *reinterpret_cast<uint2*>(a.p16[r] + off) = pk;
      }
    } else {
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r) {
        if (r < W) {
          *reinterpret_cast<float4*>(a.p[r] + off) = pp;
          if (a.p16[r]) *reinterpret_cast<uint2*>(a.p16[r] + off) = pk;
        }
      }
    }
  }

  // ---- phase B: my stores are visible everywhere; nobody reads my gradients any more -------
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < W) {
    const int peer = threadIdx.x;
    st_release_sys(a.flags[peer] + ((1 * W + rank) * kFlagStride + b), ep);
    const uint32_t* mine = a.flags[rank] + ((1 * W + peer) * kFlagStride + b);
    const long long t0 = clock64();
    while (ld_acquire_sys(mine) < ep) {
      if (clock64() - t0 > 20000000000LL) { printf("zoo rs_opt_ag: phase B timeout rank %d cta %d peer %d\n", rank, b, peer); __trap(); }
    }
  }
  __syncthreads();
  if (a.zero_grad) {
    // unit u of MY gradient arena was read only by CTA (u - shard start) % G ... of its owner; all
    // owners' CTAs with my index have signalled, so zero exactly the units those CTAs read.
    float* gl = a.g[rank];
    for (int s = 0; s < (a.replicated ? 1 : W); ++s) {
      const long long s0 = (long long)s * shard_units;
      for (long long u = s0 + b * (long long)kRsThreads + threadIdx.x; u < s0 + shard_units; u += (long long)G * kRsThreads)
        *reinterpret_cast<float4*>(gl + a.start + (u << 2)) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  if (threadIdx.x == 0) a.epoch[rank][b] = ep;
}

// Standalone barrier across ranks with the same flag protocol (one CTA, phase slot 0 of a
// dedicated flag array).  Everything enqueued before it on every rank's stream is complete and
// visible system-wide when it returns on any rank.
struct BarArgs {
  uint32_t* flags[kMaxRanks];   // per-rank [world] uint32
  uint32_t* epoch[kMaxRanks];   // per-rank [1]
  int world, rank, emulate;
};

__global__ void p2p_barrier_kernel(const BarArgs a) {
  const int rank = a.emulate ? (int)blockIdx.y : a.rank;
  __shared__ uint32_t s_ep;
  if (threadIdx.x == 0) s_ep = a.epoch[rank][0] + 1u;
  __syncthreads();
  const uint32_t ep = s_ep;
  __threadfence_system();
  if ((int)threadIdx.x < a.world) {
    const int peer = threadIdx.x;
    st_release_sys(a.flags[peer] + rank, ep);
    const long long t0 = clock64();
    while (ld_acquire_sys(a.flags[rank] + peer) < ep) {
      if (clock64() - t0 > 20000000000LL) { printf("zoo p2p_barrier: timeout rank %d peer %d\n", rank, peer); __trap(); }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) a.epoch[rank][0] = ep;
}

}  // namespace zoo

using namespace zoo;

// ------------------------------------------------------------------ symmetric heap (C1) --------
// cudaMalloc'ed slabs exported with CUDA IPC so every rank maps every other rank's slab
// (replaces the Spark BlockManager as the transport for gradients / weights / embedding rows).
ZOO_API int zoo_symm_alloc(long long bytes, void** ptr, void* handle64) {
  if (!ptr || !handle64 || bytes <= 0) return ZOO_ERR_ARG;
  ZOO_CUDA_CHECK(cudaMalloc(ptr, (size_t)bytes));
  ZOO_CUDA_CHECK(cudaMemset(*ptr, 0, (size_t)bytes));
  cudaIpcMemHandle_t h;
  ZOO_CUDA_CHECK(cudaIpcGetMemHandle(&h, *ptr));
  memcpy(handle64, &h, sizeof(h));
  return ZOO_OK;
}
ZOO_API int zoo_symm_open(const void* handle64, void** ptr) {
  if (!ptr || !handle64) return ZOO_ERR_ARG;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  ZOO_CUDA_CHECK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return ZOO_OK;
}
ZOO_API int zoo_symm_close(void* ptr) {
  if (!ptr) return ZOO_ERR_ARG;
  ZOO_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
  return ZOO_OK;
}
ZOO_API int zoo_symm_free(void* ptr) {
  if (!ptr) return ZOO_ERR_ARG;
  ZOO_CUDA_CHECK(cudaFree(ptr));
  return ZOO_OK;
}
// Plain device allocation (emulation mode in tests needs memory that is NOT from torch's pool).
ZOO_API int zoo_ipc_handle_bytes() { return (int)sizeof(cudaIpcMemHandle_t); }

// ptrs: HOST arrays of `world` device pointers each (m/v/p16 entries may be null).
// zoo_rs_opt_ag plus: NVLS multicast addresses of the LOCAL arenas (g_mc / p_mc / p16_mc, each may be null: peer loads / stores are
// used for that arena) and constant clipping of the summed gradient inside the kernel (clip != 0: clamp to [clip_lo, clip_hi]).
ZOO_API int zoo_rs_opt_ag_ex(void* const* g, void* const* p, void* const* p16, void* const* m, void* const* v,
                             void* const* flags, void* const* epoch, void* const* hyper, long long start, long long end,
                             int world, int rank, int emulate, int opt, int zero_grad, int grid, const void* g_mc, void* p_mc,
                             void* p16_mc, int clip, float clip_lo, float clip_hi, void* stream);

ZOO_API int zoo_rs_opt_ag(void* const* g, void* const* p, void* const* p16, void* const* m, void* const* v,
                          void* const* flags, void* const* epoch, void* const* hyper, long long start, long long end,
                          int world, int rank, int emulate, int opt, int zero_grad, int grid, void* stream) {
  return zoo_rs_opt_ag_ex(g, p, p16, m, v, flags, epoch, hyper, start, end, world, rank, emulate, opt, zero_grad, grid, nullptr, nullptr,
                          nullptr, 0, 0.f, 0.f, stream);
}

ZOO_API int zoo_rs_opt_ag_ex(void* const* g, void* const* p, void* const* p16, void* const* m, void* const* v,
                             void* const* flags, void* const* epoch, void* const* hyper, long long start, long long end,
                             int world, int rank, int emulate, int opt, int zero_grad, int grid, const void* g_mc, void* p_mc,
                             void* p16_mc, int clip, float clip_lo, float clip_hi, void* stream) {
  if (!g || !p || !flags || !epoch || !hyper || world < 1 || world > kMaxRanks || rank < 0 || rank >= world || grid < 1 || grid > kFlagStride ||
      end <= start || ((end - start) % (4LL * world)) || (start % 4)) {
    zoo_set_last_error("rs_opt_ag: bad arguments (range must be a multiple of 4*world floats)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  RsArgs a{};
  for (int r = 0; r < world; ++r) {
    a.g[r] = (float*)g[r]; a.p[r] = (float*)p[r];
    a.p16[r] = p16 ? (__nv_bfloat16*)p16[r] : nullptr;
    a.m[r] = m ? (float*)m[r] : nullptr; a.v[r] = v ? (float*)v[r] : nullptr;
    a.flags[r] = (uint32_t*)flags[r]; a.epoch[r] = (uint32_t*)epoch[r]; a.hyper[r] = (const float*)hyper[r];
  }
  if (!opt_needs_m(opt) && opt != OPT_SGD) for (int r = 0; r < world; ++r) a.m[r] = nullptr;
  if (!opt_needs_v(opt)) for (int r = 0; r < world; ++r) a.v[r] = nullptr;
  a.start = start; a.end = end; a.world = world; a.rank = rank; a.emulate = emulate; a.opt = opt;
  a.zero_grad = zero_grad & 1; a.replicated = (zero_grad >> 1) & 1;     // bit 0: zero the gradient range afterwards, bit 1: replicated update
  a.g_mc = emulate ? nullptr : (const float*)g_mc;
  a.p_mc = (emulate || a.replicated) ? nullptr : (float*)p_mc;
  a.p16_mc = (emulate || a.replicated) ? nullptr : (__nv_bfloat16*)p16_mc;
  a.clip = clip ? 1 : 0; a.clip_lo = clip_lo; a.clip_hi = clip_hi;
  cudaStream_t s = (cudaStream_t)stream;
  if (emulate) {
    void* args[] = {(void*)&a};
    ZOO_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)rs_opt_ag_kernel, dim3(grid, world), dim3(kRsThreads), args, 0, s));
  } else {
    rs_opt_ag_kernel<<<dim3(grid, 1), kRsThreads, 0, s>>>(a);
    ZOO_CUDA_CHECK(cudaGetLastError());
  }
  return ZOO_OK;
}

// One-to-all push over peer memory: dst[p][dst_off ...] = src[0 .. bytes) for every rank p (own slab included).  Used to
// replicate small per-step metadata (the id matrix of a sharded lookup) so later kernels scan it from local memory.
struct BcastArgs { unsigned char* dst[kMaxRanks]; const unsigned char* src; long long bytes; int world; };
__global__ void __launch_bounds__(256) symm_bcast_kernel(const BcastArgs a) {
  const long long n16 = a.bytes >> 4;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
    const uint4 v = reinterpret_cast<const uint4*>(a.src)[i];
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p)
      if (p < a.world) reinterpret_cast<uint4*>(a.dst[p])[i] = v;
  }
  for (long long i = (n16 << 4) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.bytes; i += stride)
    for (int p = 0; p < a.world; ++p) a.dst[p][i] = a.src[i];
}

ZOO_API int zoo_symm_bcast(void* const* dst, long long dst_off, const void* src, long long bytes, int world, void* stream) {
  if (!dst || !src || bytes < 0 || world < 1 || world > kMaxRanks || (dst_off & 15) || ((uintptr_t)src & 15)) {
    zoo_set_last_error("symm_bcast: bad arguments (offset and source must be 16-byte aligned)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  if (bytes == 0) return ZOO_OK;
  BcastArgs a{};
  for (int p = 0; p < world; ++p) {
    if (!dst[p] || ((uintptr_t)dst[p] & 15)) { zoo_set_last_error("symm_bcast: null or unaligned destination", __FILE__, __LINE__); return ZOO_ERR_ARG; }
    a.dst[p] = (unsigned char*)dst[p] + dst_off;
  }
  a.src = (const unsigned char*)src; a.bytes = bytes; a.world = world;
  long long blocks = ((bytes >> 4) + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 296) blocks = 296;
  symm_bcast_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

ZOO_API int zoo_p2p_barrier(void* const* flags, void* const* epoch, int world, int rank, int emulate, void* stream) {
  if (!flags || !epoch || world < 1 || world > kMaxRanks || rank < 0 || rank >= world) {
    zoo_set_last_error("p2p_barrier: bad arguments", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  BarArgs a{};
  for (int r = 0; r < world; ++r) { a.flags[r] = (uint32_t*)flags[r]; a.epoch[r] = (uint32_t*)epoch[r]; }
  a.world = world; a.rank = rank; a.emulate = emulate;
  cudaStream_t s = (cudaStream_t)stream;
  if (emulate) {
    void* args[] = {(void*)&a};
    ZOO_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)p2p_barrier_kernel, dim3(1, world), dim3(32), args, 0, s));
  } else {
    p2p_barrier_kernel<<<1, 32, 0, s>>>(a);
    ZOO_CUDA_CHECK(cudaGetLastError());
  }
  return ZOO_OK;
}
