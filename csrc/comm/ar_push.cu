// K1, small-range form: one-handshake gradient all-reduce over NVLink peer memory with a bf16 wire format.
//
// The dense part of a recommender is a few hundred KB (NeuralCF towers: 115 K parameters).  The owner-computes kernel of
// rs_opt_ag.cu needs two flag round trips per step for it (gradients ready -> reduce -> publish -> done), i.e. 40-60 us of pure
// NVLink latency for 3.6 us of payload.  Here every rank PUSHES its gradient range, rounded to bf16 (the reference ships
// its blocks FP16-compressed: Topology.scala:1043-1047), into a mailbox on every peer with posted stores and raises one flag
// per CTA; much later in the step -- after the sparse embedding updates, which need no dense gradients -- a second kernel
// checks the flags (they arrived long ago), sums the `world` copies in rank order in fp32 and leaves the SUM in the local
// gradient arena, where the ordinary fused optimizer pass (clipping included) picks it up.  Every rank adds the same bf16
// values in the same order, so the replicas stay bitwise identical.  One NVLink latency per step, none of it exposed.
//
// Mailboxes are double-buffered by step parity: rank A can only write step s+2 into parity p after its own reduce of step
// s+1, which waited for B's push of s+1, which B issued after finishing its reduce of step s (the last reader of parity p).
//
// Replaces: AllReduceParameter's put / aggregate of gradient slices through the Spark BlockManager
// (zoo/.../pipeline/api/keras/models/Topology.scala:1128-1206, docs/docs/wp-bigdl.md:146-156).
#include "../common/zoo_common.cuh"

namespace zoo {

constexpr int kArMaxRanks = 8;
constexpr int kArThreads = 256;
constexpr int kArMaxCtas = 1024;

struct ArArgs {
  float* g;                           // local gradient range [n]
  __nv_bfloat16* mail[kArMaxRanks];   // every rank's mailbox: [2 parities][world sources][n] bf16
  uint32_t* flags[kArMaxRanks];       // every rank's flags:   [2 parities][world sources][kArMaxCtas]
  uint32_t* epoch;                    // local [2][kArMaxCtas]: steps pushed / reduced so far, per CTA
  long long n;                        // floats, multiple of 8
  int world, rank;
};

__global__ void __launch_bounds__(kArThreads)
ar_push_kernel(const __grid_constant__ ArArgs a) {
  __shared__ uint32_t s_ep;
  const int b = blockIdx.x, G = gridDim.x;
  if (threadIdx.x == 0) s_ep = a.epoch[b] + 1u;
  __syncthreads();
  const uint32_t ep = s_ep;
  const long long slot = ((long long)(ep & 1u) * a.world + a.rank) * a.n;     // [parity][my rank] in every mailbox
  const long long units = a.n >> 3;
  for (long long u = (long long)b * kArThreads + threadIdx.x; u < units; u += (long long)G * kArThreads) {
    const float4 x = *reinterpret_cast<const float4*>(a.g + (u << 3));
    const float4 y = *reinterpret_cast<const float4*>(a.g + (u << 3) + 4);
    uint4 pk;
    pk.x = pack_bf16(x.x, x.y); pk.y = pack_bf16(x.z, x.w); pk.z = pack_bf16(y.x, y.y); pk.w = pack_bf16(y.z, y.w);
#pragma unroll
    for (int r = 0; r < kArMaxRanks; ++r)
      if (r < a.world) *reinterpret_cast<uint4*>(a.mail[r] + slot + (u << 3)) = pk;       // own mailbox included
  }
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < a.world)
    st_release_sys(a.flags[threadIdx.x] + ((long long)(ep & 1u) * a.world + a.rank) * kArMaxCtas + b, ep);
  if (threadIdx.x == 0) a.epoch[b] = ep;
}

__global__ void __launch_bounds__(kArThreads)
ar_reduce_kernel(const __grid_constant__ ArArgs a) {
  __shared__ uint32_t s_ep;
  const int b = blockIdx.x, G = gridDim.x;
  if (threadIdx.x == 0) s_ep = a.epoch[kArMaxCtas + b] + 1u;
  __syncthreads();
  const uint32_t ep = s_ep;
  const uint32_t par = ep & 1u;
  if ((int)threadIdx.x < a.world) {
    const uint32_t* f = a.flags[a.rank] + ((long long)par * a.world + threadIdx.x) * kArMaxCtas + b;
    const long long t0 = clock64();
    while (ld_acquire_sys(f) < ep) {
      if (clock64() - t0 > 20000000000LL) { printf("zoo ar_reduce: timeout rank %d cta %d source %d step %u\n", a.rank, b, (int)threadIdx.x, ep); __trap(); }
    }
  }
  __syncthreads();
  const __nv_bfloat16* box = a.mail[a.rank] + (long long)par * a.world * a.n;
  const long long units = a.n >> 3;
  for (long long u = (long long)b * kArThreads + threadIdx.x; u < units; u += (long long)G * kArThreads) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < kArMaxRanks; ++r) {
      if (r < a.world) {
        const uint4 q = *reinterpret_cast<const uint4*>(box + (long long)r * a.n + (u << 3));
        const float2 c0 = unpack_bf16(q.x), c1 = unpack_bf16(q.y), c2 = unpack_bf16(q.z), c3 = unpack_bf16(q.w);
        acc[0] += c0.x; acc[1] += c0.y; acc[2] += c1.x; acc[3] += c1.y; acc[4] += c2.x; acc[5] += c2.y; acc[6] += c3.x; acc[7] += c3.y;
      }
    }
    *reinterpret_cast<float4*>(a.g + (u << 3)) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    *reinterpret_cast<float4*>(a.g + (u << 3) + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
  if (threadIdx.x == 0) a.epoch[kArMaxCtas + b] = ep;
}

}  // namespace zoo

using namespace zoo;

static int ar_fill(ArArgs& a, float* g, void* const* mail, void* const* flags, unsigned int* epoch, long long n, int world, int rank,
                   int grid) {
  if (!g || !mail || !flags || !epoch || n <= 0 || (n & 7) || ((uintptr_t)g & 15) || world < 1 || world > kArMaxRanks || rank < 0 ||
      rank >= world || grid < 1 || grid > kArMaxCtas) {
    zoo_set_last_error("ar_push/ar_reduce: bad arguments (n must be a multiple of 8 floats, grid <= 1024)", __FILE__, __LINE__);
    return ZOO_ERR_ARG;
  }
  for (int r = 0; r < world; ++r) {
    if (!mail[r] || !flags[r] || ((uintptr_t)mail[r] & 15)) { zoo_set_last_error("ar_push/ar_reduce: null or unaligned mailbox", __FILE__, __LINE__); return ZOO_ERR_ARG; }
    a.mail[r] = (__nv_bfloat16*)mail[r];
    a.flags[r] = (uint32_t*)flags[r];
  }
  a.g = g; a.epoch = epoch; a.n = n; a.world = world; a.rank = rank;
  return ZOO_OK;
}

// Push this rank's gradient range (bf16) into every rank's mailbox and flag it.  mail / flags: HOST arrays of `world` device
// pointers into the symmetric heap ([2][world][n] bf16 and [2][world][1024] uint32, zero at start); epoch: local uint32 [2][1024],
// zero at start.  Both kernels of a step must use the same `grid` on every rank.
ZOO_API int zoo_ar_push(float* g, void* const* mail, void* const* flags, unsigned int* epoch, long long n, int world, int rank,
                        int grid, void* stream) {
  ArArgs a{};
  const int rc = ar_fill(a, g, mail, flags, epoch, n, world, rank, grid);
  if (rc) return rc;
  ar_push_kernel<<<grid, kArThreads, 0, (cudaStream_t)stream>>>(a);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}

// Wait for every rank's push of this step and leave the sum of all `world` bf16 copies (rank order, fp32) in g.
ZOO_API int zoo_ar_reduce(float* g, void* const* mail, void* const* flags, unsigned int* epoch, long long n, int world, int rank,
                          int grid, void* stream) {
  ArArgs a{};
  const int rc = ar_fill(a, g, mail, flags, epoch, n, world, rank, grid);
  if (rc) return rc;
  ar_reduce_kernel<<<grid, kArThreads, 0, (cudaStream_t)stream>>>(a);
  ZOO_CUDA_CHECK(cudaGetLastError());
  return ZOO_OK;
}
