// Arena allocator + variable-length record store (component C3) and CRC32C / TFRecord framing.
//
// The reference caches training data off-heap: `NativeArray` / `VarLenBytesArray` over a memkind
// (PMEM) or direct-memory allocator (zoo/.../feature/pmem/NativeArray.scala:25-100,
// zoo/src/main/java/.../pmem/PersistentMemoryAllocator.java:37-43), selected by
// `MemoryType` = DRAM / PMEM / DIRECT / DISK_AND_DRAM(n).  Here the tiers are: anonymous
// page-aligned host memory (the caller pins it for DMA with cudaHostRegister), a file-backed
// mapping (the DISK tier: the page cache holds the resident slice), and HBM (a torch tensor, not
// managed here).  A record store appends byte records into an arena and serves batched gathers
// into a contiguous staging buffer with a few copy threads.
// TensorBoard's event files need masked CRC32C framing (zoo/.../tensorboard/RecordWriter.scala);
// the hardware crc32 instruction is used when built with SSE4.2.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#if defined(__SSE4_2__)
#include <nmmintrin.h>
#endif

#define ZOO_RT_API extern "C" __attribute__((visibility("default")))

namespace {

struct Arena {
  uint8_t* base = nullptr;
  size_t capacity = 0;
  std::atomic<size_t> used{0};
  int fd = -1;
  std::string path;
};

struct Records {
  Arena* arena;
  std::vector<uint64_t> offset;
  std::vector<uint32_t> length;
  std::mutex mu;
};

uint32_t crc_table[8][256];
std::once_flag crc_once;
void crc_init() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
    crc_table[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i)
    for (int t = 1; t < 8; ++t) crc_table[t][i] = (crc_table[t - 1][i] >> 8) ^ crc_table[0][crc_table[t - 1][i] & 0xff];
}

uint32_t crc32c(const uint8_t* p, size_t n) {
  uint32_t c = 0xFFFFFFFFu;
#if defined(__SSE4_2__)
  while (n >= 8) { uint64_t v; memcpy(&v, p, 8); c = (uint32_t)_mm_crc32_u64(c, v); p += 8; n -= 8; }
  while (n--) c = _mm_crc32_u8(c, *p++);
#else
  std::call_once(crc_once, crc_init);
  while (n >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4); memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = crc_table[7][lo & 0xff] ^ crc_table[6][(lo >> 8) & 0xff] ^ crc_table[5][(lo >> 16) & 0xff] ^ crc_table[4][lo >> 24] ^
        crc_table[3][hi & 0xff] ^ crc_table[2][(hi >> 8) & 0xff] ^ crc_table[1][(hi >> 16) & 0xff] ^ crc_table[0][hi >> 24];
    p += 8; n -= 8;
  }
  while (n--) c = crc_table[0][(c ^ *p++) & 0xff] ^ (c >> 8);
#endif
  return c ^ 0xFFFFFFFFu;
}

uint32_t masked(uint32_t crc) { return ((crc >> 15) | (crc << 17)) + 0xA282EAD8u; }

}  // namespace

// kind 0: anonymous host memory; kind 1: file-backed mapping at `path` (created/extended to `capacity`)
ZOO_RT_API void* zoo_arena_create(int kind, long long capacity, const char* path) {
  if (capacity <= 0) return nullptr;
  Arena* a = new Arena();
  a->capacity = (size_t)capacity;
  void* p = MAP_FAILED;
  if (kind == 1) {
    if (!path) { delete a; return nullptr; }
    a->path = path;
    a->fd = ::open(path, O_RDWR | O_CREAT, 0644);
    if (a->fd < 0 || ftruncate(a->fd, (off_t)capacity) != 0) { if (a->fd >= 0) ::close(a->fd); delete a; return nullptr; }
    p = mmap(nullptr, a->capacity, PROT_READ | PROT_WRITE, MAP_SHARED, a->fd, 0);
  } else {
    p = mmap(nullptr, a->capacity, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
  }
  if (p == MAP_FAILED) { if (a->fd >= 0) ::close(a->fd); delete a; return nullptr; }
  a->base = (uint8_t*)p;
  return a;
}
ZOO_RT_API void zoo_arena_destroy(void* h) {
  Arena* a = (Arena*)h;
  if (!a) return;
  if (a->base) munmap(a->base, a->capacity);
  if (a->fd >= 0) ::close(a->fd);
  delete a;
}
ZOO_RT_API void* zoo_arena_base(void* h) { return ((Arena*)h)->base; }
ZOO_RT_API long long zoo_arena_capacity(void* h) { return (long long)((Arena*)h)->capacity; }
ZOO_RT_API long long zoo_arena_used(void* h) { return (long long)((Arena*)h)->used.load(); }
ZOO_RT_API void zoo_arena_reset(void* h) { ((Arena*)h)->used.store(0); }
// Lock-free bump allocation; returns the byte offset or -1 when the arena is full.
ZOO_RT_API long long zoo_arena_alloc(void* h, long long nbytes, int align) {
  Arena* a = (Arena*)h;
  if (nbytes < 0) return -1;
  const size_t al = align > 0 ? (size_t)align : 64;
  size_t cur = a->used.load();
  while (true) {
    size_t start = (cur + al - 1) / al * al;
    if (start + (size_t)nbytes > a->capacity) return -1;
    if (a->used.compare_exchange_weak(cur, start + (size_t)nbytes)) return (long long)start;
  }
}
// Drop the resident pages of [offset, offset+len) (DISK tier: keep only the active slice in DRAM).
ZOO_RT_API int zoo_arena_evict(void* h, long long offset, long long len) {
  Arena* a = (Arena*)h;
  const size_t pg = (size_t)sysconf(_SC_PAGESIZE);
  size_t lo = ((size_t)offset + pg - 1) / pg * pg, hi = ((size_t)offset + (size_t)len) / pg * pg;
  if (hi <= lo) return 0;
  if (a->fd >= 0) msync(a->base + lo, hi - lo, MS_SYNC);
  return madvise(a->base + lo, hi - lo, a->fd >= 0 ? MADV_DONTNEED : MADV_FREE);
}

ZOO_RT_API void* zoo_records_create(void* arena) { Records* r = new Records(); r->arena = (Arena*)arena; return r; }
ZOO_RT_API void zoo_records_destroy(void* h) { delete (Records*)h; }
ZOO_RT_API long long zoo_records_count(void* h) { return (long long)((Records*)h)->offset.size(); }
ZOO_RT_API long long zoo_records_append(void* h, const void* data, long long len) {
  Records* r = (Records*)h;
  if (len < 0 || len > 0xFFFFFFFFLL) return -1;
  long long off = zoo_arena_alloc(r->arena, len, 16);
  if (off < 0) return -1;
  memcpy(r->arena->base + off, data, (size_t)len);
  std::lock_guard<std::mutex> lk(r->mu);
  r->offset.push_back((uint64_t)off);
  r->length.push_back((uint32_t)len);
  return (long long)r->offset.size() - 1;
}
ZOO_RT_API long long zoo_records_get(void* h, long long i, const void** ptr) {
  Records* r = (Records*)h;
  if (i < 0 || (size_t)i >= r->offset.size()) return -1;
  *ptr = r->arena->base + r->offset[(size_t)i];
  return (long long)r->length[(size_t)i];
}
// Copy records idx[0..n) back to back into dst; out_off[k] = start of record k, out_off[n] = total.
// Returns total bytes, or -1 if dst_cap is too small / an index is out of range.
ZOO_RT_API long long zoo_records_gather(void* h, const long long* idx, long long n, void* dst, long long dst_cap, long long* out_off, int threads) {
  Records* r = (Records*)h;
  long long total = 0;
  for (long long k = 0; k < n; ++k) {
    if (idx[k] < 0 || (size_t)idx[k] >= r->offset.size()) return -1;
    out_off[k] = total;
    total += r->length[(size_t)idx[k]];
  }
  out_off[n] = total;
  if (total > dst_cap) return -1;
  auto work = [&](long long lo, long long hi) {
    for (long long k = lo; k < hi; ++k)
      memcpy((uint8_t*)dst + out_off[k], r->arena->base + r->offset[(size_t)idx[k]], r->length[(size_t)idx[k]]);
  };
  if (threads <= 1 || total < (1 << 20) || n < 2 * threads) { work(0, n); return total; }
  std::vector<std::thread> ts;
  const long long per = (n + threads - 1) / threads;
  for (int t = 0; t < threads; ++t) { long long lo = t * per, hi = std::min(n, lo + per); if (lo < hi) ts.emplace_back(work, lo, hi); }
  for (auto& t : ts) t.join();
  return total;
}

ZOO_RT_API unsigned int zoo_crc32c(const void* data, long long n) { return crc32c((const uint8_t*)data, (size_t)n); }
ZOO_RT_API unsigned int zoo_masked_crc32c(const void* data, long long n) { return masked(crc32c((const uint8_t*)data, (size_t)n)); }
// TFRecord frame: u64 length | u32 masked_crc(length) | payload | u32 masked_crc(payload); out has n + 16 bytes.
ZOO_RT_API long long zoo_tfrecord_frame(const void* data, long long n, void* out) {
  uint8_t* o = (uint8_t*)out;
  uint64_t len = (uint64_t)n;
  memcpy(o, &len, 8);
  uint32_t c = masked(crc32c(o, 8));
  memcpy(o + 8, &c, 4);
  memcpy(o + 12, data, (size_t)n);
  c = masked(crc32c((const uint8_t*)data, (size_t)n));
  memcpy(o + 12 + n, &c, 4);
  return n + 16;
}

// ------------------------------------------------------------------------------------------------------------------------
// Batch base64 decoding on native threads.  The serving client ships images base64-encoded inside stream records
// (pyzoo/zoo/serving/client.py:87-126); CPython's decoder holds the interpreter lock, so a thread pool decoded a 64-image batch
// (~6.7 MB of text) serially in ~8 ms.  Here `n` strings are decoded by `threads` native threads into one output buffer.
//   srcs / lens : the encoded strings;  out + out_off[i] : where record i goes (capacity 3/4 of its length, rounded up);
//   out_len[i]  : decoded size, or -1 for a malformed string.
#include <thread>
#include <vector>

namespace {
static int8_t g_b64[256];
static bool g_b64_init = [] {
  for (int i = 0; i < 256; ++i) g_b64[i] = -1;
  const char* abc = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
  for (int i = 0; i < 64; ++i) g_b64[(unsigned char)abc[i]] = (int8_t)i;
  g_b64[(unsigned char)'-'] = 62; g_b64[(unsigned char)'_'] = 63;      // url-safe alphabet too
  return true;
}();

long long b64_decode_one(const unsigned char* s, long long n, unsigned char* o) {
  while (n > 0 && (s[n - 1] == '=' || s[n - 1] == '\n' || s[n - 1] == '\r')) --n;
  long long w = 0;
  uint32_t acc = 0;
  int bits = 0;
  for (long long i = 0; i < n; ++i) {
    const int8_t v = g_b64[s[i]];
    if (v < 0) { if (s[i] == '\n' || s[i] == '\r') continue; return -1; }
    acc = (acc << 6) | (uint32_t)v;
    bits += 6;
    if (bits >= 8) { bits -= 8; o[w++] = (unsigned char)(acc >> bits); }
  }
  return w;
}
}  // namespace

ZOO_RT_API int zoo_b64_decode_batch(const unsigned char* const* srcs, const long long* lens, int n, unsigned char* out,
                                    const long long* out_off, long long* out_len, int threads) {
  if (!srcs || !lens || !out || !out_off || !out_len || n < 0) return 1;
  if (threads < 1) threads = 1;
  if (threads > n) threads = n > 0 ? n : 1;
  auto work = [&](int t) {
    for (int i = t; i < n; i += threads) out_len[i] = b64_decode_one(srcs[i], lens[i], out + out_off[i]);
  };
  if (threads == 1) { work(0); return 0; }
  std::vector<std::thread> ts;
  for (int t = 1; t < threads; ++t) ts.emplace_back(work, t);
  work(0);
  for (auto& t : ts) t.join();
  return 0;
}
