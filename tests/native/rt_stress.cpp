// Thread-sanitizer stress of the C++ runtime (stream broker + arena + record store) through its C ABI.
// Build and run:  dev/tsan_runtime.sh   (g++ -fsanitize=thread over csrc/runtime/*.cpp + this file; output kept in profiles/)
// Workload: producers XADD into one stream while consumer-group readers drain it (XREADGROUP / read_pack), writers HSET results,
// a trimmer XTRIMs, pollers XLEN / INFO; in parallel threads bump-allocate from one arena and append / gather records.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <arpa/inet.h>
#include <sys/socket.h>
#include <unistd.h>

extern "C" {
int zoo_broker_serve(void* b, const char* host, int port);
void zoo_broker_stop_server(void* b);
void* zoo_broker_create(long long maxmemory);
void zoo_broker_destroy(void* b);
int zoo_broker_command(void* b, int argc, const char* const* argv, const long long* arglen, char** out, long long* outlen);
void zoo_broker_free(char* p);
long long zoo_broker_read_pack(void* b, const char* key, const char* group, long long count, long long block_ms, const char* field, char* data,
                               long long data_cap, long long* rec_len, char* uris, long long uris_cap, long long* uris_len);
long long zoo_broker_used_memory(void* b);
void* zoo_arena_create(int kind, long long capacity, const char* path);
void zoo_arena_destroy(void* h);
long long zoo_arena_alloc(void* h, long long nbytes, int align);
long long zoo_arena_used(void* h);
void* zoo_arena_base(void* h);
void* zoo_records_create(void* arena);
void zoo_records_destroy(void* h);
long long zoo_records_append(void* h, const void* data, long long len);
long long zoo_records_count(void* h);
long long zoo_records_gather(void* h, const long long* idx, long long n, void* dst, long long dst_cap, long long* out_off, int threads);
}

static std::string cmd(void* b, std::vector<std::string> args) {
  std::vector<const char*> argv;
  std::vector<long long> len;
  for (auto& a : args) { argv.push_back(a.data()); len.push_back((long long)a.size()); }
  char* out = nullptr;
  long long n = 0;
  zoo_broker_command(b, (int)argv.size(), argv.data(), len.data(), &out, &n);
  std::string r(out ? out : "", (size_t)(out ? n : 0));
  if (out) zoo_broker_free(out);
  return r;
}

// One RESP client over a TCP connection: pipelines `n` XADDs and an XLEN, reads until the replies are in; then a malformed frame,
// which the server must answer with -ERR and a closed connection instead of buffering forever.
static long long resp_client(int port, int id, int n) {
  int fd = socket(AF_INET, SOCK_STREAM, 0);
  sockaddr_in a{};
  a.sin_family = AF_INET;
  a.sin_port = htons((uint16_t)port);
  inet_pton(AF_INET, "127.0.0.1", &a.sin_addr);
  if (connect(fd, (sockaddr*)&a, sizeof a) != 0) return -1;
  std::string req;
  for (int i = 0; i < n; ++i) {
    std::string uri = "c" + std::to_string(id) + "_" + std::to_string(i);
    req += "*5\r\n$4\r\nXADD\r\n$4\r\nsock\r\n$1\r\n*\r\n$3\r\nuri\r\n$" + std::to_string(uri.size()) + "\r\n" + uri + "\r\n";
  }
  req += "*2\r\n$4\r\nXLEN\r\n$4\r\nsock\r\n";
  size_t sent = 0;
  while (sent < req.size()) {
    ssize_t k = send(fd, req.data() + sent, req.size() - sent, 0);
    if (k <= 0) break;
    sent += (size_t)k;
  }
  long long replies = 0;
  char buf[4096];
  while (replies < 2 * n + 1) {                       // an XADD answers with a bulk string (two line ends), XLEN with an integer (one)
    ssize_t k = recv(fd, buf, sizeof buf, 0);
    if (k <= 0) break;
    for (ssize_t i = 0; i + 1 < k; ++i) replies += (buf[i] == '\r' && buf[i + 1] == '\n');
  }
  const char* bad = "*2\r\nthis is not a bulk string\r\n";
  send(fd, bad, strlen(bad), 0);
  std::string tail;                                    // -ERR ... then EOF
  for (ssize_t k = recv(fd, buf, sizeof buf, 0); k > 0; k = recv(fd, buf, sizeof buf, 0)) tail.append(buf, (size_t)k);
  bool err = tail.rfind("-ERR", 0) == 0;
  close(fd);
  return err ? replies : -replies;
}

int main() {
  void* b = zoo_broker_create(0);
  cmd(b, {"XGROUP", "CREATE", "s", "g", "0", "MKSTREAM"});
  const int producers = 4, per = 2000;
  std::atomic<long long> consumed{0}, produced{0};
  std::atomic<bool> stop{false};
  std::vector<std::thread> ts;
  for (int p = 0; p < producers; ++p)
    ts.emplace_back([&, p] {
      std::string payload(64, (char)('a' + p));
      for (int i = 0; i < per; ++i) {
        cmd(b, {"XADD", "s", "*", "uri", "u" + std::to_string(p) + "_" + std::to_string(i), "image", payload});
        produced++;
      }
    });
  for (int c = 0; c < 3; ++c)
    ts.emplace_back([&, c] {
      std::vector<char> data(64 * 64), uris(64 * 32);
      while (consumed.load() < (long long)producers * per) {
        long long rec_len = 0, uris_len = 0;
        long long n = zoo_broker_read_pack(b, "s", "g", 32, 2, "image", data.data(), (long long)data.size(), &rec_len, uris.data(), (long long)uris.size(), &uris_len);
        if (n > 0) {
          consumed += n;
          cmd(b, {"HSET", "result:c" + std::to_string(c) + "_" + std::to_string(consumed.load()), "value", "ok"});
        }
      }
    });
  ts.emplace_back([&] { while (!stop.load()) { cmd(b, {"XLEN", "s"}); cmd(b, {"INFO", "memory"}); zoo_broker_used_memory(b); } });
  ts.emplace_back([&] { while (!stop.load()) { cmd(b, {"XTRIM", "s", "MAXLEN", "100000"}); cmd(b, {"KEYS", "result:*"}); } });

  void* arena = zoo_arena_create(0, 64ll << 20, nullptr);
  void* recs = zoo_records_create(arena);
  std::atomic<long long> appended{0};
  for (int w = 0; w < 4; ++w)
    ts.emplace_back([&, w] {
      char buf[256];
      memset(buf, w, sizeof buf);
      for (int i = 0; i < 3000; ++i) {
        zoo_arena_alloc(arena, 100 + (i % 50), 16);
        if (zoo_records_append(recs, buf, 32 + (i % 200)) >= 0) appended++;
      }
    });
  for (size_t i = 0; i < ts.size() - 2 - 4; ++i) ts[i].join();          // producers and consumers
  stop = true;
  for (size_t i = ts.size() - 2 - 4; i < ts.size(); ++i) ts[i].join();
  int port = zoo_broker_serve(b, "127.0.0.1", 0);
  std::vector<long long> sock_replies(3, 0);
  {
    std::vector<std::thread> cs;
    for (int c = 0; c < 3; ++c) cs.emplace_back([&, c] { sock_replies[c] = resp_client(port, c, 300); });
    for (auto& t : cs) t.join();
  }
  std::string xlen = cmd(b, {"XLEN", "sock"});
  zoo_broker_stop_server(b);
  printf("RESP server on port %d: replies per client %lld %lld %lld (negative = malformed frame not answered with -ERR), XLEN sock -> %s", port, sock_replies[0],
         sock_replies[1], sock_replies[2], xlen.c_str());
  long long n = zoo_records_count(recs);
  std::vector<long long> idx, off((size_t)n + 1);
  for (long long i = 0; i < n; i += 7) idx.push_back(i);
  off.resize(idx.size() + 1);
  std::vector<char> dst(idx.size() * 256);
  long long got = zoo_records_gather(recs, idx.data(), (long long)idx.size(), dst.data(), (long long)dst.size(), off.data(), 4);
  printf("produced %lld consumed %lld records appended %lld (count %lld) gathered %lld bytes arena used %lld\n", produced.load(), consumed.load(), appended.load(), n, got,
         zoo_arena_used(arena));
  int ok = produced.load() == (long long)producers * per && consumed.load() == produced.load() && appended.load() == n && got > 0 && port > 0 &&
           sock_replies[0] > 0 && sock_replies[1] > 0 && sock_replies[2] > 0 && xlen.find("900") != std::string::npos;
  zoo_records_destroy(recs);
  zoo_arena_destroy(arena);
  zoo_broker_destroy(b);
  printf(ok ? "RT_STRESS_OK\n" : "RT_STRESS_MISMATCH\n");
  return ok ? 0 : 1;
}
