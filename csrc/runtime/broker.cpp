// Stream broker (component C4): the data plane of Cluster Serving.
//
// The reference moves serving traffic through a Redis server: clients XADD images to the stream
// `image_stream`, Spark Structured Streaming reads it with a consumer group, results are written as
// hashes `result:<uri>` (zoo/.../serving/ClusterServing.scala:107-118,278-283;
// pyzoo/zoo/serving/client.py:87-151).  On one box that is an extra process and two TCP hops per
// record, so the broker lives INSIDE the serving process: the engine calls it directly
// (zoo_broker_command, no socket), and remote clients reach the same object through a small RESP2
// listener, so any Redis client speaking XADD / XREADGROUP / HGETALL keeps working.
// Implemented verbs: PING ECHO XADD XLEN XRANGE XREAD XGROUP(CREATE|DESTROY) XREADGROUP XACK XTRIM
// XDEL HSET HGET HGETALL HDEL SET GET KEYS DEL EXISTS DBSIZE FLUSHALL INFO CONFIG(SET|GET maxmemory).
// Memory accounting mirrors what the reference relies on: `INFO memory` reports used_memory /
// maxmemory so the 60 % client back-pressure and 48 % server trim rules work unchanged.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <charconv>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#define ZOO_RT_API extern "C" __attribute__((visibility("default")))

namespace {

struct StreamId {
  uint64_t ms = 0, seq = 0;
  bool operator<(const StreamId& o) const { return ms < o.ms || (ms == o.ms && seq < o.seq); }
  bool operator==(const StreamId& o) const { return ms == o.ms && seq == o.seq; }
  std::string str() const { return std::to_string(ms) + "-" + std::to_string(seq); }
};

struct Entry {
  StreamId id;
  std::vector<std::pair<std::string, std::string>> fields;
  size_t bytes = 0;
};

struct Group {
  StreamId last_delivered;
  std::set<std::pair<uint64_t, uint64_t>> pending;
};

struct Stream {
  std::deque<Entry> entries;
  StreamId last_id;
  std::map<std::string, Group> groups;
};

bool parse_id(const std::string& s, StreamId* id) {
  if (s.empty()) return false;
  size_t dash = s.find('-');
  try {
    id->ms = std::stoull(s.substr(0, dash));
    id->seq = dash == std::string::npos ? 0 : std::stoull(s.substr(dash + 1));
  } catch (...) { return false; }
  return true;
}

bool glob_match(const char* p, const char* s) {
  for (; *p; ++p, ++s) {
    if (*p == '*') {
      while (*p == '*') ++p;
      if (!*p) return true;
      for (; *s; ++s) if (glob_match(p, s)) return true;
      return false;
    }
    if (!*s) return false;
    if (*p != '?' && *p != *s) return false;
  }
  return !*s;
}

std::string upper(std::string s) { for (auto& c : s) c = (char)toupper((unsigned char)c); return s; }

// ---- RESP2 encoding -----------------------------------------------------------------------------
void r_simple(std::string& o, const char* s) { o += '+'; o += s; o += "\r\n"; }
void r_err(std::string& o, const std::string& s) { o += "-ERR "; o += s; o += "\r\n"; }
void r_int(std::string& o, long long v) { o += ':'; o += std::to_string(v); o += "\r\n"; }
void r_bulk(std::string& o, const std::string& s) { o += '$'; o += std::to_string(s.size()); o += "\r\n"; o += s; o += "\r\n"; }
void r_null(std::string& o) { o += "$-1\r\n"; }
void r_nullarr(std::string& o) { o += "*-1\r\n"; }
void r_arr(std::string& o, size_t n) { o += '*'; o += std::to_string(n); o += "\r\n"; }

void r_entry(std::string& o, const Entry& e) {
  r_arr(o, 2);
  r_bulk(o, e.id.str());
  r_arr(o, e.fields.size() * 2);
  for (auto& f : e.fields) { r_bulk(o, f.first); r_bulk(o, f.second); }
}

class Broker {
 public:
  explicit Broker(long long maxmem) : maxmemory_(maxmem) {}
  ~Broker() { stop_server(); }

  std::string execute(const std::vector<std::string>& a) {
    std::string out;
    if (a.empty()) { r_err(out, "empty command"); return out; }
    const std::string cmd = upper(a[0]);
    if (cmd == "XREAD" || cmd == "XREADGROUP") return xread(a, cmd == "XREADGROUP");
    std::lock_guard<std::mutex> lk(mu_);
    if (cmd == "PING") { if (a.size() > 1) r_bulk(out, a[1]); else r_simple(out, "PONG"); }
    else if (cmd == "ECHO" && a.size() == 2) r_bulk(out, a[1]);
    else if (cmd == "XADD") xadd(a, out);
    else if (cmd == "XLEN" && a.size() == 2) { auto it = streams_.find(a[1]); r_int(out, it == streams_.end() ? 0 : (long long)it->second.entries.size()); }
    else if (cmd == "XRANGE" && a.size() >= 4) xrange(a, out);
    else if (cmd == "XGROUP") xgroup(a, out);
    else if (cmd == "XACK" && a.size() >= 4) xack(a, out);
    else if (cmd == "XTRIM" && a.size() >= 4) xtrim(a, out);
    else if (cmd == "XDEL" && a.size() >= 3) xdel(a, out);
    else if (cmd == "HSET" && a.size() >= 4 && a.size() % 2 == 0) hset(a, out);
    else if (cmd == "HGET" && a.size() == 3) { auto it = hashes_.find(a[1]); if (it == hashes_.end() || !it->second.count(a[2])) r_null(out); else r_bulk(out, it->second[a[2]]); }
    else if (cmd == "HGETALL" && a.size() == 2) { auto it = hashes_.find(a[1]); if (it == hashes_.end()) r_arr(out, 0); else { r_arr(out, it->second.size() * 2); for (auto& kv : it->second) { r_bulk(out, kv.first); r_bulk(out, kv.second); } } }
    else if (cmd == "HDEL" && a.size() >= 3) { long long n = 0; auto it = hashes_.find(a[1]); if (it != hashes_.end()) { for (size_t i = 2; i < a.size(); ++i) { auto f = it->second.find(a[i]); if (f != it->second.end()) { used_ -= f->first.size() + f->second.size(); it->second.erase(f); ++n; } } if (it->second.empty()) hashes_.erase(it); } r_int(out, n); }
    else if (cmd == "SET" && a.size() >= 3) { auto& v = strings_[a[1]]; used_ += a[2].size(); used_ -= v.size(); v = a[2]; r_simple(out, "OK"); }
    else if (cmd == "GET" && a.size() == 2) { auto it = strings_.find(a[1]); if (it == strings_.end()) r_null(out); else r_bulk(out, it->second); }
    else if (cmd == "KEYS" && a.size() == 2) keys(a[1], out);
    else if (cmd == "DEL" && a.size() >= 2) { long long n = 0; for (size_t i = 1; i < a.size(); ++i) n += del_key(a[i]); r_int(out, n); }
    else if (cmd == "EXISTS" && a.size() >= 2) { long long n = 0; for (size_t i = 1; i < a.size(); ++i) n += streams_.count(a[i]) + hashes_.count(a[i]) + strings_.count(a[i]); r_int(out, n); }
    else if (cmd == "DBSIZE") r_int(out, (long long)(streams_.size() + hashes_.size() + strings_.size()));
    else if (cmd == "FLUSHALL" || cmd == "FLUSHDB") { streams_.clear(); hashes_.clear(); strings_.clear(); used_ = 0; r_simple(out, "OK"); }
    else if (cmd == "INFO") { std::string s = "# Memory\r\nused_memory:" + std::to_string(used_) + "\r\nmaxmemory:" + std::to_string(maxmemory_) + "\r\n# Stats\r\ntotal_commands_processed:" + std::to_string(ncmd_) + "\r\n"; r_bulk(out, s); }
    else if (cmd == "CONFIG" && a.size() >= 3) config(a, out);
    else if (cmd == "COMMAND" || cmd == "CLIENT" || cmd == "SELECT") r_simple(out, "OK");
    else r_err(out, "unknown command or wrong number of arguments for '" + a[0] + "'");
    ++ncmd_;
    return out;
  }

  // ---- RESP listener ------------------------------------------------------------------------------
  int serve(const char* host, int port) {
    if (listen_fd_ >= 0) return port_;
    int fd = ::socket(AF_INET, SOCK_STREAM, 0);
    if (fd < 0) return -1;
    int one = 1;
    setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_port = htons((uint16_t)port);
    if (inet_pton(AF_INET, host && *host ? host : "127.0.0.1", &addr.sin_addr) != 1) { ::close(fd); return -1; }
    if (::bind(fd, (sockaddr*)&addr, sizeof(addr)) != 0 || ::listen(fd, 128) != 0) { ::close(fd); return -1; }
    socklen_t len = sizeof(addr);
    getsockname(fd, (sockaddr*)&addr, &len);
    port_ = ntohs(addr.sin_port);
    listen_fd_ = fd;
    running_ = true;
    acceptor_ = std::thread([this] { accept_loop(); });
    return port_;
  }

  void stop_server() {
    if (listen_fd_ < 0) return;
    running_ = false;
    stopping_ = true;
    ::shutdown(listen_fd_, SHUT_RDWR);
    ::close(listen_fd_);
    listen_fd_ = -1;
    cv_.notify_all();
    if (acceptor_.joinable()) acceptor_.join();
    {
      std::lock_guard<std::mutex> lk(conn_mu_);
      for (int c : conns_) ::shutdown(c, SHUT_RDWR);
    }
    for (int spin = 0; active_conns_.load() > 0 && spin < 5000; ++spin) std::this_thread::sleep_for(std::chrono::milliseconds(1));
    stopping_ = false;
  }

  long long used() { std::lock_guard<std::mutex> lk(mu_); return used_; }

  // ---- batch paths for the in-process serving engine (no RESP encoding, one call per batch) ----------------
  // Read up to `count` new entries of `key` for `group`, copying the bytes of `field` of each entry back to back into
  // `data` (entries are taken while they have the same length as the first one and fit) and their `uri` fields,
  // '\n'-separated, into `uris`.  Blocks up to block_ms for the first entry.  Returns the number of entries.
  long long read_pack(const std::string& key, const std::string& group, long long count, long long block_ms, const std::string& field,
                      char* data, long long data_cap, long long* rec_len, char* uris, long long uris_cap, long long* uris_len) {
    std::unique_lock<std::mutex> lk(mu_);
    ++ncmd_;
    auto take = [&]() -> long long {
      auto it = streams_.find(key);
      if (it == streams_.end()) return 0;
      Stream& s = it->second;
      auto gi = s.groups.find(group);
      if (gi == s.groups.end()) return -1;
      Group& g = gi->second;
      auto first = std::upper_bound(s.entries.begin(), s.entries.end(), g.last_delivered, [](const StreamId& v, const Entry& e) { return v < e.id; });
      long long n = 0, doff = 0, uoff = 0, len = -1;
      for (auto e = first; e != s.entries.end() && n < count; ++e) {
        const std::string* val = nullptr; const std::string* uri = nullptr;
        for (auto& f : e->fields) { if (f.first == field) val = &f.second; else if (f.first == "uri") uri = &f.second; }
        if (!val || !uri) { if (n == 0) { g.last_delivered = e->id; continue; } break; }   // skip a malformed head entry
        if (len < 0) len = (long long)val->size();
        if ((long long)val->size() != len || doff + len > data_cap || uoff + (long long)uri->size() + 1 > uris_cap) break;
        memcpy(data + doff, val->data(), (size_t)len); doff += len;
        memcpy(uris + uoff, uri->data(), uri->size()); uoff += (long long)uri->size(); uris[uoff++] = '\n';
        g.last_delivered = e->id;
        ++n;
      }
      *rec_len = len < 0 ? 0 : len;
      *uris_len = uoff;
      return n;
    };
    long long n = take();
    if (n != 0 || block_ms < 0) return n;
    auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(block_ms == 0 ? 3600 * 1000 : block_ms);
    while (true) {
      if (cv_.wait_until(lk, deadline) == std::cv_status::timeout) return take();
      n = take();
      if (n != 0 || stopping_) return n;
    }
  }

  // HSET <prefix><key_k> <field> <value_k> for k in [0, n); keys / values are blobs with n+1 offsets each.
  long long hset_many(const std::string& prefix, const std::string& field, long long n, const char* keys, const long long* koff, const char* vals,
                      const long long* voff) {
    std::lock_guard<std::mutex> lk(mu_);
    ++ncmd_;
    long long bytes = 0;
    for (long long k = 0; k < n; ++k) bytes += voff[k + 1] - voff[k] + (long long)field.size();
    if (maxmemory_ > 0 && used_ + bytes > maxmemory_) return -1;
    for (long long k = 0; k < n; ++k) {
      std::string name = prefix;
      name.append(keys + koff[k], (size_t)(koff[k + 1] - koff[k]));
      auto& h = hashes_[name];
      auto f = h.find(field);
      std::string v(vals + voff[k], (size_t)(voff[k + 1] - voff[k]));
      if (f == h.end()) { used_ += (long long)(field.size() + v.size()); h.emplace(field, std::move(v)); }
      else { used_ += (long long)v.size() - (long long)f->second.size(); f->second = std::move(v); }
    }
    return n;
  }

  // Results of one batch: HSET <prefix><uri_k> <field> "[[idx,prob],...]" for the n records whose uris are the '\n'-separated
  // blob `uris` (as returned by read_pack); idx / val are row-major [n, k].  The result string is the reference's
  // (PostProcessing.scala:83-92) with shortest round-trip floats, formatted here instead of once per record in Python.
  long long hset_topn(const std::string& prefix, const std::string& field, long long n, const char* uris, long long uris_len, int k,
                      const int* idx, const float* val) {
    std::vector<std::string> vals((size_t)n);
    char num[64];
    for (long long r = 0; r < n; ++r) {                      // format outside the lock
      std::string& v = vals[(size_t)r];
      v.reserve(16 * (size_t)k + 4);
      v.push_back('[');
      for (int j = 0; j < k; ++j) {
        if (j) v.push_back(',');
        v.push_back('[');
        auto r1 = std::to_chars(num, num + sizeof(num), idx[r * k + j]);
        v.append(num, r1.ptr);
        v.push_back(',');
        auto r2 = std::to_chars(num, num + sizeof(num), val[r * k + j]);
        v.append(num, r2.ptr);
        v.push_back(']');
      }
      v.push_back(']');
    }
    std::lock_guard<std::mutex> lk(mu_);
    ++ncmd_;
    long long bytes = 0;
    for (auto& v : vals) bytes += (long long)(v.size() + field.size());
    if (maxmemory_ > 0 && used_ + bytes > maxmemory_) return -1;
    const char* p = uris;
    const char* end = uris + uris_len;
    long long done = 0;
    for (; done < n && p < end; ++done) {
      const char* q = (const char*)memchr(p, '\n', (size_t)(end - p));
      if (!q) q = end;
      std::string name = prefix;
      name.append(p, (size_t)(q - p));
      auto& h = hashes_[name];
      auto f = h.find(field);
      std::string& v = vals[(size_t)done];
      if (f == h.end()) { used_ += (long long)(field.size() + v.size()); h.emplace(field, std::move(v)); }
      else { used_ += (long long)v.size() - (long long)f->second.size(); f->second = std::move(v); }
      p = q + 1;
    }
    return done;
  }

  // XADD key * uri <uri_k> <field> <record_k> [extra fields...] for k in [0, n); records have a fixed length.
  long long xadd_many(const std::string& key, long long n, const char* uris, const long long* uoff, const std::string& field, const char* recs,
                      long long rec_len, const std::vector<std::pair<std::string, std::string>>& extra) {
    std::lock_guard<std::mutex> lk(mu_);
    ++ncmd_;
    Stream& s = streams_[key];
    long long added = 0;
    for (long long k = 0; k < n; ++k) {
      Entry e;
      e.id = next_id(s);
      e.fields.emplace_back("uri", std::string(uris + uoff[k], (size_t)(uoff[k + 1] - uoff[k])));
      e.fields.emplace_back(field, std::string(recs + k * rec_len, (size_t)rec_len));
      size_t bytes = 64 + e.fields[0].second.size() + (size_t)rec_len + field.size();
      for (auto& x : extra) { e.fields.push_back(x); bytes += x.first.size() + x.second.size() + 16; }
      if (maxmemory_ > 0 && used_ + (long long)bytes > maxmemory_) break;
      e.bytes = bytes;
      used_ += (long long)bytes;
      s.last_id = e.id;
      s.entries.push_back(std::move(e));
      ++added;
    }
    if (added) cv_.notify_all();
    return added;
  }

 private:
  StreamId next_id(Stream& s) {
    using namespace std::chrono;
    uint64_t now = (uint64_t)duration_cast<milliseconds>(system_clock::now().time_since_epoch()).count();
    StreamId id;
    if (now > s.last_id.ms) { id.ms = now; id.seq = 0; } else { id.ms = s.last_id.ms; id.seq = s.last_id.seq + 1; }
    return id;
  }

  void xadd(const std::vector<std::string>& a, std::string& out) {
    // XADD key [MAXLEN [~|=] n] id field value [field value ...]
    size_t i = 2;
    long long maxlen = -1;
    if (a.size() > i && upper(a[i]) == "MAXLEN") {
      ++i;
      if (a.size() > i && (a[i] == "~" || a[i] == "=")) ++i;
      if (a.size() <= i) { r_err(out, "syntax error"); return; }
      maxlen = atoll(a[i].c_str());
      ++i;
    }
    if (a.size() < i + 3 || (a.size() - i - 1) % 2 != 0) { r_err(out, "wrong number of arguments for 'xadd'"); return; }
    Stream& s = streams_[a[1]];
    Entry e;
    if (a[i] == "*") e.id = next_id(s);
    else if (!parse_id(a[i], &e.id) || !(s.last_id < e.id)) { r_err(out, "The ID specified in XADD is equal or smaller than the target stream top item"); return; }
    ++i;
    size_t bytes = 32;
    for (; i + 1 < a.size(); i += 2) {
      // packed batch reads join the uris of a batch with '\n' (read_pack / hset_topn): a uri may not contain one
      if (a[i] == "uri" && a[i + 1].find('\n') != std::string::npos) { r_err(out, "the uri field must not contain a newline"); return; }
      e.fields.emplace_back(a[i], a[i + 1]); bytes += a[i].size() + a[i + 1].size() + 16;
    }
    if (maxmemory_ > 0 && used_ + (long long)bytes > maxmemory_) { out += "-OOM command not allowed when used memory > 'maxmemory'.\r\n"; return; }
    e.bytes = bytes;
    used_ += bytes;
    s.last_id = e.id;
    s.entries.push_back(std::move(e));
    if (maxlen >= 0) trim(s, (size_t)maxlen);
    r_bulk(out, s.last_id.str());
    cv_.notify_all();
  }

  long long trim(Stream& s, size_t maxlen) {
    long long n = 0;
    while (s.entries.size() > maxlen) { used_ -= s.entries.front().bytes; s.entries.pop_front(); ++n; }
    return n;
  }

  void xtrim(const std::vector<std::string>& a, std::string& out) {
    size_t i = 2;
    if (upper(a[i]) != "MAXLEN") { r_err(out, "syntax error"); return; }
    ++i;
    if (a[i] == "~" || a[i] == "=") ++i;
    if (i >= a.size()) { r_err(out, "syntax error"); return; }
    auto it = streams_.find(a[1]);
    r_int(out, it == streams_.end() ? 0 : trim(it->second, (size_t)atoll(a[i].c_str())));
  }

  void xdel(const std::vector<std::string>& a, std::string& out) {
    long long n = 0;
    auto it = streams_.find(a[1]);
    if (it != streams_.end()) {
      for (size_t i = 2; i < a.size(); ++i) {
        StreamId id;
        if (!parse_id(a[i], &id)) continue;
        auto& es = it->second.entries;
        for (auto e = es.begin(); e != es.end(); ++e) if (e->id == id) { used_ -= e->bytes; es.erase(e); ++n; break; }
      }
    }
    r_int(out, n);
  }

  void xrange(const std::vector<std::string>& a, std::string& out) {
    auto it = streams_.find(a[1]);
    StreamId lo, hi;
    if (a[2] == "-") lo = StreamId{0, 0}; else parse_id(a[2], &lo);
    if (a[3] == "+") hi = StreamId{UINT64_MAX, UINT64_MAX}; else { parse_id(a[3], &hi); if (a[3].find('-') == std::string::npos) hi.seq = UINT64_MAX; }
    long long count = -1;
    if (a.size() >= 6 && upper(a[4]) == "COUNT") count = atoll(a[5].c_str());
    std::vector<const Entry*> sel;
    if (it != streams_.end())
      for (auto& e : it->second.entries) {
        if (e.id < lo) continue;
        if (hi < e.id) break;
        sel.push_back(&e);
        if (count >= 0 && (long long)sel.size() >= count) break;
      }
    r_arr(out, sel.size());
    for (auto* e : sel) r_entry(out, *e);
  }

  void xgroup(const std::vector<std::string>& a, std::string& out) {
    if (a.size() >= 5 && upper(a[1]) == "CREATE") {
      bool mk = a.size() >= 6 && upper(a[5]) == "MKSTREAM";
      auto it = streams_.find(a[2]);
      if (it == streams_.end()) {
        if (!mk) { r_err(out, "The XGROUP subcommand requires the key to exist. Note that for CREATE you may want to use the MKSTREAM option to create an empty stream automatically."); return; }
        it = streams_.emplace(a[2], Stream{}).first;
      }
      if (it->second.groups.count(a[3])) { out += "-BUSYGROUP Consumer Group name already exists\r\n"; return; }
      Group g;
      if (a[4] == "$") g.last_delivered = it->second.last_id; else parse_id(a[4], &g.last_delivered);
      it->second.groups[a[3]] = g;
      r_simple(out, "OK");
    } else if (a.size() >= 4 && upper(a[1]) == "DESTROY") {
      auto it = streams_.find(a[2]);
      r_int(out, it != streams_.end() ? (long long)it->second.groups.erase(a[3]) : 0);
    } else r_err(out, "unsupported XGROUP subcommand");
  }

  void xack(const std::vector<std::string>& a, std::string& out) {
    long long n = 0;
    auto it = streams_.find(a[1]);
    if (it != streams_.end()) {
      auto g = it->second.groups.find(a[2]);
      if (g != it->second.groups.end())
        for (size_t i = 3; i < a.size(); ++i) { StreamId id; if (parse_id(a[i], &id)) n += (long long)g->second.pending.erase({id.ms, id.seq}); }
    }
    r_int(out, n);
  }

  void hset(const std::vector<std::string>& a, std::string& out) {
    auto& h = hashes_[a[1]];
    long long added = 0;
    for (size_t i = 2; i + 1 < a.size(); i += 2) {
      auto f = h.find(a[i]);
      if (f == h.end()) { h[a[i]] = a[i + 1]; used_ += a[i].size() + a[i + 1].size(); ++added; }
      else { used_ += (long long)a[i + 1].size() - (long long)f->second.size(); f->second = a[i + 1]; }
    }
    r_int(out, added);
  }

  void keys(const std::string& pat, std::string& out) {
    std::vector<const std::string*> ks;
    for (auto& kv : streams_) if (glob_match(pat.c_str(), kv.first.c_str())) ks.push_back(&kv.first);
    for (auto& kv : hashes_) if (glob_match(pat.c_str(), kv.first.c_str())) ks.push_back(&kv.first);
    for (auto& kv : strings_) if (glob_match(pat.c_str(), kv.first.c_str())) ks.push_back(&kv.first);
    r_arr(out, ks.size());
    for (auto* k : ks) r_bulk(out, *k);
  }

  long long del_key(const std::string& k) {
    long long n = 0;
    auto s = streams_.find(k);
    if (s != streams_.end()) { for (auto& e : s->second.entries) used_ -= e.bytes; streams_.erase(s); ++n; }
    auto h = hashes_.find(k);
    if (h != hashes_.end()) { for (auto& kv : h->second) used_ -= kv.first.size() + kv.second.size(); hashes_.erase(h); ++n; }
    auto t = strings_.find(k);
    if (t != strings_.end()) { used_ -= t->second.size(); strings_.erase(t); ++n; }
    return n;
  }

  void config(const std::vector<std::string>& a, std::string& out) {
    const std::string sub = upper(a[1]);
    if (sub == "SET" && a.size() == 4 && upper(a[2]) == "MAXMEMORY") {
      std::string v = a[3];
      long long mul = 1;
      if (!v.empty()) {
        std::string u = upper(v);
        auto ends = [&](const char* s) { size_t n = strlen(s); return u.size() > n && u.compare(u.size() - n, n, s) == 0; };
        if (ends("GB")) { mul = 1LL << 30; v.resize(v.size() - 2); } else if (ends("MB")) { mul = 1LL << 20; v.resize(v.size() - 2); }
        else if (ends("KB")) { mul = 1LL << 10; v.resize(v.size() - 2); } else if (ends("G")) { mul = 1LL << 30; v.resize(v.size() - 1); }
        else if (ends("M")) { mul = 1LL << 20; v.resize(v.size() - 1); } else if (ends("K")) { mul = 1LL << 10; v.resize(v.size() - 1); }
      }
      maxmemory_ = atoll(v.c_str()) * mul;
      r_simple(out, "OK");
    } else if (sub == "GET" && a.size() == 3 && upper(a[2]) == "MAXMEMORY") {
      r_arr(out, 2); r_bulk(out, "maxmemory"); r_bulk(out, std::to_string(maxmemory_));
    } else if (sub == "SET") r_simple(out, "OK");
    else r_arr(out, 0);
  }

  // XREAD [COUNT n] [BLOCK ms] STREAMS key id   /   XREADGROUP GROUP g c [COUNT n] [BLOCK ms] [NOACK] STREAMS key >
  std::string xread(const std::vector<std::string>& a, bool grouped) {
    std::string out;
    size_t i = 1;
    std::string group, consumer;
    if (grouped) {
      if (a.size() < 7 || upper(a[1]) != "GROUP") { r_err(out, "syntax error"); return out; }
      group = a[2]; consumer = a[3]; i = 4;
    }
    long long count = -1, block = -1;
    bool noack = false;
    for (; i < a.size(); ++i) {
      std::string u = upper(a[i]);
      if (u == "COUNT" && i + 1 < a.size()) count = atoll(a[++i].c_str());
      else if (u == "BLOCK" && i + 1 < a.size()) block = atoll(a[++i].c_str());
      else if (u == "NOACK") noack = true;
      else if (u == "STREAMS") { ++i; break; }
      else { r_err(out, "syntax error"); return out; }
    }
    const size_t nk = (a.size() - i) / 2;
    if (nk == 0 || (a.size() - i) % 2 != 0) { r_err(out, "Unbalanced XREAD list of streams: for each stream key an ID or '$' must be specified."); return out; }
    std::unique_lock<std::mutex> lk(mu_);
    ++ncmd_;
    std::vector<StreamId> from(nk);
    for (size_t k = 0; k < nk; ++k) {
      const std::string& key = a[i + k], &ids = a[i + nk + k];
      auto it = streams_.find(key);
      if (grouped) {
        if (it == streams_.end() || !it->second.groups.count(group)) { out += "-NOGROUP No such key '" + key + "' or consumer group '" + group + "'\r\n"; return out; }
        if (ids != ">") parse_id(ids, &from[k]);
      } else if (ids == "$") from[k] = it == streams_.end() ? StreamId{} : it->second.last_id;
      else parse_id(ids, &from[k]);
    }
    auto collect = [&](std::string& o) -> size_t {
      size_t total = 0;
      std::string body;
      size_t nstreams = 0;
      for (size_t k = 0; k < nk; ++k) {
        auto it = streams_.find(a[i + k]);
        if (it == streams_.end()) continue;
        Stream& s = it->second;
        Group* g = grouped ? &s.groups[group] : nullptr;
        const bool fresh = grouped && a[i + nk + k] == ">";
        StreamId after = fresh ? g->last_delivered : from[k];
        // entries are id-ordered: binary search the first one > after
        auto first = std::upper_bound(s.entries.begin(), s.entries.end(), after, [](const StreamId& v, const Entry& e) { return v < e.id; });
        std::string eb;
        size_t n = 0;
        for (auto e = first; e != s.entries.end(); ++e) {
          if (count >= 0 && (long long)n >= count) break;
          r_entry(eb, *e);
          ++n;
          if (fresh) { g->last_delivered = e->id; if (!noack) g->pending.insert({e->id.ms, e->id.seq}); }
        }
        if (n) { ++nstreams; r_arr(body, 2); r_bulk(body, a[i + k]); r_arr(body, n); body += eb; total += n; }
      }
      if (total) { r_arr(o, nstreams); o += body; }
      return total;
    };
    if (collect(out)) return out;
    if (block < 0) { r_nullarr(out); return out; }
    auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(block == 0 ? 3600 * 1000 : block);
    while (true) {
      if (cv_.wait_until(lk, deadline) == std::cv_status::timeout) { if (!collect(out)) r_nullarr(out); return out; }
      if (collect(out)) return out;
      if (stopping_) { r_nullarr(out); return out; }
    }
  }

  void accept_loop() {
    while (running_) {
      int c = ::accept(listen_fd_, nullptr, nullptr);
      if (c < 0) { if (!running_) break; continue; }
      int one = 1;
      setsockopt(c, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      std::lock_guard<std::mutex> lk(conn_mu_);
      conns_.insert(c);
      ++active_conns_;
      std::thread([this, c] { conn_loop(c); --active_conns_; }).detach();      // finished connections leave nothing behind
    }
  }

  // Limits of the socket protocol: a client that exceeds them (or sends something that is not RESP) gets -ERR and is disconnected,
  // instead of growing the connection buffer without bound.
  static constexpr long long kMaxBulk = 512LL << 20;      // one argument (a base64 image batch is a few MB)
  static constexpr long long kMaxArgs = 1 << 20;          // arguments per command
  static constexpr size_t kMaxInline = 64 << 10;          // an inline (non-array) command line
  static constexpr size_t kMaxBuffered = 1ULL << 30;      // unparsed bytes per connection

  // 1 = one command parsed (pos advanced), 0 = need more bytes, -1 = protocol error
  static int parse_one(const std::string& buf, size_t& pos, std::vector<std::string>& args) {
    args.clear();
    if (pos >= buf.size()) return 0;
    size_t p = pos;
    bool too_long = false;
    auto line = [&](std::string& l, size_t limit) -> bool {
      size_t e = buf.find("\r\n", p);
      if (e == std::string::npos) { if (buf.size() - p > limit) too_long = true; return false; }
      if (e - p > limit) { too_long = true; return false; }
      l = buf.substr(p, e - p); p = e + 2; return true;
    };
    auto number = [](const std::string& l, long long& out) -> bool {
      if (l.size() < 2 || l.size() > 20) return false;
      size_t i = 1;
      if (l[i] == '-') ++i;
      if (i >= l.size()) return false;
      for (size_t k = i; k < l.size(); ++k) if (l[k] < '0' || l[k] > '9') return false;
      out = atoll(l.c_str() + 1);
      return true;
    };
    if (buf[p] != '*') {
      std::string l;
      if (!line(l, kMaxInline)) return too_long ? -1 : 0;
      size_t s = 0;
      while (s < l.size()) { size_t e = l.find(' ', s); if (e == std::string::npos) e = l.size(); if (e > s) args.push_back(l.substr(s, e - s)); s = e + 1; }
      pos = p;
      return 1;
    }
    std::string l;
    if (!line(l, 32)) return too_long ? -1 : 0;
    long long n = 0;
    if (!number(l, n) || n < 0 || n > kMaxArgs) return -1;
    for (long long k = 0; k < n; ++k) {
      if (!line(l, 32)) return too_long ? -1 : 0;
      long long len = 0;
      if (l.empty() || l[0] != '$' || !number(l, len) || len > kMaxBulk) return -1;
      if (len < 0) { args.emplace_back(); continue; }
      if (p + (size_t)len + 2 > buf.size()) return 0;
      if (buf[p + (size_t)len] != '\r' || buf[p + (size_t)len + 1] != '\n') return -1;
      args.emplace_back(buf, p, (size_t)len);
      p += (size_t)len + 2;
    }
    pos = p;
    return 1;
  }

  void conn_loop(int fd) {
    std::string buf;
    std::vector<std::string> args;
    char tmp[1 << 16];
    while (running_) {
      ssize_t n = ::recv(fd, tmp, sizeof(tmp), 0);
      if (n <= 0) break;
      buf.append(tmp, (size_t)n);
      size_t pos = 0;
      std::string reply;
      int st;
      while ((st = parse_one(buf, pos, args)) == 1) {
        if (args.empty()) continue;
        if (upper(args[0]) == "QUIT") { reply += "+OK\r\n"; running_conn_quit(fd, reply); goto done; }
        reply += execute(args);
      }
      buf.erase(0, pos);
      if (st < 0 || buf.size() > kMaxBuffered) {
        reply += "-ERR Protocol error: malformed or oversized request\r\n";
        send_all(fd, reply);
        break;
      }
      if (!reply.empty() && !send_all(fd, reply)) break;
    }
  done:
    ::close(fd);
    std::lock_guard<std::mutex> lk(conn_mu_);
    conns_.erase(fd);
  }

  static bool send_all(int fd, const std::string& s) {
    size_t off = 0;
    while (off < s.size()) { ssize_t n = ::send(fd, s.data() + off, s.size() - off, MSG_NOSIGNAL); if (n <= 0) return false; off += (size_t)n; }
    return true;
  }
  static void running_conn_quit(int fd, const std::string& reply) { send_all(fd, reply); }

  std::mutex mu_;
  std::condition_variable cv_;
  std::map<std::string, Stream> streams_;
  std::map<std::string, std::map<std::string, std::string>> hashes_;
  std::map<std::string, std::string> strings_;
  long long used_ = 0, maxmemory_ = 0, ncmd_ = 0;
  std::atomic<bool> running_{false}, stopping_{false};
  int listen_fd_ = -1, port_ = 0;
  std::thread acceptor_;
  std::mutex conn_mu_;
  std::set<int> conns_;
  std::atomic<int> active_conns_{0};
};

}  // namespace

ZOO_RT_API void* zoo_broker_create(long long maxmemory) { return new Broker(maxmemory); }
ZOO_RT_API void zoo_broker_destroy(void* b) { delete static_cast<Broker*>(b); }
ZOO_RT_API int zoo_broker_serve(void* b, const char* host, int port) { return static_cast<Broker*>(b)->serve(host, port); }
ZOO_RT_API void zoo_broker_stop_server(void* b) { static_cast<Broker*>(b)->stop_server(); }
ZOO_RT_API long long zoo_broker_used_memory(void* b) { return static_cast<Broker*>(b)->used(); }

// Execute one command in-process.  argv/arglen: argc byte strings.  The RESP2 reply is returned in a
// malloc'ed buffer (*out, *outlen) the caller releases with zoo_broker_free.
ZOO_RT_API int zoo_broker_command(void* b, int argc, const char* const* argv, const long long* arglen, char** out, long long* outlen) {
  std::vector<std::string> a;
  a.reserve((size_t)argc);
  for (int i = 0; i < argc; ++i) a.emplace_back(argv[i], (size_t)arglen[i]);
  std::string r = static_cast<Broker*>(b)->execute(a);
  char* buf = (char*)malloc(r.size() + 1);
  if (!buf) return 1;
  memcpy(buf, r.data(), r.size());
  buf[r.size()] = 0;
  *out = buf;
  *outlen = (long long)r.size();
  return 0;
}
ZOO_RT_API void zoo_broker_free(char* p) { free(p); }

ZOO_RT_API long long zoo_broker_read_pack(void* b, const char* key, const char* group, long long count, long long block_ms, const char* field,
                                          char* data, long long data_cap, long long* rec_len, char* uris, long long uris_cap, long long* uris_len) {
  return static_cast<Broker*>(b)->read_pack(key, group, count, block_ms, field, data, data_cap, rec_len, uris, uris_cap, uris_len);
}
ZOO_RT_API long long zoo_broker_hset_many(void* b, const char* prefix, const char* field, long long n, const char* keys, const long long* koff,
                                          const char* vals, const long long* voff) {
  return static_cast<Broker*>(b)->hset_many(prefix, field, n, keys, koff, vals, voff);
}
ZOO_RT_API long long zoo_broker_hset_topn(void* b, const char* prefix, const char* field, long long n, const char* uris, long long uris_len,
                                          int k, const int* idx, const float* val) {
  return static_cast<Broker*>(b)->hset_topn(prefix, field, n, uris, uris_len, k, idx, val);
}
ZOO_RT_API long long zoo_broker_xadd_many(void* b, const char* key, long long n, const char* uris, const long long* uoff, const char* field,
                                          const char* recs, long long rec_len, int n_extra, const char* const* extra_names,
                                          const char* const* extra_vals) {
  std::vector<std::pair<std::string, std::string>> extra;
  for (int i = 0; i < n_extra; ++i) extra.emplace_back(extra_names[i], extra_vals[i]);
  return static_cast<Broker*>(b)->xadd_many(key, n, uris, uoff, field, recs, rec_len, extra);
}
