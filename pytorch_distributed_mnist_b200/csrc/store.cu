// store.cu -- a small TCP key-value rendezvous store (host code only).
//
// Role: the c10d TCPStore behind the reference's `tcp://` init method (ref:167-168,
// ref:323-327) as far as this framework needs it -- exchanging the 64-byte CUDA IPC
// handles of the peer-memory communicator and counting arrivals.  Rank 0 hosts a poll()-driven
// server thread; every rank (rank 0 included) is a client.
//
// Wire format, little-endian, one request per message:
//   u8 op | u32 key_len | key bytes | u64 val_len | val bytes
//   op 1 SET   -> reply u8 status
//   op 2 GET   -> reply u8 status | u64 len | bytes   (parked until the key exists)
//   op 3 ADD   -> val = i64 delta; reply u8 status | i64 new value (missing key counts as 0)
// A GET for a missing key is not answered until a SET/ADD creates it, which is what makes
// `get` a rendezvous; the client enforces the timeout on its side.
#include "common.cuh"

#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <map>
#include <new>
#include <string>
#include <thread>
#include <vector>

namespace {

enum { OP_SET = 1, OP_GET = 2, OP_ADD = 3 };

bool send_all(int fd, const void* buf, size_t n) {
  const char* p = (const char*)buf;
  while (n > 0) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    p += k;
    n -= (size_t)k;
  }
  return true;
}

// Blocking receive with an absolute deadline (ms since steady epoch; <0 = none).
bool recv_all(int fd, void* buf, size_t n, long long deadline_ms) {
  char* p = (char*)buf;
  while (n > 0) {
    if (deadline_ms >= 0) {
      const long long now = std::chrono::duration_cast<std::chrono::milliseconds>(
                                std::chrono::steady_clock::now().time_since_epoch()).count();
      const long long left = deadline_ms - now;
      if (left <= 0) { errno = ETIMEDOUT; return false; }
      struct pollfd pf = {fd, POLLIN, 0};
      int r = ::poll(&pf, 1, (int)std::min<long long>(left, 1000));
      if (r < 0 && errno != EINTR) return false;
      if (r <= 0) continue;
    }
    ssize_t k = ::recv(fd, p, n, 0);
    if (k == 0) { errno = ECONNRESET; return false; }
    if (k < 0) {
      if (errno == EINTR || errno == EAGAIN) continue;
      return false;
    }
    p += k;
    n -= (size_t)k;
  }
  return true;
}

struct Server {
  int listen_fd = -1;
  std::thread thread;
  std::atomic<bool> stop{false};
  std::map<std::string, std::string> kv;
  std::map<std::string, std::vector<int>> waiters;   // key -> fds parked in GET
  std::vector<int> clients;

  void reply_get(int fd, const std::string& val) {
    uint8_t ok = 0;
    uint64_t len = val.size();
    send_all(fd, &ok, 1);
    send_all(fd, &len, 8);
    if (len) send_all(fd, val.data(), len);
  }
  void wake(const std::string& key) {
    auto it = waiters.find(key);
    if (it == waiters.end()) return;
    for (int fd : it->second) reply_get(fd, kv[key]);
    waiters.erase(it);
  }
  void drop(int fd) {
    for (auto& w : waiters)
      for (size_t i = 0; i < w.second.size();)
        if (w.second[i] == fd) w.second.erase(w.second.begin() + i); else ++i;
    for (size_t i = 0; i < clients.size(); ++i)
      if (clients[i] == fd) { clients.erase(clients.begin() + i); break; }
    ::close(fd);
  }
  bool handle(int fd) {
    uint8_t op;
    uint32_t klen;
    uint64_t vlen;
    const long long dl = std::chrono::duration_cast<std::chrono::milliseconds>(
                             std::chrono::steady_clock::now().time_since_epoch()).count() + 10000;
    if (!recv_all(fd, &op, 1, dl) || !recv_all(fd, &klen, 4, dl) || klen > (1u << 16)) return false;
    std::string key(klen, '\0');
    if (klen && !recv_all(fd, &key[0], klen, dl)) return false;
    if (!recv_all(fd, &vlen, 8, dl) || vlen > (1ull << 26)) return false;
    std::string val((size_t)vlen, '\0');
    if (vlen && !recv_all(fd, &val[0], (size_t)vlen, dl)) return false;
    uint8_t ok = 0;
    if (op == OP_SET) {
      kv[key] = val;
      send_all(fd, &ok, 1);
      wake(key);
    } else if (op == OP_GET) {
      auto it = kv.find(key);
      if (it != kv.end()) reply_get(fd, it->second);
      else waiters[key].push_back(fd);
    } else if (op == OP_ADD) {
      int64_t cur = 0, delta = 0;
      auto it = kv.find(key);
      if (it != kv.end() && it->second.size() == 8) memcpy(&cur, it->second.data(), 8);
      if (val.size() == 8) memcpy(&delta, val.data(), 8);
      cur += delta;
      kv[key] = std::string((const char*)&cur, 8);
      send_all(fd, &ok, 1);
      send_all(fd, &cur, 8);
      wake(key);
    } else {
      return false;
    }
    return true;
  }
  void loop() {
    while (!stop.load()) {
      std::vector<struct pollfd> pfs;
      pfs.push_back({listen_fd, POLLIN, 0});
      for (int fd : clients) pfs.push_back({fd, POLLIN, 0});
      int r = ::poll(pfs.data(), (nfds_t)pfs.size(), 100);
      if (r <= 0) continue;
      if (pfs[0].revents & POLLIN) {
        int c = ::accept(listen_fd, nullptr, nullptr);
        if (c >= 0) {
          int one = 1;
          setsockopt(c, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
          clients.push_back(c);
        }
      }
      for (size_t i = 1; i < pfs.size(); ++i) {
        if (pfs[i].revents & (POLLIN | POLLHUP | POLLERR)) {
          if (!handle(pfs[i].fd)) drop(pfs[i].fd);
        }
      }
    }
    for (int fd : clients) ::close(fd);
    clients.clear();
    ::close(listen_fd);
  }
};

long long now_ms() {
  return std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

struct pdm_store {
  Server* server = nullptr;
  int fd = -1;
  double timeout_s = 300.0;
};

extern "C" {

int pdm_store_destroy(pdm_store* s);

// is_server != 0: also host the store on `port` (rank 0).  Every rank connects as a client,
// retrying until the server is up or `timeout_s` passes.
int pdm_store_create(pdm_store** out, const char* host, int port, int is_server, double timeout_s) {
  PDM_REQUIRE(out && host && port > 0 && port < 65536, PDM_ERR_ARG, "pdm_store_create: bad arguments");
  pdm_store* s = new (std::nothrow) pdm_store();
  PDM_REQUIRE(s != nullptr, PDM_ERR_STATE, "pdm_store_create: out of memory");
  s->timeout_s = timeout_s > 0 ? timeout_s : 300.0;
  if (is_server) {
    int lf = ::socket(AF_INET, SOCK_STREAM, 0);
    int one = 1;
    setsockopt(lf, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    struct sockaddr_in addr;
    memset(&addr, 0, sizeof(addr));
    addr.sin_family = AF_INET;
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    addr.sin_port = htons((uint16_t)port);
    if (lf < 0 || ::bind(lf, (struct sockaddr*)&addr, sizeof(addr)) != 0 || ::listen(lf, 64) != 0) {
      pdm_set_error("pdm_store_create: cannot listen on port %d: %s", port, strerror(errno));
      if (lf >= 0) ::close(lf);
      delete s;
      return PDM_ERR_IO;
    }
    s->server = new Server();
    s->server->listen_fd = lf;
    s->server->thread = std::thread([srv = s->server] { srv->loop(); });
  }
  struct addrinfo hints, *res = nullptr;
  memset(&hints, 0, sizeof(hints));
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  char portstr[16];
  snprintf(portstr, sizeof(portstr), "%d", port);
  if (getaddrinfo(host, portstr, &hints, &res) != 0 || !res) {
    pdm_set_error("pdm_store_create: cannot resolve %s", host);
    pdm_store_destroy(s);
    return PDM_ERR_IO;
  }
  const long long deadline = now_ms() + (long long)(s->timeout_s * 1000);
  for (;;) {
    int fd = ::socket(AF_INET, SOCK_STREAM, 0);
    if (fd >= 0 && ::connect(fd, res->ai_addr, res->ai_addrlen) == 0) {
      int one = 1;
      setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      s->fd = fd;
      break;
    }
    if (fd >= 0) ::close(fd);
    if (now_ms() > deadline) {
      freeaddrinfo(res);
      pdm_set_error("pdm_store_create: no store at %s:%d after %.0f s", host, port, s->timeout_s);
      pdm_store_destroy(s);
      return PDM_ERR_TIMEOUT;
    }
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
  freeaddrinfo(res);
  *out = s;
  return PDM_OK;
}

static int request(pdm_store* s, uint8_t op, const char* key, const void* val, uint64_t vlen) {
  const uint32_t klen = (uint32_t)strlen(key);
  if (!send_all(s->fd, &op, 1) || !send_all(s->fd, &klen, 4) || !send_all(s->fd, key, klen) ||
      !send_all(s->fd, &vlen, 8) || (vlen && !send_all(s->fd, val, (size_t)vlen))) {
    pdm_set_error("pdm_store: send failed: %s", strerror(errno));
    return PDM_ERR_IO;
  }
  return PDM_OK;
}

int pdm_store_set(pdm_store* s, const char* key, const void* val, size_t len) {
  PDM_REQUIRE(s && key && (val || len == 0), PDM_ERR_ARG, "pdm_store_set: bad arguments");
  int rc = request(s, OP_SET, key, val, len);
  if (rc != PDM_OK) return rc;
  uint8_t st;
  PDM_REQUIRE(recv_all(s->fd, &st, 1, now_ms() + (long long)(s->timeout_s * 1000)) && st == 0, PDM_ERR_IO,
              "pdm_store_set: no acknowledgement: %s", strerror(errno));
  return PDM_OK;
}

// Blocks until `key` exists (or the store's timeout).  *len receives the value size; at most `cap` bytes are copied.
int pdm_store_get(pdm_store* s, const char* key, void* buf, size_t cap, size_t* len) {
  PDM_REQUIRE(s && key && len, PDM_ERR_ARG, "pdm_store_get: bad arguments");
  int rc = request(s, OP_GET, key, nullptr, 0);
  if (rc != PDM_OK) return rc;
  const long long dl = now_ms() + (long long)(s->timeout_s * 1000);
  uint8_t st;
  uint64_t n;
  if (!recv_all(s->fd, &st, 1, dl) || !recv_all(s->fd, &n, 8, dl)) {
    const bool to = errno == ETIMEDOUT;
    pdm_set_error("pdm_store_get('%s'): %s", key, to ? "timed out waiting for a peer to publish it" : strerror(errno));
    return to ? PDM_ERR_TIMEOUT : PDM_ERR_IO;
  }
  std::string tmp((size_t)n, '\0');
  PDM_REQUIRE(n == 0 || recv_all(s->fd, &tmp[0], (size_t)n, dl), PDM_ERR_IO, "pdm_store_get: truncated value");
  *len = (size_t)n;
  if (buf && cap) memcpy(buf, tmp.data(), std::min<size_t>(cap, (size_t)n));
  return PDM_OK;
}

int pdm_store_add(pdm_store* s, const char* key, int64_t delta, int64_t* result) {
  PDM_REQUIRE(s && key, PDM_ERR_ARG, "pdm_store_add: bad arguments");
  int rc = request(s, OP_ADD, key, &delta, 8);
  if (rc != PDM_OK) return rc;
  const long long dl = now_ms() + (long long)(s->timeout_s * 1000);
  uint8_t st;
  int64_t v;
  PDM_REQUIRE(recv_all(s->fd, &st, 1, dl) && recv_all(s->fd, &v, 8, dl), PDM_ERR_IO, "pdm_store_add: no reply: %s",
              strerror(errno));
  if (result) *result = v;
  return PDM_OK;
}

int pdm_store_destroy(pdm_store* s) {
  if (!s) return PDM_OK;
  if (s->fd >= 0) ::close(s->fd);
  if (s->server) {
    s->server->stop.store(true);
    if (s->server->thread.joinable()) s->server->thread.join();
    delete s->server;
  }
  delete s;
  return PDM_OK;
}

}  // extern "C"
