// Bulk TCP data plane (C++): what the reference moves through the go-libp2p daemon's stream mux.
// Length-prefixed frames over blocking sockets with TCP_NODELAY and large kernel buffers, plus an epoll-based
// acceptor thread that hands established connections to the caller.  The Python side (transport/bulk.py) calls
// these with the GIL released; tensor parts are sent straight from (pinned) tensor memory without copies.
#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/epoll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>

#include "b200_host_api.h"

namespace {

void tune(int fd) {
  int one = 1, buf = 8 << 20;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  setsockopt(fd, SOL_SOCKET, SO_SNDBUF, &buf, sizeof(buf));
  setsockopt(fd, SOL_SOCKET, SO_RCVBUF, &buf, sizeof(buf));
}

struct Listener {
  int fd = -1, ep = -1, port = 0;
  std::thread th;
  std::atomic<bool> stop{false};
  std::mutex mu;
  std::condition_variable cv;
  std::deque<int> ready;

  void run() {
    epoll_event ev{};
    while (!stop.load()) {
      const int n = epoll_wait(ep, &ev, 1, 100);
      if (n <= 0) continue;
      for (;;) {
        const int c = accept(fd, nullptr, nullptr);
        if (c < 0) break;
        tune(c);
        {
          std::lock_guard<std::mutex> lk(mu);
          ready.push_back(c);
        }
        cv.notify_one();
      }
    }
  }
};

long long fail_after = -1;   // fault injection: DALLE_B200_FAIL_DROP_AFTER_BYTES
std::atomic<long long> sent_total{0};

bool send_all(int fd, const void* p, long long n) {
  const char* c = static_cast<const char*>(p);
  while (n > 0) {
    if (fail_after >= 0 && sent_total.load() > fail_after) { shutdown(fd, SHUT_RDWR); return false; }
    const ssize_t k = send(fd, c, static_cast<size_t>(n > (1 << 24) ? (1 << 24) : n), MSG_NOSIGNAL);
    if (k <= 0) { if (errno == EINTR) continue; return false; }
    c += k; n -= k; sent_total += k;
  }
  return true;
}

bool recv_all(int fd, void* p, long long n, int timeout_ms) {
  char* c = static_cast<char*>(p);
  while (n > 0) {
    if (timeout_ms >= 0) {
      pollfd pf{fd, POLLIN, 0};
      const int r = poll(&pf, 1, timeout_ms);
      if (r <= 0) return false;
    }
    const ssize_t k = recv(fd, c, static_cast<size_t>(n > (1 << 24) ? (1 << 24) : n), 0);
    if (k <= 0) { if (k < 0 && errno == EINTR) continue; return false; }
    c += k; n -= k;
  }
  return true;
}

}  // namespace

extern "C" void* b200h_listen(const char* host, int port, int* out_port) {
  if (const char* e = getenv("DALLE_B200_FAIL_DROP_AFTER_BYTES")) fail_after = atoll(e);
  auto* l = new Listener();
  l->fd = socket(AF_INET, SOCK_STREAM | SOCK_NONBLOCK, 0);
  int one = 1;
  setsockopt(l->fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in a{};
  a.sin_family = AF_INET;
  a.sin_port = htons(static_cast<uint16_t>(port));
  a.sin_addr.s_addr = (host && *host) ? inet_addr(host) : INADDR_ANY;
  if (bind(l->fd, reinterpret_cast<sockaddr*>(&a), sizeof(a)) != 0 || listen(l->fd, 128) != 0) {
    close(l->fd);
    delete l;
    return nullptr;
  }
  socklen_t len = sizeof(a);
  getsockname(l->fd, reinterpret_cast<sockaddr*>(&a), &len);
  l->port = ntohs(a.sin_port);
  if (out_port) *out_port = l->port;
  l->ep = epoll_create1(0);
  epoll_event ev{};
  ev.events = EPOLLIN;
  ev.data.fd = l->fd;
  epoll_ctl(l->ep, EPOLL_CTL_ADD, l->fd, &ev);
  l->th = std::thread([l] { l->run(); });
  return l;
}

extern "C" int b200h_accept(void* listener, int timeout_ms) {
  auto* l = static_cast<Listener*>(listener);
  std::unique_lock<std::mutex> lk(l->mu);
  if (!l->cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), [l] { return !l->ready.empty(); })) return -1;
  const int c = l->ready.front();
  l->ready.pop_front();
  return c;
}

extern "C" void b200h_close_listener(void* listener) {
  auto* l = static_cast<Listener*>(listener);
  l->stop.store(true);
  if (l->th.joinable()) l->th.join();
  for (int c : l->ready) close(c);
  close(l->ep);
  close(l->fd);
  delete l;
}

extern "C" int b200h_connect(const char* host, int port, int timeout_ms) {
  if (const char* e = getenv("DALLE_B200_FAIL_DROP_AFTER_BYTES")) fail_after = atoll(e);
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  char ps[16];
  snprintf(ps, sizeof(ps), "%d", port);
  if (getaddrinfo(host, ps, &hints, &res) != 0 || !res) return -1;
  const int fd = socket(res->ai_family, res->ai_socktype, res->ai_protocol);
  fcntl(fd, F_SETFL, fcntl(fd, F_GETFL) | O_NONBLOCK);
  int rc = connect(fd, res->ai_addr, res->ai_addrlen);
  freeaddrinfo(res);
  if (rc != 0 && errno == EINPROGRESS) {
    pollfd pf{fd, POLLOUT, 0};
    if (poll(&pf, 1, timeout_ms) <= 0) { close(fd); return -1; }
    int err = 0;
    socklen_t len = sizeof(err);
    getsockopt(fd, SOL_SOCKET, SO_ERROR, &err, &len);
    if (err != 0) { close(fd); return -1; }
  } else if (rc != 0) {
    close(fd);
    return -1;
  }
  fcntl(fd, F_SETFL, fcntl(fd, F_GETFL) & ~O_NONBLOCK);
  tune(fd);
  return fd;
}

// frame = u64 tag | u64 nbytes | payload
extern "C" int b200h_send_frame(int fd, unsigned long long tag, const void* data, long long nbytes) {
  uint64_t hdr[2] = {tag, static_cast<uint64_t>(nbytes)};
  return (send_all(fd, hdr, sizeof(hdr)) && send_all(fd, data, nbytes)) ? 0 : -1;
}

extern "C" long long b200h_recv_header(int fd, unsigned long long* tag, int timeout_ms) {
  uint64_t hdr[2];
  if (!recv_all(fd, hdr, sizeof(hdr), timeout_ms)) return -1;
  *tag = hdr[0];
  return static_cast<long long>(hdr[1]);
}

extern "C" int b200h_recv_payload(int fd, void* data, long long nbytes, int timeout_ms) {
  return recv_all(fd, data, nbytes, timeout_ms) ? 0 : -1;
}

extern "C" void b200h_close(int fd) { close(fd); }
