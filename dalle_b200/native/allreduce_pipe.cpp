// Butterfly all-reduce round as a C++ thread pipeline over the bulk TCP transport (include/b200_host_api.h,
// b200h_allreduce_round).  The reference moves every tensor part through the native go-libp2p daemon and
// hivemind's AllReduceRunner / TensorPartReducer coroutines; the first version here did the
// part loop in Python asyncio with msgpack frames.  This file is the data plane without the interpreter:
//
//   per other owner   tx thread: compress part from the flat buffer -> frame -> socket
//                     rx thread: frame -> decode delta -> flat += delta
//   per other sender  rx thread: frame -> decode -> keep the lossy copy -> acc += w * x
//   (own span)        tx thread: wait until the part is complete -> delta = acc / denom - lossy -> compress -> frame
//
// so sockets stay full in both directions while the codecs run on other cores.  The flat buffer is the caller's
// (pinned, when it is the staging buffer of a GPU trainer); nothing is copied except into codec output.
#include <immintrin.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "b200_host_api.h"

namespace {

constexpr uint64_t kHello = 0xB200A11D00000001ull;
constexpr uint64_t kDone = 0xB200A11D000000FFull;
constexpr float kFp16Max = 65504.0f;

struct Hello {
  unsigned char gid[16];
  int32_t sender;
  float weight;
};

// ---- codecs (wire formats of collab/compression.py) -----------------------------------------------------------
void encode(int codec, const float* x, long long n, std::vector<unsigned char>& out) {
  if (codec == 0) {
    out.resize(static_cast<size_t>(n) * 4);
    std::memcpy(out.data(), x, out.size());
  } else if (codec == 1) {
    out.resize(static_cast<size_t>(n) * 2);
    uint16_t* h = reinterpret_cast<uint16_t*>(out.data());
    const __m256 hi = _mm256_set1_ps(kFp16Max), lo = _mm256_set1_ps(-kFp16Max);
    long long i = 0;
    for (; i + 8 <= n; i += 8) {
      __m256 v = _mm256_loadu_ps(x + i);
      v = _mm256_min_ps(_mm256_max_ps(v, lo), hi);     // (NaN stays NaN: the owner refuses the part)
      _mm_storeu_si128(reinterpret_cast<__m128i*>(h + i), _mm256_cvtps_ph(v, _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC));
    }
    for (; i < n; ++i) {
      float v = x[i];
      v = v > kFp16Max ? kFp16Max : (v < -kFp16Max ? -kFp16Max : v);
      h[i] = _cvtss_sh(v, _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC);
    }
  } else {
    out.resize(1024 + static_cast<size_t>(n));
    b200h_uniform8_quantize(x, n, out.data() + 1024, reinterpret_cast<float*>(out.data()));
  }
}

long long encoded_bytes(int codec, long long n) { return codec == 0 ? 4 * n : (codec == 1 ? 2 * n : 1024 + n); }

// returns false on a size mismatch or a non-finite value
bool decode(int codec, const unsigned char* in, long long nbytes, float* out, long long n) {
  if (nbytes != encoded_bytes(codec, n)) return false;
  if (codec == 0) {
    std::memcpy(out, in, static_cast<size_t>(n) * 4);
  } else if (codec == 1) {
    const uint16_t* h = reinterpret_cast<const uint16_t*>(in);
    long long i = 0;
    for (; i + 8 <= n; i += 8)
      _mm256_storeu_ps(out + i, _mm256_cvtph_ps(_mm_loadu_si128(reinterpret_cast<const __m128i*>(h + i))));
    for (; i < n; ++i) out[i] = _cvtsh_ss(h[i]);
  } else {
    b200h_uniform8_dequantize(in + 1024, reinterpret_cast<const float*>(in), out, n);
  }
  float bad = 0.f;
  for (long long i = 0; i < n; ++i) bad += out[i] * 0.0f;   // 0 unless some value is Inf / NaN
  return bad == 0.f;
}

bool recv_frame(int fd, uint64_t* tag, std::vector<unsigned char>& buf, int timeout_ms, long long max_bytes) {
  unsigned long long t = 0;
  const long long n = b200h_recv_header(fd, &t, timeout_ms);
  if (n < 0 || n > max_bytes) return false;
  buf.resize(static_cast<size_t>(n));
  if (n && b200h_recv_payload(fd, buf.data(), n, timeout_ms) != 0) return false;
  *tag = t;
  return true;
}

}  // namespace

namespace {

struct OwnPart {
  long long start = 0, n = 0;
  int codec = 0;
  std::vector<float> acc;
  double denom = 0.0;
  bool done = false;
  std::vector<std::vector<float>> lossy;   // per member index: what that sender's part decoded to
  std::vector<char> have;
};

struct Round {
  const B200hAllreduce* d;
  int p0 = 0, p1 = 0;                      // my span
  std::vector<OwnPart> own;
  std::vector<char> is_sender, active;     // per member
  std::mutex mu;
  std::condition_variable cv;
  std::atomic<int> failures{0};
  long long max_frame = 0;

  explicit Round(const B200hAllreduce* desc) : d(desc) {
    const int n = d->n_members;
    is_sender.assign(n, 0);
    active.assign(n, 0);
    for (int k = 0; k < n; ++k) {
      is_sender[k] = d->modes[k] != 2;
      active[k] = is_sender[k] && k != d->my_index;
    }
    p0 = d->span_begin[d->my_index];
    p1 = d->span_end[d->my_index];
    own.resize(std::max(0, p1 - p0));
    for (int j = 0; j < p1 - p0; ++j) {
      OwnPart& o = own[j];
      o.start = d->part_start[p0 + j];
      o.n = d->part_end[p0 + j] - o.start;
      o.codec = d->part_codec[p0 + j];
      o.acc.assign(static_cast<size_t>(o.n), 0.f);
      o.lossy.resize(n);
      o.have.assign(n, 0);
      if (is_sender[d->my_index] && d->weight > 0.f) {
        b200h_weighted_add(o.acc.data(), d->flat + o.start, d->weight, o.n);
        o.denom = d->weight;
      }
    }
    for (int i = 0; i < d->n_parts; ++i)
      max_frame = std::max(max_frame, 4 * (d->part_end[i] - d->part_start[i]) + 1024);
    std::lock_guard<std::mutex> lk(mu);
    recheck_locked();
  }

  void mark_failed(int member) {
    if (d->failed_out) d->failed_out[member] = 1;
    failures.fetch_add(1);
  }

  // a part is complete when every still-active sender has delivered it
  void recheck_locked() {
    for (OwnPart& o : own) {
      if (o.done) continue;
      bool all = true;
      for (int k = 0; k < d->n_members; ++k)
        if (active[k] && !o.have[k]) { all = false; break; }
      if (all) o.done = true;
    }
    cv.notify_all();
  }

  void drop_sender(int k) {
    std::lock_guard<std::mutex> lk(mu);
    if (active[k]) {
      active[k] = 0;
      recheck_locked();
    }
  }

  // ---- owner side: one connection from sender k ----
  void serve_rx(int fd, int k, float w, std::atomic<bool>* broken) {
    std::vector<unsigned char> buf;
    for (;;) {
      uint64_t tag = 0;
      if (!recv_frame(fd, &tag, buf, d->timeout_ms, max_frame)) break;
      if (tag == kDone) return;
      const int i = static_cast<int>(tag & 0xffffffffu), codec = static_cast<int>(tag >> 32);
      if (i < p0 || i >= p1 || codec != own[i - p0].codec) break;
      OwnPart& o = own[i - p0];
      std::vector<float> x(static_cast<size_t>(o.n));
      if (!decode(codec, buf.data(), static_cast<long long>(buf.size()), x.data(), o.n)) break;
      std::lock_guard<std::mutex> lk(mu);
      if (o.have[k] || o.done) continue;     // duplicate, or the part was closed without this sender
      if (w > 0.f) {
        b200h_weighted_add(o.acc.data(), x.data(), w, o.n);
        o.denom += w;
      }
      o.lossy[k] = std::move(x);
      o.have[k] = 1;
      recheck_locked();
    }
    if (!broken->exchange(true)) mark_failed(k);
    drop_sender(k);
  }

  void serve_tx(int fd, int k, std::atomic<bool>* broken) {
    std::vector<unsigned char> enc;
    std::vector<float> delta;
    for (int j = 0; j < p1 - p0; ++j) {
      OwnPart& o = own[j];
      {
        std::unique_lock<std::mutex> lk(mu);
        const bool ok = cv.wait_for(lk, std::chrono::milliseconds(4ll * d->timeout_ms),
                                    [&] { return o.done || broken->load(); });
        if (!ok || broken->load() || !o.have[k]) return;   // this sender dropped out: nothing to send back
      }
      delta.resize(static_cast<size_t>(o.n));
      b200h_delta(o.acc.data(), o.denom > 0 ? static_cast<float>(1.0 / o.denom) : 1.0f, o.lossy[k].data(), delta.data(), o.n);
      encode(o.codec, delta.data(), o.n, enc);
      const uint64_t tag = static_cast<uint64_t>(p0 + j) | (static_cast<uint64_t>(o.codec) << 32);
      if (b200h_send_frame(fd, tag, enc.data(), static_cast<long long>(enc.size())) != 0) {
        if (!broken->exchange(true)) mark_failed(k);
        drop_sender(k);
        return;
      }
      std::vector<float>().swap(o.lossy[k]);
    }
  }

  // ---- sender side: my parts of owner k's span ----
  void exchange_with_owner(int k) {
    const int q0 = d->span_begin[k], q1 = d->span_end[k];
    const int fd = b200h_connect(d->hosts[k], d->ports[k], d->timeout_ms);
    if (fd < 0) { mark_failed(k); return; }
    Hello h;
    std::memcpy(h.gid, d->group_id, 16);
    h.sender = d->my_index;
    h.weight = d->weight;
    std::atomic<bool> bad{false};
    if (b200h_send_frame(fd, kHello, &h, sizeof(h)) != 0) bad.store(true);
    std::thread tx([&] {
      std::vector<unsigned char> enc;
      for (int i = q0; i < q1 && !bad.load(); ++i) {
        const long long s = d->part_start[i], n = d->part_end[i] - s;
        encode(d->part_codec[i], d->flat + s, n, enc);
        const uint64_t tag = static_cast<uint64_t>(i) | (static_cast<uint64_t>(d->part_codec[i]) << 32);
        if (b200h_send_frame(fd, tag, enc.data(), static_cast<long long>(enc.size())) != 0) bad.store(true);
      }
      if (!bad.load() && b200h_send_frame(fd, kDone, nullptr, 0) != 0) bad.store(true);
    });
    std::vector<unsigned char> buf;
    std::vector<float> delta;
    int received = 0;
    while (received < q1 - q0 && !bad.load()) {
      uint64_t tag = 0;
      if (!recv_frame(fd, &tag, buf, 4 * d->timeout_ms, max_frame)) { bad.store(true); break; }
      const int i = static_cast<int>(tag & 0xffffffffu), codec = static_cast<int>(tag >> 32);
      if (i < q0 || i >= q1 || codec != d->part_codec[i]) { bad.store(true); break; }
      const long long s = d->part_start[i], n = d->part_end[i] - s;
      delta.resize(static_cast<size_t>(n));
      if (!decode(codec, buf.data(), static_cast<long long>(buf.size()), delta.data(), n)) { bad.store(true); break; }
      b200h_weighted_add(d->flat + s, delta.data(), 1.0f, n);   // x += delta (keeps the local rounding error out)
      ++received;
    }
    tx.join();
    b200h_close(fd);
    if (bad.load()) mark_failed(k);
  }
};

}  // namespace

extern "C" int b200h_allreduce_round(const B200hAllreduce* d) {
  if (!d || d->n_members < 1 || d->my_index < 0 || d->my_index >= d->n_members) return -1;
  const int me = d->my_index, n = d->n_members;
  if (d->failed_out) std::memset(d->failed_out, 0, sizeof(int) * n);
  Round r(d);
  std::vector<std::thread> threads;
  // outgoing: my contribution to every other owner's span
  if (r.is_sender[me])
    for (int k = 0; k < n; ++k)
      if (k != me && d->span_end[k] > d->span_begin[k]) threads.emplace_back([&r, k] { r.exchange_with_owner(k); });
  // incoming: the other senders' contributions to my span
  int expected = 0;
  for (int k = 0; k < n; ++k) expected += r.active[k];
  std::vector<std::unique_ptr<std::atomic<bool>>> flags;
  std::vector<int> fds;
  std::vector<char> connected(n, 0);
  if (r.p1 > r.p0 && expected > 0) {
    if (!d->listener) return -2;
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(d->timeout_ms);
    int got = 0;
    std::vector<unsigned char> buf;
    while (got < expected) {
      const auto left = std::chrono::duration_cast<std::chrono::milliseconds>(deadline - std::chrono::steady_clock::now()).count();
      if (left <= 0) break;
      const int fd = b200h_accept(d->listener, static_cast<int>(std::min<long long>(left, 200)));
      if (fd < 0) continue;
      uint64_t tag = 0;
      Hello h;
      if (!recv_frame(fd, &tag, buf, d->timeout_ms, 64) || tag != kHello || buf.size() != sizeof(Hello)) { b200h_close(fd); continue; }
      std::memcpy(&h, buf.data(), sizeof(h));
      if (std::memcmp(h.gid, d->group_id, 16) != 0 || h.sender < 0 || h.sender >= n || !r.active[h.sender] ||
          connected[h.sender] || !(h.weight >= 0.f)) {
        b200h_close(fd);      // another round's straggler, or not a member
        continue;
      }
      connected[h.sender] = 1;
      ++got;
      flags.emplace_back(new std::atomic<bool>(false));
      std::atomic<bool>* broken = flags.back().get();
      fds.push_back(fd);
      const int k = h.sender;
      const float w = h.weight;
      threads.emplace_back([&r, fd, k, w, broken] { r.serve_rx(fd, k, w, broken); });
      threads.emplace_back([&r, fd, k, broken] { r.serve_tx(fd, k, broken); });
    }
    for (int k = 0; k < n; ++k)
      if (r.active[k] && !connected[k]) {   // never showed up: average over the others
        r.mark_failed(k);
        r.drop_sender(k);
      }
  }
  for (std::thread& t : threads) t.join();
  for (int fd : fds) b200h_close(fd);
  // my own span: the average replaces my values (an auxiliary peer has none of its own)
  if (r.is_sender[me])
    for (OwnPart& o : r.own)
      if (o.denom > 0) {
        const float inv = static_cast<float>(1.0 / o.denom);
        float* dst = d->flat + o.start;
        for (long long i = 0; i < o.n; ++i) dst[i] = o.acc[i] * inv;
      }
  return r.failures.load();
}
