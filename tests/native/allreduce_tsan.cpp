// ThreadSanitizer / AddressSanitizer driver for the C++ data plane (native/allreduce_pipe.cpp + transport.cpp +
// quant_host.cpp): three peers inside one process - a node, a node and an auxiliary member; the second run adds a
// client-mode sender - run b200h_allreduce_round concurrently over loopback sockets with all three codecs in the part
// table.  Built by tests/test_native_host.py with -fsanitize=thread (and again with address,undefined); exits 0 when
// the averages are right; the sanitizer itself fails the process on a report.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "b200_host_api.h"

static int run_case(bool with_client) {
  const int n = with_client ? 4 : 3;
  // modes: 0 node, 1 client, 2 aux
  std::vector<int> modes = {0, 0, 2};
  if (with_client) modes.push_back(1);
  const long long total = 3 * 4096 + 1000;
  const int n_parts = 4;
  long long ps[n_parts] = {0, 4096, 8192, 12288}, pe[n_parts] = {4096, 8192, 12288, total};
  int codec[n_parts] = {0, 1, 2, 1};
  // owners: members 0, 1, 2 (the client owns nothing)
  std::vector<int> sb = {0, 1, 3}, se = {1, 3, 4};
  if (with_client) { sb.push_back(4); se.push_back(4); }
  std::vector<void*> listeners(n, nullptr);
  std::vector<int> ports(n, 0);
  for (int k = 0; k < n; ++k)
    if (modes[k] != 1) {
      listeners[k] = b200h_listen("127.0.0.1", 0, &ports[k]);
      if (!listeners[k]) { fprintf(stderr, "listen failed\n"); return 2; }
    }
  std::vector<const char*> hosts(n, "127.0.0.1");
  std::vector<std::vector<float>> flat(n, std::vector<float>(total));
  std::vector<float> weight(n, 0.f);
  for (int k = 0; k < n; ++k) {
    weight[k] = modes[k] == 2 ? 0.f : static_cast<float>(k + 1);
    for (long long i = 0; i < total; ++i) flat[k][i] = std::sin(0.001f * static_cast<float>(i)) + 0.1f * static_cast<float>(k);
  }
  unsigned char gid[16];
  for (int i = 0; i < 16; ++i) gid[i] = static_cast<unsigned char>(17 * i + (with_client ? 1 : 0));
  std::vector<std::vector<int>> failed(n, std::vector<int>(n, 0));
  std::vector<int> rc(n, -100);
  std::vector<std::thread> th;
  for (int k = 0; k < n; ++k)
    th.emplace_back([&, k] {
      B200hAllreduce d;
      std::memset(&d, 0, sizeof(d));
      d.n_members = n; d.my_index = k; d.modes = modes.data(); d.hosts = hosts.data(); d.ports = ports.data();
      d.span_begin = sb.data(); d.span_end = se.data(); d.n_parts = n_parts; d.part_start = ps; d.part_end = pe;
      d.part_codec = codec; d.flat = flat[k].data(); d.weight = weight[k]; d.group_id = gid;
      d.listener = listeners[k]; d.timeout_ms = 5000; d.failed_out = failed[k].data();
      rc[k] = b200h_allreduce_round(&d);
    });
  for (auto& t : th) t.join();
  for (int k = 0; k < n; ++k)
    if (listeners[k]) b200h_close_listener(listeners[k]);
  int bad = 0;
  for (int k = 0; k < n; ++k) bad += rc[k] != 0;
  // expected weighted mean over the senders (everyone but the aux member)
  double wsum = 0, off = 0;
  for (int k = 0; k < n; ++k) if (modes[k] != 2) { wsum += weight[k]; off += weight[k] * 0.1 * k; }
  off /= wsum;
  for (int k = 0; k < n; ++k) {
    if (modes[k] == 2) continue;
    for (long long i = 0; i < total; i += 97) {
      const double want = std::sin(0.001 * static_cast<double>(i)) + off;
      const int part = i < 4096 ? 0 : (i < 8192 ? 1 : (i < 12288 ? 2 : 3));
      const double tol = codec[part] == 0 ? 1e-5 : (codec[part] == 1 ? 2e-3 : 4e-2);
      if (std::fabs(flat[k][i] - want) > tol) {
        if (bad < 5) fprintf(stderr, "member %d elem %lld: %f vs %f (codec %d)\n", k, i, flat[k][i], want, codec[part]);
        ++bad;
      }
    }
  }
  return bad;
}

int main() {
  const int a = run_case(false);
  const int b = run_case(true);
  printf("allreduce sanitizer driver: %d + %d problems\n", a, b);
  return (a || b) ? 1 : 0;
}
