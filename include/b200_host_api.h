/* C ABI of libdalle_b200_host.so - host-side C++ (OpenMP numeric twins + bulk TCP transport). */
#ifndef B200_HOST_API_H_
#define B200_HOST_API_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Block-wise 8-bit quantisation with a sorted 256-entry code (bitsandbytes cquantize_blockwise_cpu_fp32 twin;
 * used by reference lib/training/lamb_8bit.py:181-192). */
void b200h_quantize_blockwise(const float* code, const float* x, float* absmax, uint8_t* out, long long n, int blocksize);
void b200h_dequantize_blockwise(const float* code, const uint8_t* q, const float* absmax, float* out, long long n,
                                int blocksize);
/* hivemind Uniform8BitQuantization (reference task.py:125-126): q[n] + 256-entry lookup of bucket means. */
void b200h_uniform8_quantize(const float* x, long long n, uint8_t* q, float* lookup);
void b200h_uniform8_dequantize(const uint8_t* q, const float* lookup, float* out, long long n);
/* Owner-side part reduction of the butterfly all-reduce (hivemind TensorPartReducer). */
void b200h_weighted_add(float* acc, const float* x, float w, long long n);
void b200h_delta(const float* acc, float inv_denom, const float* mine, float* out, long long n);

/* Bulk TCP transport (replaces the go-libp2p daemon data path): framed blocking sockets + epoll acceptor. */
void* b200h_listen(const char* host, int port, int* out_port);
int b200h_accept(void* listener, int timeout_ms);
void b200h_close_listener(void* listener);
int b200h_connect(const char* host, int port, int timeout_ms);
int b200h_send_frame(int fd, unsigned long long tag, const void* data, long long nbytes);
long long b200h_recv_header(int fd, unsigned long long* tag, int timeout_ms);
int b200h_recv_payload(int fd, void* data, long long nbytes, int timeout_ms);
void b200h_close(int fd);

/* One peer's side of a butterfly all-reduce round, run by C++ threads over the bulk transport (the caller's GIL is
 * released for the whole call): for every other owner, parts of its span are compressed from `flat` and sent while
 * the returned deltas are decoded and added in place; for this peer's own span, parts arriving from the other
 * senders are decoded, accumulated with their weights, averaged when complete, and the per-sender deltas are
 * compressed and sent back.  Semantics of hivemind's AllReduceRunner / TensorPartReducer: a failing or
 * late sender is dropped and the part is averaged over the rest; auxiliary members own spans and contribute nothing;
 * client-mode members contribute and own nothing.
 * modes: 0 node, 1 client, 2 auxiliary.  part_codec: 0 fp32, 1 fp16 (clamped), 2 uniform 8-bit (hivemind).
 * Returns the number of exchanges that failed (0 = clean), < 0 on a setup error; failed_out[k] = 1 for every member
 * whose exchange with this peer broke. */
typedef struct {
  int n_members, my_index;
  const int* modes;
  const char* const* hosts;
  const int* ports;
  const int* span_begin;
  const int* span_end;
  int n_parts;
  const long long* part_start;
  const long long* part_end;
  const int* part_codec;
  float* flat;
  float weight;
  const unsigned char* group_id; /* 16 bytes */
  void* listener;                /* this peer's b200h_listen handle; NULL for a client-mode peer */
  int timeout_ms;
  int* failed_out;
} B200hAllreduce;
int b200h_allreduce_round(const B200hAllreduce* d);

#ifdef __cplusplus
}
#endif
#endif
