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

#ifdef __cplusplus
}
#endif
#endif
