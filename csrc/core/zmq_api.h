// zmq_api.h — the slice of the libzmq C API the transfer layer uses. This image has no zmq.h;
// the host core links against the libzmq 4.3.5 bundled with pyzmq (see build.py find_libzmq),
// so the prototypes are declared here (ABI-stable since libzmq 4.0).
#pragma once
#include <cstddef>
extern "C" {
typedef struct zmq_msg_t {
  alignas(void*) unsigned char _[64];
} zmq_msg_t;
void* zmq_ctx_new(void);
int zmq_ctx_term(void* context);
int zmq_ctx_set(void* context, int option, int optval);
void* zmq_socket(void* context, int type);
int zmq_close(void* s);
int zmq_bind(void* s, const char* addr);
int zmq_connect(void* s, const char* addr);
int zmq_setsockopt(void* s, int option, const void* optval, size_t optvallen);
int zmq_getsockopt(void* s, int option, void* optval, size_t* optvallen);
int zmq_msg_init(zmq_msg_t* msg);
int zmq_msg_init_size(zmq_msg_t* msg, size_t size);
int zmq_msg_close(zmq_msg_t* msg);
int zmq_msg_move(zmq_msg_t* dest, zmq_msg_t* src);
size_t zmq_msg_size(const zmq_msg_t* msg);
void* zmq_msg_data(zmq_msg_t* msg);
int zmq_msg_send(zmq_msg_t* msg, void* s, int flags);
int zmq_msg_recv(zmq_msg_t* msg, void* s, int flags);
int zmq_msg_more(const zmq_msg_t* msg);
int zmq_errno(void);
const char* zmq_strerror(int errnum);
}
#define ZMQ_PULL 7
#define ZMQ_PUSH 8
#define ZMQ_SNDMORE 2
#define ZMQ_DONTWAIT 1
#define ZMQ_LINGER 17
#define ZMQ_SNDHWM 23
#define ZMQ_RCVHWM 24
#define ZMQ_RCVTIMEO 27
#define ZMQ_LAST_ENDPOINT 32
#define ZMQ_IO_THREADS 1
