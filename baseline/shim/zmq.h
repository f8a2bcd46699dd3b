// zmq.h stand-in: the image has libzmq 4.3.5 (bundled with pyzmq) but no header.
//
// The reference was written against libzmq 4.0.5 and reads a received meta frame by copying the
// first bytes of the zmq_msg_t STRUCT (core/Message.h:156) — true for 4.0.x, where a small
// message's payload sits at offset 0 of the struct, false for libzmq >= 4.1.
// Rather than patch the reference, this shim presents a zmq_msg_t with the 4.0.x property:
// `head` mirrors the first 32 payload bytes, `real` is the genuine 4.3.5 message the calls are
// forwarded to. The reference source in baseline/_ref stays byte-identical to /root/reference.
#pragma once
#include <cstddef>
#include <cstring>
extern "C" {
typedef struct ssb_real_zmq_msg_t { alignas(void*) unsigned char _[64]; } ssb_real_zmq_msg_t;
typedef void(zmq_free_fn)(void* data, void* hint);
void* zmq_ctx_new(void);
int zmq_ctx_destroy(void* context);
void* zmq_socket(void* context, int type);
int zmq_close(void* s);
int zmq_bind(void* s, const char* addr);
int zmq_connect(void* s, const char* addr);
int zmq_errno(void);
const char* zmq_strerror(int errnum);
int ssb_real_msg_init(ssb_real_zmq_msg_t*) __asm__("zmq_msg_init");
int ssb_real_msg_init_size(ssb_real_zmq_msg_t*, size_t) __asm__("zmq_msg_init_size");
int ssb_real_msg_init_data(ssb_real_zmq_msg_t*, void*, size_t, zmq_free_fn*, void*) __asm__("zmq_msg_init_data");
int ssb_real_msg_close(ssb_real_zmq_msg_t*) __asm__("zmq_msg_close");
int ssb_real_msg_move(ssb_real_zmq_msg_t*, ssb_real_zmq_msg_t*) __asm__("zmq_msg_move");
size_t ssb_real_msg_size(const ssb_real_zmq_msg_t*) __asm__("zmq_msg_size");
void* ssb_real_msg_data(ssb_real_zmq_msg_t*) __asm__("zmq_msg_data");
int ssb_real_msg_send(ssb_real_zmq_msg_t*, void*, int) __asm__("zmq_msg_send");
int ssb_real_msg_recv(ssb_real_zmq_msg_t*, void*, int) __asm__("zmq_msg_recv");
int ssb_real_msg_more(const ssb_real_zmq_msg_t*) __asm__("zmq_msg_more");
}
#define ZMQ_PULL 7
#define ZMQ_PUSH 8
#define ZMQ_SNDMORE 2
#define ZMQ_DONTWAIT 1

struct zmq_msg_t {
  unsigned char head[32];   // mirror of the first payload bytes (libzmq 4.0.x inline layout)
  ssb_real_zmq_msg_t real;
};
static inline void ssb_zmq_sync_head(zmq_msg_t* m) {
  size_t n = ssb_real_msg_size(&m->real);
  std::memset(m->head, 0, sizeof(m->head));
  if (n) std::memcpy(m->head, ssb_real_msg_data(&m->real), n < sizeof(m->head) ? n : sizeof(m->head));
}
static inline int zmq_msg_init(zmq_msg_t* m) { std::memset(m->head, 0, sizeof(m->head)); return ssb_real_msg_init(&m->real); }
static inline int zmq_msg_init_size(zmq_msg_t* m, size_t n) { std::memset(m->head, 0, sizeof(m->head)); return ssb_real_msg_init_size(&m->real, n); }
static inline int zmq_msg_init_data(zmq_msg_t* m, void* d, size_t n, zmq_free_fn* f, void* h) {
  int r = ssb_real_msg_init_data(&m->real, d, n, f, h);
  if (r == 0) ssb_zmq_sync_head(m);
  return r;
}
static inline int zmq_msg_close(zmq_msg_t* m) { return ssb_real_msg_close(&m->real); }
static inline int zmq_msg_move(zmq_msg_t* d, zmq_msg_t* s) {
  int r = ssb_real_msg_move(&d->real, &s->real);
  std::memcpy(d->head, s->head, sizeof(d->head));
  return r;
}
static inline size_t zmq_msg_size(zmq_msg_t* m) { return ssb_real_msg_size(&m->real); }
static inline void* zmq_msg_data(zmq_msg_t* m) { return ssb_real_msg_data(&m->real); }
static inline int zmq_msg_send(zmq_msg_t* m, void* s, int flags) { return ssb_real_msg_send(&m->real, s, flags); }
static inline int zmq_msg_recv(zmq_msg_t* m, void* s, int flags) {
  int r = ssb_real_msg_recv(&m->real, s, flags);
  if (r >= 0) ssb_zmq_sync_head(m);
  return r;
}
static inline int zmq_msg_more(zmq_msg_t* m) { return ssb_real_msg_more(&m->real); }
