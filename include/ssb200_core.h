/*
 * ssb200_core.h — C ABI of the C++17 host core (libssb200_core.so): config parser, hash
 * fragments, the reference-format text dump, the CPU sparse table, and the master / server /
 * worker roles over ZeroMQ.  `ref:` citations point into /root/reference/src.
 */
#pragma once
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int ssb_core_version(void);

/* config — ref: utils/ConfigParser.h:25-129 */
void ssb_config_clear(void);
int ssb_config_load(const char *path);                      /* 0 ok, -1 cannot open          */
void ssb_config_set(const char *key, const char *value);
int ssb_config_parse_line(const char *line);
int ssb_config_get(const char *key, char *out, int out_len); /* length, -1 missing, -2 too small */
int ssb_config_keys(char *out, int out_len);                /* newline-separated key list     */

/* hashing — ref: utils/HashFunction.h:16-24, parameter/hashfrag.h:36-53 */
uint64_t ssb_hash_code(uint64_t x);
int ssb_hashfrag_build(int num_nodes, int frag_num, uint32_t *map_out);
void ssb_hashfrag_route(const uint32_t *keys, int64_t n, const uint32_t *map, int frag_num,
                        int32_t *node_out);

/* BinaryBuffer self test — ref: utils/Buffer.h:174-234 */
int ssb_buffer_selftest(void);

/* text parameter dump (`key \t value \n`) and its loader — ref: parameter/sparsetable.h:49-56,
 * utils/vec1.h:106-112; the loader has no counterpart in the reference (it never reads a dump). */
int64_t ssb_dump_write_text(const char *path, const uint32_t *keys, const float *rows, int64_t n,
                            int dim, int append);
int64_t ssb_dump_read_text(const char *path, uint32_t *keys, float *rows, int64_t max_n, int dim);
/* Param-style dump lines "<key>\tVec:\t v…\tAcc:\t a…" (row + optimizer state, the whole Param as the reference's
 * operator<< would print it, parameter/sparsetable.h:49-56); state may be NULL. digits = significant digits
 * (6 = the reference's ostream default, 9 = exact float round trip). */
int64_t ssb_dump_write_text2(const char *path, const uint32_t *keys, const float *rows, const float *state, int64_t n,
                             int dim, int append, int digits);
int64_t ssb_dump_read_text2(const char *path, uint32_t *keys, float *rows, float *state, int64_t max_n, int dim,
                            int *has_state);
/* The same with state_dim floats of optimizer state per entry (state words x dim: AdaGrad dim, FTRL 2 x dim). */
int64_t ssb_dump_write_text3(const char *path, const uint32_t *keys, const float *rows, const float *state, int64_t n,
                             int dim, int state_dim, int append, int digits);
int64_t ssb_dump_read_text3(const char *path, uint32_t *keys, float *rows, float *state, int64_t max_n, int dim,
                            int state_dim, int *has_state);

/* CPU sparse table with the stock SGD / AdaGrad access methods — ref: parameter/sparsetable.h */
void *ssb_cpu_table_new(int shard_num, int dim, int adagrad, float lr, float eps, uint64_t seed,
                        float init_scale);
void ssb_cpu_table_free(void *t);
void ssb_cpu_table_pull(void *t, const uint32_t *keys, int64_t n, float *rows_out);
int64_t ssb_cpu_table_push(void *t, const uint32_t *keys, const float *grads, int64_t n);
int64_t ssb_cpu_table_size(void *t);
int ssb_cpu_table_dump(void *t, const char *path);

/* corpus loader: background parser of `one sentence of integer word ids per line` files into
 * fixed-size token batches (sentences padded to sent_len with pad_id). next() returns the number
 * of real tokens in the batch or -1 at the end. ref: utils/file.h:14-33 (scan_file_by_line). */
void *ssb_corpus_open(const char *path, int batch_tokens, int sent_len, uint32_t pad_id, int epochs);
/* Same, keeping only the non-empty lines whose index % split_n == split_i (per-worker data split). */
void *ssb_corpus_open_split(const char *path, int batch_tokens, int sent_len, uint32_t pad_id, int epochs,
                            int split_n, int split_i);
/* full form: `threads` parser threads (<= 0: chosen from the host), ids >= max_id become padding (0 = no bound) */
void *ssb_corpus_open_ex(const char *path, int batch_tokens, int sent_len, uint32_t pad_id, int epochs, int split_n,
                         int split_i, int threads, uint32_t max_id);
/* out[0] = sentence lines per pass (this reader's share), out[1] = minibatches, out[2] = ids replaced by padding */
void ssb_corpus_stats(void *h, long long *out);
/* counts[vocab] += word frequencies (for the unigram^0.75 noise table); returns the number of ids >= vocab, -1 = cannot open */
long long ssb_corpus_count_words(const char *path, uint32_t vocab, unsigned long long *counts, int threads);
long ssb_corpus_next(void *h, uint32_t *out);
void ssb_corpus_close(void *h);

/* sparse-LR example files `label key:value …` -> padded arrays for the GPU step (keys == NULL: count only);
 * the format of the CPU worker's lr app (csrc/apps/algorithms.h). Lines with index % split_n == split_i. */
long long ssb_lr_parse_file(const char *path, int max_feat, uint32_t *keys, float *vals, float *labels,
                            long long max_examples, unsigned long long key_mod, int split_n, int split_i,
                            long long *out_stats);

/* roles — ref: core/framework/SwiftMaster.h, SwiftServer.h, SwiftWorker.h. One role per process. */
int ssb_run_master(const char *config_path);
int ssb_run_server(const char *config_path);
int ssb_run_worker(const char *config_path, const char *app, const char *data_path);

/* A GPU rank's control plane (system/group_exchange.h): the process hosts server shard g and worker g. It registers
 * with the master as a server over ZeroMQ (NODE_INIT_ADDRESS; rank = assigned id - 1), adopts the route, fetches
 * the fragment map (NODE_ASKFOR_HASHFRAG), all-gathers small byte strings through the master (NODE_GROUP_EXCHANGE:
 * socket names for the VMM fd exchange, barriers, max-over-ranks), and shuts down with WORKER_FINISH_WORK ->
 * SERVER_TOLD_TO_TERMINATE -> dump -> ack. ref: system/node_init.h:16-152, master/terminate.h, server/terminate.h */
int ssb_node_start(const char *master_addr, const char *listen_addr, int init_timeout_s);  /* rank, or < 0 */
int ssb_node_rank(void);
int ssb_node_world(void);
int ssb_node_hashfrag(int *out2, uint32_t *map_out, int cap);
long ssb_node_exchange(const void *buf, long len, void *out, long out_cap, long *lens, int timeout_s);
int ssb_node_finish(int timeout_s);
void ssb_node_dump_done(void);
void ssb_node_stop(void);

#ifdef __cplusplus
}
#endif
