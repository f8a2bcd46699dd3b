/*
 * ssb200.h — C ABI of the swiftsnails_b200 device library (libssb200_kernels.so).
 *
 * Every entry point takes raw device pointers plus a cudaStream_t (passed as void*),
 * launches on that stream and returns a cudaError_t as int (0 = success); nothing here
 * synchronises the stream unless the comment says so.  The kernels are written for
 * sm_100a only.  Citations `ref:` point into /root/reference/src (the CPU reference whose
 * behaviour the kernel reproduces); see SURVEY.md §8 for the K-numbering.
 */
#pragma once
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSB_MAX_SHARDS 16
#define SSB_EMPTY_KEY 0xFFFFFFFFu  /* ref: core/parameter/sparsetable.h:13 (empty key = max()) */
#define SSB_LOCKED_KEY 0xFFFFFFFEu /* slot claimed, row being initialised (device-table only)   */
#define SSB_BUCKET 32              /* keys per probe bucket = one 128-byte line                   */
#define SSB_INVALID_SLOT 0xFFFFFFFFu

/* A view of the sharded sparse table: one shard per GPU ("server"), every shard mapped into
 * this process (own HBM or a peer's over NVLink).  Global slot index = shard * capacity + slot.
 * ref: core/parameter/sparsetable.h:69-121 (SparseTable), hashfrag.h:12-116 (BasicHashFrag). */
typedef struct SsbTableView {
  uint32_t *keys[SSB_MAX_SHARDS];  /* [capacity] per shard, SSB_EMPTY_KEY when free          */
  float *rows[SSB_MAX_SHARDS];     /* [capacity][dim] parameter rows                          */
  float *state[SSB_MAX_SHARDS];    /* [capacity][state_words][dim] optimizer state (AdaGrad
                                      accumulator = 1 word; FTRL z, n = 2 words), or NULL        */
  void *shadow[SSB_MAX_SHARDS];    /* [capacity][dim] bf16 mirror of rows, or NULL. Kept in step by
                                      every owner-side writer (insert, load, local/inbox push); a
                                      pull that only wants bf16 reads it: half the row bytes over
                                      HBM / NVLink. Not maintained by REMOTE direct pushes.       */
  const uint32_t *frag_map;        /* [frag_num] fragment -> node id (1..num_shards), device  */
  uint32_t frag_num;
  uint32_t capacity;               /* slots per shard, multiple of SSB_BUCKET                 */
  int32_t dim;                     /* floats per row                                          */
  int32_t num_shards;
  uint64_t seed;                   /* deterministic lazy init: row = f(key, seed)             */
  float init_scale;                /* init value = (u01 - 0.5) * init_scale                   */
  int32_t state_words;             /* floats of optimizer state per weight (0 reads as 1)     */
  int32_t init_kind;               /* SSB_INIT_*: how the pull kernels create a missing row   */
  int32_t pad_;
} SsbTableView;

/* Lazy row init kinds (ref: PullAccessMethod::init_param, sparse_access_method.h:21). Element j of key's row:
 * UNIFORM (u01(key, j, seed) - 0.5) * init_scale (the word2vec app's randInit, vec1.h:223-226); ZERO 0 (sparse LR,
 * FTRL); CONSTANT init_scale; UNIFORM_POS u01 * init_scale. */
#define SSB_INIT_UNIFORM 0
#define SSB_INIT_ZERO 1
#define SSB_INIT_CONSTANT 2
#define SSB_INIT_UNIFORM_POS 3

/* ---- K1: key hashing / routing / bucketing ------------------------------------------------
 * ref: utils/HashFunction.h:16-24, parameter/hashfrag.h:48-53, global_pull_access.h:58-72 */
int ssb_fmix64(const uint64_t *in, uint64_t *out, int64_t n, void *stream);
/* owner[i] = frag_map[fmix64(key) % frag_num] - 1; counts[o] += 1 (counts must be zeroed). */
int ssb_hash_route(const uint32_t *keys, int64_t n, const uint32_t *frag_map, uint32_t frag_num,
                   int32_t num_owners, int32_t *owner, uint32_t *counts, void *stream);
/* Partition keys by owner: offsets[o] = exclusive scan of counts (device, num_owners+1),
 * cursor[o] zeroed scratch; bucketed[offsets[o] + k] = key, perm[...] = original index. */
int ssb_bucket_scatter(const uint32_t *keys, const int32_t *owner, int64_t n,
                       const uint32_t *offsets, uint32_t *cursor, uint32_t *bucketed,
                       uint32_t *perm, void *stream);

/* ---- K2: de-duplication ------------------------------------------------------------------
 * ref: parameter/global_param_cache.h:107-108 (grads is a map => unique keys).
 * Scratch: set[set_cap] uint64 (epoch-tagged, never needs clearing between calls as long as
 * `epoch` changes; zero-fill once), set_ids[set_cap] uint32.  set_cap must be a power of two
 * >= 2*n.  Outputs: unique[<=n], inverse[n] (index into unique), *n_unique (device, must be
 * zeroed by the caller or by passing reset=1). Keys equal to SSB_EMPTY_KEY map to
 * inverse = SSB_INVALID_SLOT and are not emitted. */
int ssb_dedup(const uint32_t *keys, int64_t n, uint64_t *set, uint32_t *set_ids, uint32_t set_cap,
              uint32_t epoch, uint32_t *unique, uint32_t *inverse, uint32_t *n_unique, int reset,
              void *stream);

/* Allocation-free variant of ssb_dedup (no counterpart in the reference, which has no device
 * code): the scratch hash set is a fixed 1.5 MiB __device__ array inside the library, one copy per
 * GPU that loads it, held until the process exits. For callers that may not allocate device
 * memory — hosts that own all of it through their own arena/pool, and latency-critical paths
 * where cudaMalloc would stall the device. At most SSB_DEDUP_STATIC_MAX_KEYS keys per call (more
 * is refused with cudaErrorInvalidValue and a message on stderr). The first call on each device
 * resolves the array's address (cudaGetSymbolAddress); make it outside any stream capture. Calls
 * share the array and each one clears it on the caller's stream first; calls that overlap on
 * the GPU race. Same outputs as ssb_dedup (the order of `unique` is unspecified in both). */
#define SSB_DEDUP_STATIC_MAX_KEYS 65536
int ssb_dedup_static(const uint32_t *keys, int64_t n, uint32_t *unique, uint32_t *inverse,
                     uint32_t *n_unique, void *stream);

/* ---- K3/K4/K10: table probe+insert, pull gather, cache merge -------------------------------
 * ref: sparsetable.h:142-149 (lazy init_param on miss), global_pull_access.h:40-120,
 * system/server/init.h:49-69.
 * n_dev (device uint32, may be NULL) overrides n when smaller (size known only on device).
 * gslot[i] = shard*capacity + slot.  rows_out [n][dim] fp32 and/or rows_bf16 [n][dim] bf16 (the
 * tensor-core cache) may be NULL.  grads_zero [n][dim] (may be NULL) is cleared in the same pass.
 * insert=0: missing keys return SSB_INVALID_SLOT and the deterministic init row.        */
int ssb_pull(const SsbTableView *view, const uint32_t *keys, int64_t n, const uint32_t *n_dev,
             uint32_t *gslot, float *rows_out, void *rows_bf16, float *grads_zero, int insert,
             uint32_t *stats, void *stream);
/* Same, with a worker-side slot cache: slot_cache[2^cache_bits] 64-bit entries in the caller's own
 * device memory, initialised to all ones (0xFF bytes) and owned by ONE table instance. A key found in
 * the cache skips the probe of the owner's key bucket (one NVLink round trip less per remote key);
 * keys resolved the slow way are remembered. Legal because slots never move and never die. */
int ssb_pull_cached(const SsbTableView *view, const uint32_t *keys, int64_t n, const uint32_t *n_dev,
                    uint32_t *gslot, float *rows_out, void *rows_bf16, float *grads_zero,
                    int insert, uint32_t *stats, uint64_t *slot_cache, int32_t cache_bits,
                    void *stream);
/* Gather by already-resolved global slots (worker-side slot cache hit path). */
int ssb_gather_slots(const SsbTableView *view, const uint32_t *gslot, int64_t n,
                     const uint32_t *n_dev, float *rows_out, void *stream);

/* ---- K7/K8: push = scatter-add into the owner shard fused with the server update ----------
 * ref: global_push_access.h:36-149, sparsetable.h:181-192, sparse_access_method.h:33-46.
 * SGD:     w -= lr * g                    (one vector `red` per 16 bytes, remote or local)
 * AdaGrad: acc += g*g; w -= lr*g/sqrt(acc+eps)   (vector `atom` on acc, `red` on w)
 * zero_grads=1 clears grads after reading (Grad::reset(), global_push_access.h:78).  */
int ssb_push_sgd(const SsbTableView *view, const uint32_t *gslot, float *grads, int64_t n,
                 const uint32_t *n_dev, float lr, int zero_grads, void *stream);
int ssb_push_adagrad(const SsbTableView *view, const uint32_t *gslot, float *grads, int64_t n,
                     const uint32_t *n_dev, float lr, float eps, int zero_grads, void *stream);

/* Inbox push (multi-GPU AdaGrad): ship (slot, grad row) once into the owner's inbox with peer
 * stores; rows owned by `self_shard` are applied directly. After a cross-GPU barrier the owner
 * applies its inbox with ssb_push_adagrad / ssb_push_sgd on a one-shard local view
 * (gslot = inbox slots, grads = inbox rows, n_dev = inbox count) and clears the count.
 * ref: global_push_access.h:123-149 + server/init.h:116-132 (send (key, grad)*, the owner applies). */
typedef struct SsbInboxView {
  float *rows[SSB_MAX_SHARDS];     /* [capacity][dim] per owner (peer-mapped)        */
  uint32_t *slots[SSB_MAX_SHARDS]; /* [capacity] slot within the owner shard        */
  uint32_t *count[SSB_MAX_SHARDS]; /* entries claimed so far                        */
  uint32_t *flags[SSB_MAX_SHARDS]; /* [num_shards] per owner: last step each sender finished */
  uint32_t capacity;
  uint32_t pad_;
} SsbInboxView;
int ssb_push_send(const SsbTableView *view, const SsbInboxView *inbox, const uint32_t *gslot,
                  const float *grads, int64_t n, const uint32_t *n_dev, int32_t self_shard, float lr,
                  float eps, int32_t adagrad, int32_t bf16_rows, uint32_t *overflow, void *stream);
/* Private-region form: `inbox` describes THIS sender's sub-region of every owner's inbox (rows / slots
 * offset to it, capacity = rows per sub-region, count[o] = where owner o expects this sender's row
 * count). Space is claimed with local atomics on cursor[SSB_MAX_SHARDS] (device memory, zero before
 * the first call); ssb_inbox_publish then stores the counts to the owners with plain peer stores,
 * resets the cursor and, when value != 0, raises the hand-off flags. No remote atomic is issued. */
int ssb_push_send2(const SsbTableView *view, const SsbInboxView *inbox, const uint32_t *gslot,
                   const float *grads, int64_t n, const uint32_t *n_dev, int32_t self_shard, float lr,
                   float eps, int32_t adagrad, int32_t bf16_rows, uint32_t *overflow, uint32_t *cursor,
                   void *stream);
int ssb_inbox_publish(const SsbInboxView *inbox, int32_t num_shards, int32_t self, uint32_t *cursor,
                      uint32_t value, void *stream);
/* Owner-side apply of num_segs sub-regions in one launch: segment y = rows [y*seg_rows, y*seg_rows +
 * counts[y]) of gslot / grads_bf16. */
int ssb_push_bf16_seg(const SsbTableView *view, const uint32_t *gslot, const void *grads_bf16,
                      int64_t seg_rows, int32_t num_segs, const uint32_t *counts, float lr, float eps,
                      int32_t adagrad, void *stream);
/* Owner-side apply when the inbox rows are bf16 (bf16_rows=1 above): fp32 update from bf16 gradients. */
int ssb_push_bf16(const SsbTableView *view, const uint32_t *gslot, const void *grads_bf16, int64_t n,
                  const uint32_t *n_dev, float lr, float eps, int32_t adagrad, void *stream);
/* Hand-off flags instead of a collective: signal = store `value` into flags[owner][self] on every
 * owner (after a system fence); wait = spin (bounded) until all of my flags reached `value`. */
int ssb_flags_signal(const SsbInboxView *inbox, int32_t num_shards, int32_t self, uint32_t value, void *stream);
int ssb_flags_wait(const uint32_t *my_flags, int32_t num_shards, uint32_t value, void *stream);
/* Same with an explicit bound (iterations of ~0.2 us; 0 = default ~27 s, or SSB200_WAIT_LIMIT) and an error word: when
 * the wait runs out err[0] is set to 1 and the kernel returns (err == NULL: it traps, as ssb_flags_wait does). */
int ssb_flags_wait2(const uint32_t *my_flags, int32_t num_shards, uint32_t value, uint32_t spin_limit, uint32_t *err,
                    void *stream);

/* Chunked inbox push: ship and apply overlap. The view describes THIS sender's sub-region of every owner's inbox:
 * rows[o] (bf16 gradient rows [seg_rows][dim]), slots[o] ([seg_rows]), ready[o] ([max_chunks] words: chunk c of this
 * sender is complete when it holds this step's tag), total[o] (2 words: {rows shipped, tag}). A chunk is 256 rows.
 * ssb_push_send_chunks applies the rows this GPU owns directly and ships the others; `cursor` [SSB_MAX_SHARDS],
 * `done` [SSB_MAX_SHARDS * max_chunks] and `finished` [1] are sender-local device words, zero before the first call
 * (the kernel leaves them zero). ssb_push_apply_chunks runs on the owner, over ALL senders' sub-regions of its own
 * inbox (slots / rows / ready / total are the owner's local arrays, sender-major), and applies every chunk as soon as
 * its flag or the sender's total carries `tag` (bounded wait: on expiry err[0] is set to 1 and the item is skipped).
 * Launch it on another stream than the send so that the two overlap. `tag` must differ from the tags of the two
 * previous steps (the step number). ref: global_push_access.h:123-149 + server/init.h:116-132. */
typedef struct SsbChunkInboxView {
  void *rows[SSB_MAX_SHARDS];
  uint32_t *slots[SSB_MAX_SHARDS];
  uint32_t *ready[SSB_MAX_SHARDS];
  uint32_t *total[SSB_MAX_SHARDS];
  uint32_t seg_rows;    /* rows per (owner, sender) sub-region, a multiple of 256 */
  uint32_t max_chunks;  /* seg_rows / 256 */
} SsbChunkInboxView;
int ssb_push_send_chunks(const SsbTableView *view, const SsbChunkInboxView *inbox, const uint32_t *gslot,
                         const float *grads, int64_t n, const uint32_t *n_dev, int32_t self_shard, float lr, float eps,
                         int32_t adagrad, uint32_t tag, uint32_t *overflow, uint32_t *cursor, uint32_t *done,
                         uint32_t *finished, void *stream);
int ssb_push_apply_chunks(const SsbTableView *local_view, const uint32_t *slots, const void *rows_bf16,
                          const uint32_t *ready, const uint32_t *total, uint32_t seg_rows, int32_t world, int32_t self,
                          uint32_t tag, float lr, float eps, int32_t adagrad, uint32_t spin_limit, uint32_t *err,
                          void *stream);

/* keys[0..n) = tokens, keys[n..2n) = tokens + vocab (padding 0xFFFFFFFF stays padding): the word2vec step's input and
 * output-vector keys in one pass. */
int ssb_build_keys(const uint32_t *tokens, int64_t n, uint32_t vocab, uint32_t *keys, void *stream);
/* ---- K9: negative sampling (alias method, Philox4x32-10 counter RNG) ---------------------- */
int ssb_neg_sample(const float *prob, const uint32_t *alias, uint32_t vocab, uint64_t seed,
                   uint64_t offset, uint32_t key_offset, uint32_t *out, int64_t n, void *stream);

/* ---- word2vec SGNS (apps are absent from the reference; SURVEY.md §3.4 hot loop I) --------
 * Pair generation: tokens[n_tok] in sentences of `sent_len`; dynamic window shrink like
 * classic word2vec.  inv_in/inv_out map token position -> row index in the worker cache.
 * pairs_in/out are compacted (order unspecified); *n_pairs device counter (zeroed by reset). */
int ssb_sgns_make_pairs(const uint32_t *inv_in, const uint32_t *inv_out, int64_t n_tok,
                        int32_t sent_len, int32_t window, uint64_t seed, uint64_t offset,
                        uint32_t *pairs_in, uint32_t *pairs_out, uint32_t *n_pairs, void *stream);
/* CUDA-core tile kernel: tile = 128 pairs sharing n_neg negatives (rows neg_idx[tile*n_neg+..]).
 * loss_out[0] += sum of losses, loss_out[1] += number of pairs.  neg_weight scales the
 * negative term (K/n_neg keeps the K-negatives objective in expectation). */
int ssb_sgns_tiles(const float *params, float *grads, int32_t dim, const uint32_t *pairs_in,
                   const uint32_t *pairs_out, const uint32_t *n_pairs, int64_t max_pairs,
                   const uint32_t *neg_idx, int32_t n_neg, float neg_weight, float *loss_out,
                   void *stream);

/* ---- sparse logistic regression (K6) -------------------------------------------------------
 * inv[n_ex*n_feat] row index into the cache (dim 1), x values, labels in {0,1}.          */
int ssb_lr_fwd_bwd(const float *params, float *grads, const uint32_t *inv, const float *x,
                   const float *labels, int64_t n_ex, int32_t n_feat, float *loss_out,
                   void *stream);
/* K6 fused: one sparse-LR training step (resolve keys -> gather each touched weight once -> forward / backward ->
 * apply the update rule once per touched key) as ONE persistent cooperative kernel; see csrc/kernels/lr.cu.
 * view->dim must be 1. keys [n_ex][n_feat] (SSB_EMPTY_KEY = empty position), x [n_ex][n_feat] or NULL (ones),
 * labels [n_ex]. Worker-side scratch, allocated once: entries = 16 bytes {tag, idx, slot, -} per feature id
 * (dense_keys > 0: keys must be < dense_keys) or per global slot (dense_keys == 0: every instance probes the
 * table), initialised to {0, 0, SSB_INVALID_SLOT, 0}; stash[n_feat * n_ex]; touched / wc / gc [max_touched]
 * (gc zeroed once; max_touched >= n_ex * n_feat is always enough); counters[16] (out: [0] compact indices used,
 * [2] index overflow, [3] keys without a slot — both must stay 0; [8..12] globaltimer stamps of the phases).
 * step: non-zero, different from every earlier call's. rule_id: 0 SGD, 1 AdaGrad, 2 FTRL; hyper points to an
 * ssb::rule::Hyper {lr, eps, a, b, c, step}. blocks_per_sm: 0 = as many as fit; bit 30 set: keys / x are FEATURE-major
 * ([n_feat][n_ex]: the 32 examples a warp works on are contiguous, so its loads are coalesced).
 * ref: core/framework/SwiftWorker.h:19-57 (the worker loop), parameter/sparse_access_method.h:33-46. */
int ssb_lr_fused_step(const SsbTableView *view, const uint32_t *keys, const float *x, const float *labels, int64_t n_ex,
                      int32_t n_feat, void *entries, uint64_t dense_keys, uint32_t *stash, uint32_t *touched, float *wc,
                      float *gc, uint32_t max_touched, uint32_t *counters, uint32_t step, float *loss_out,
                      int32_t rule_id, const void *hyper, int32_t blocks_per_sm, void *stream);

/* ---- K5 on tensor cores (tcgen05 / TMEM / TMA gather4), dim == 128 only ---------------------
 * One tile = 128 consecutive centre tokens against their 144 context rows + 16 shared negatives.
 * params: bf16 worker cache [n_rows][128]; grads: fp32 [n_rows][128] (accumulated with vector REDs).
 * inv_in / inv_out [n_tok]: cache row of each token's input / output vector; neg_idx
 * [ceil(n_tok/128)*16]: cache rows of each tile's negatives; window <= 8.
 * mode: 0 = vector-load gather, 1 = TMA tile::gather4. box_rows must be 1; debug may be NULL. */
/* Number of persistent CTAs of the tensor-core kernel (0 = one per SM = 148). Fewer leaves whole SMs to the pull /
 * push kernels that run next to it in the pipelined multi-GPU step. Also: environment SSB200_TC_CTAS. */
int ssb_sgns_tc_set_ctas(int32_t n);
int ssb_sgns_block_tc(const void *params, float *grads, int64_t n_rows, int32_t dim,
                      const uint32_t *inv_in, const uint32_t *inv_out, int64_t n_tok,
                      int32_t sent_len, int32_t window, uint64_t seed, uint64_t offset,
                      const uint32_t *neg_idx, float neg_k, float *loss_out, int32_t mode,
                      int32_t box_rows, float *debug, void *stream);
/* Fused-pull variant: operand rows are gathered by TMA straight from the bf16 mirror of the
 * sharded table (local or peer HBM); gslot[] maps a cache row to its global table slot.
 * mirror_base = start of the virtual range holding every shard segment; shard_rows and
 * mirror_row_off are in 256-byte rows; total_rows = rows spanned by the tensor map. */
int ssb_sgns_block_tc_direct(const void *mirror_base, uint64_t total_rows, uint32_t capacity,
                             uint32_t shard_rows, uint32_t mirror_row_off, const uint32_t *gslot,
                             float *grads, int32_t dim, const uint32_t *inv_in,
                             const uint32_t *inv_out, int64_t n_tok, int32_t sent_len,
                             int32_t window, uint64_t seed, uint64_t offset,
                             const uint32_t *neg_idx, float neg_k, float *loss_out, int32_t mode,
                             void *stream);

/* ---- peer memory (CUDA VMM; these DO allocate / map device memory) --------------------------
 * One physical segment per GPU, exported as a POSIX fd, imported by the peers and mapped side by
 * side in one reserved virtual range, so shard s lives at base + s * stride on every rank.    */
int ssb_vmm_granularity(int device, uint64_t *out);
int ssb_vmm_create(int device, uint64_t bytes, uint64_t *handle, int *fd);
int ssb_vmm_import(int fd, uint64_t *handle);
int ssb_vmm_reserve(uint64_t bytes, uint64_t align, uint64_t *va);
int ssb_vmm_map(uint64_t va, uint64_t bytes, uint64_t handle, int device);
int ssb_vmm_unmap(uint64_t va, uint64_t bytes);
int ssb_vmm_release(uint64_t handle);
int ssb_vmm_free_va(uint64_t va, uint64_t bytes);
int ssb_memset_async(void *p, int value, uint64_t bytes, void *stream);

/* Chunked dump for checkpoints: compacts the live slots in [slot_lo, slot_hi) (multiples of 32) of one shard into
 * out_keys / out_rows and, when out_state is given, the optimizer-state rows into out_state at the SAME indices.
 * n_live (device) receives the count; at most max_out entries are written.
 * ref: parameter/sparsetable.h:49-56 — the reference dumps the whole Param (row + whatever state it carries). */
int ssb_dump_compact_range(const uint32_t *keys, const float *rows, const float *state, uint32_t capacity, int32_t dim,
                           uint32_t slot_lo, uint32_t slot_hi, uint32_t *out_keys, float *out_rows, float *out_state,
                           uint32_t *n_live, uint32_t max_out, int32_t state_dim /* state_words * dim; 0 = dim */, void *stream);
/* ---- K11: parameter dump -----------------------------------------------------------------
 * Compacts the live slots of one shard: out_keys[*n_live], out_rows[*n_live][dim].        */
int ssb_dump_compact(const uint32_t *keys, const float *rows, uint32_t capacity, int32_t dim,
                     uint32_t *out_keys, float *out_rows, uint32_t *n_live, uint32_t max_out,
                     void *stream);
/* Bulk load (new capability: checkpoint resume): insert keys with given rows into the view. */
int ssb_table_load(const SsbTableView *view, const uint32_t *keys, const float *rows,
                   const float *state, int64_t n, uint32_t *stats, void *stream);

/* ---- K12: peer bandwidth probe -------------------------------------------------------------
 * Copies n_bytes from src to dst with a grid of vector loads/stores (either may be a peer). */
int ssb_p2p_copy(const void *src, void *dst, int64_t n_bytes, int32_t blocks, void *stream);
/* Reduction-throughput probe (benchmarks/red_probe.py): every warp of `blocks` x 512 threads adds into
 * `iters` random rows of buf [n_rows x 128 fp32]. pattern 0 = one red.v4.f32 per 512-byte row, 1 = four
 * scalar 128-byte reds per row, 2 = red.v4.f32 over four rows x 128 bytes, 3 = plain row stores. */
int ssb_red_probe(float *buf, uint32_t n_rows, int32_t iters, int32_t pattern, int32_t blocks, void *stream);

/* ---- misc ---------------------------------------------------------------------------------- */
int ssb_fill_u32(uint32_t *p, uint32_t v, int64_t n, void *stream);
/* Cap the grid of the grid-stride kernels (pull, push, dedup, sampling, route) at this many blocks
 * per SM so they can share the SMs with the persistent tensor-core kernel (0 = no cap). */
void ssb_set_blocks_per_sm_cap(int blocks_per_sm);
const char *ssb_error_string(int err);
int ssb_version(void);

#ifdef __cplusplus
}
#endif
