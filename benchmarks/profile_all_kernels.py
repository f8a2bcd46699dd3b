#!/usr/bin/env python
"""Launch every kernel of libssb200_kernels.so once at a representative size, in a fixed order, so
one `ncu --set full -c <N>` call captures all of them (profiles/prof_all_kernels.md)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from swiftsnails_b200 import _lib
from swiftsnails_b200.ops import hashing, lr as lr_ops, sgns
from swiftsnails_b200.ops._util import u32_tensor
from swiftsnails_b200.ops.dedup import Deduper
from swiftsnails_b200.ops.sampling import AliasSampler
from swiftsnails_b200.ops.table import ShardedTable
from swiftsnails_b200.parallel.hashfrag import HashFrag

dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
rng = np.random.default_rng(0)
N, D = 1 << 20, 128
keys_np = rng.integers(0, 1 << 22, N, dtype=np.int64).astype(np.uint32)
keys = u32_tensor(keys_np, dev)
t = ShardedTable.local(1 << 23, D, dev, num_shards=4, frag_num=1024, with_state=True)
ded = Deduper(N, dev)
smp = AliasSampler(1.0 / np.arange(1, 100001) ** 0.75, dev)
warm = [t.pull(keys[:1024])]
torch.cuda.synchronize()
print("== profiled launches start")
hashing.fmix64(torch.arange(N, device=dev, dtype=torch.int64))                      # fmix64_kernel
bucketed, perm, counts, offs = hashing.bucket_by_owner(keys, t.frag_map_dev, 4)     # hash_route + (torch cumsum) + bucket_scatter
uniq, inv, nu = ded(keys)                                                           # dedup_insert + dedup_resolve
rows, gslot = t.pull(uniq, n_dev=nu)                                                # pull_kernel (inserts)
rows, gslot = t.pull(uniq, n_dev=nu)                                                # pull_kernel (hits)
t.gather_slots(gslot, n_dev=nu)                                                     # gather_slots_kernel
t.enable_slot_cache(N)
t.pull(uniq, n_dev=nu)                                                              # pull_kernel<cached> (fills the slot cache)
t.pull(uniq, n_dev=nu)                                                              # pull_kernel<cached> (all hits)
t.slot_cache, t.slot_cache_bits, t.slot_cache_dense = None, 0, False
g = torch.randn(N, D, device=dev)
_lib.call("ssb_push_bf16", t._view_ref(), gslot.data_ptr(), g.to(torch.bfloat16).data_ptr(), N, nu.data_ptr(), 0.01, 1e-8, 1,
          _lib.stream_ptr())                                                         # push_bf16_kernel<AdaGrad> (inbox apply)
t.push_sgd(gslot, g, 0.01, n_dev=nu, zero_grads=False)                              # push_kernel<SGD>
t.push_adagrad(gslot, g, 0.01, n_dev=nu, zero_grads=False)                          # push_kernel<AdaGrad>
smp.sample(N)                                                                       # neg_sample_kernel
B = 262144
inv_in = inv[:B].contiguous(); inv_out = inv[B:2 * B].contiguous()
pi = torch.empty(B * 10, dtype=torch.int32, device=dev); po = torch.empty_like(pi); cnt = torch.zeros(1, dtype=torch.int32, device=dev)
sgns.make_pairs(inv_in, inv_out, 1000, 5, 1, 0, pi, po, cnt)                        # sgns_make_pairs_kernel
grads = torch.zeros_like(rows); loss = torch.zeros(2, device=dev)
neg = inv[2 * B:2 * B + (B * 10 // 128 + 1) * 16].contiguous()
sgns.sgns_tiles(rows, grads, pi, po, cnt, B * 10, neg, 16, 5 / 16, loss)            # sgns_tiles_kernel<16,1>
sgns.sgns_block_tc(rows.to(torch.bfloat16), grads, inv_in, inv_out, 1000, 5, 1, 0, neg, 5.0, loss, mode=1)  # sgns_block_tc_kernel
t1 = ShardedTable.local(1 << 22, 1, dev, with_state=True)
fk = u32_tensor(rng.integers(0, 1 << 21, (65536, 39), dtype=np.int64).astype(np.uint32), dev)
u1, i1, n1 = Deduper(fk.numel(), dev)(fk.reshape(-1))
p1, g1 = t1.pull(u1, n_dev=n1)
gr1 = torch.zeros_like(p1)
lr_ops.lr_fwd_bwd(p1, gr1, i1, torch.rand(65536, 39, device=dev), (torch.rand(65536, device=dev) < 0.3).float(), 39, loss)  # lr_fwd_bwd_kernel
t.dump_shard(t.owned[0])                                                            # dump_compact_kernel x2
t2 = ShardedTable.local(1 << 22, D, dev)
t2.load(uniq[:100000], rows[:100000])                                               # table_load_kernel
a = torch.empty(256 << 20, dtype=torch.uint8, device=dev); b = torch.empty_like(a)
_lib.call("ssb_p2p_copy", a.data_ptr(), b.data_ptr(), a.numel(), 0, _lib.stream_ptr())   # p2p_copy_kernel
torch.cuda.synchronize()
print("== done")
