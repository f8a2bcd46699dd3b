#!/usr/bin/env python
"""Timeline of one CTA of the pipelined tensor-core SGNS kernel: clock64() stamps that CTA 0 records at the
hand-over points of every role (csrc/kernels/sgns_tc.cu, SSB_STAMP) on a real bench.py step. Prints, for a few
steady-state tiles, every event in cycles relative to the tile's "gradient MMAs issued" of the previous tile, and
the mean duration of each phase. usage: python benchmarks/tc_timeline.py [tokens] [mode]"""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

NAMES = {0: "prod: loop top", 1: "prod: stage free (empty_bar)", 2: "prod: gather issued/handed over", 3: "prod: lead_bar arrive",
         4: "mma: S issued", 5: "mma: grads issued", 6: "help: idx_bar seen", 7: "help: cp.async issued + win arrive",
         8: "help: g_free seen", 9: "help: G zeroed (g_zero arrive)", 10: "e1: loop top", 11: "e1: s_full seen",
         12: "e1: S in regs (s_free arrive)", 13: "e1: rotated", 22: "e1: win+early seen", 14: "e1: sigmoids done (phase A)", 15: "e1: cells ready (phase B)", 23: "e1: plain stores issued",
         16: "e1: g_zero seen", 17: "e1: g_full arrive", 18: "e2: loop top", 19: "e2: d_full seen",
         20: "e2: d_free arrive (TMEM read)", 21: "e2: reds issued"}
SLOTS = 24


def main():
    from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker, synthetic_corpus
    from swiftsnails_b200.ops import sgns as sgns_ops
    from swiftsnails_b200.parallel.cluster import make_table

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    tps = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
    mode = int(sys.argv[2]) if len(sys.argv) > 2 else 33
    cfg = Word2VecConfig(tile_impl="tcgen05", optimizer="adagrad", lr=0.05, tokens_per_step=tps, seed=1)
    table = make_table(2 * cfg.vocab, cfg.dim, dev, with_state=True, init_scale=1.0 / cfg.dim, frag_num=1024)
    allk = torch.arange(0, 2 * cfg.vocab, dtype=torch.int32, device=dev)
    for c in range(0, allk.numel(), 1 << 20):
        table.resolve(allk[c:c + (1 << 20)])
    w = Word2VecWorker(cfg, table, dev)
    toks = torch.from_numpy(synthetic_corpus(cfg.vocab, 3 * tps, seed=50).view(np.int32)).reshape(3, -1).to(dev)
    s0, B = w.sets[0], tps
    n_my = ((tps + 127) // 128 + 147) // 148
    dbg = torch.zeros(n_my * SLOTS * 2 + 64, dtype=torch.float32, device=dev)
    for s in range(3):
        w._comm(s0, toks[s])
        dbg.zero_()
        sgns_ops.sgns_block_tc(s0.params_bf16, s0.grads, s0.inverse[:B], s0.inverse[B:2 * B], cfg.sent_len, cfg.window, cfg.seed, s * B,
                               s0.inverse[2 * B:], float(cfg.negatives), w.loss, mode=mode, debug=dbg)
        torch.cuda.synchronize()
    st = dbg.cpu().numpy().view(np.int64)[:n_my * SLOTS].reshape(n_my, SLOTS).astype(np.float64)
    t_first = st[0][st[0] > 0].min()
    t_last = st[n_my - 1].max()
    print("tiles of CTA 0: %d, cycles from first to last stamp: %.0f (%.0f per tile)" % (n_my, t_last - t_first, (t_last - t_first) / n_my))
    for it in (8, 9, 10):
        if it >= n_my:
            continue
        ref = st[it - 1][5]          # gradient MMAs of the previous tile issued
        print("\n-- tile %d (cycles after 'grads issued' of tile %d) --" % (it, it - 1))
        for slot, t in sorted(((k, st[it][k]) for k in NAMES if st[it][k] > 0), key=lambda kv: kv[1]):
            print("  %8.0f  %s" % (t - ref, NAMES[slot]))
    body = st[4:n_my - 2]
    d = lambda a, b: float(np.mean(body[:, b] - body[:, a]))
    per_tile = float(np.mean(np.diff(st[4:n_my - 2, 5])))
    out = {"mode": mode, "tokens": tps, "cycles_per_tile_grads_issue_to_issue": round(per_tile),
           "producer_wait_stage": round(d(0, 1)), "producer_publish_and_gather_issue": round(d(1, 2)), "producer_dedup_and_meta": round(d(2, 3)),
           "helper_gather_issue_and_window": round(d(6, 7)), "helper_wait_g_free": round(d(7, 8)), "helper_zero_G": round(d(8, 9)),
           "e1_wait_s_full": round(d(10, 11)), "e1_tmem_load": round(d(11, 12)), "e1_rotate": round(d(12, 13)), "e1_wait_win_early": round(d(13, 22)), "e1_sigmoids_A": round(d(22, 14)),
           "e1_wait_lead_and_cells_B": round(d(14, 15)), "e1_wait_g_zero": round(d(15, 16)),
           "e1_plain_stores": round(d(16, 23)), "e1_atomics_fence_arrive": round(d(23, 17)),
           "e2_wait_d_full": round(d(18, 19)), "e2_tmem_read_until_d_free": round(d(19, 20)), "e2_rest_of_reds": round(d(20, 21)),
           "S_issue_to_s_full_seen": round(float(np.mean(body[:, 11] - body[:, 4]))),
           "g_full_arrive_to_grads_issued": round(float(np.mean(body[:, 5] - body[:, 17]))),
           "grads_issued_to_d_full_seen": round(float(np.mean(body[:, 19] - body[:, 5]))),
           "grads_issued_to_g_free_seen": round(float(np.mean(body[1:, 8] - body[:-1, 5])))}
    print("\n" + json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
