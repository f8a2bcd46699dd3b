"""How long does the host take to enqueue one word2vec step (launch-bound check)?"""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from swiftsnails_b200.models.word2vec import Word2VecConfig, Word2VecWorker, synthetic_corpus
from swiftsnails_b200.parallel.cluster import make_table
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
for pipe in (False, True):
    cfg = Word2VecConfig(tile_impl="tcgen05", optimizer="adagrad", lr=0.05, tokens_per_step=262144)
    table = make_table(2 * cfg.vocab, cfg.dim, dev, with_state=True, init_scale=1.0 / cfg.dim)
    w = Word2VecWorker(cfg, table, dev, pipeline=pipe)
    toks = torch.from_numpy(synthetic_corpus(cfg.vocab, 10 * cfg.tokens_per_step, seed=1).view(np.int32)).reshape(10, -1).to(dev)
    for s in range(10): w.step(toks[s])
    w.flush(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(100): w.step(toks[s % 10])
    t_enq = time.perf_counter() - t0
    w.flush(); torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    print("pipeline", pipe, "host enqueue %.1f us/step, total %.1f us/step" % (t_enq * 1e4, t_all * 1e4))
    del w, table
