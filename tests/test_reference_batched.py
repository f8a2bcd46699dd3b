"""The batched fp32 oracle of the block-formulation SGNS step (utils/reference.sgns_block_ref_batched, used by
benchmarks/precision_parity.py at full size) equals the per-tile loop oracle the kernel tests use."""
import numpy as np
import torch


def test_batched_block_reference_equals_the_loop_reference():
    from swiftsnails_b200.utils.reference import EMPTY, sgns_block_ref, sgns_block_ref_batched

    torch.manual_seed(0)
    rng = np.random.default_rng(0)
    n_rows, D, n_tok, sent_len, window = 500, 32, 128 * 5 + 37, 50, 5
    params = torch.randn(n_rows, D) * 0.3
    inv_in = rng.integers(0, n_rows, n_tok).astype(np.int64)
    inv_out = rng.integers(0, n_rows, n_tok).astype(np.int64)
    inv_in[rng.random(n_tok) < 0.05] = EMPTY
    inv_out[rng.random(n_tok) < 0.05] = EMPTY
    n_tiles = (n_tok + 127) // 128
    negs = rng.integers(0, n_rows, n_tiles * 16).astype(np.int64)
    negs[rng.random(len(negs)) < 0.1] = EMPTY
    g0, l0, p0 = sgns_block_ref(params, inv_in, inv_out, negs, sent_len, window, 9, 1234, 5.0)
    for chunk in (2, 1024):
        g1, l1, p1 = sgns_block_ref_batched(params, torch.from_numpy(inv_in), torch.from_numpy(inv_out), torch.from_numpy(negs),
                                            sent_len, window, 9, 1234, 5.0, chunk_tiles=chunk)
        assert p0 == p1 and abs(l0 - l1) < 1e-6 * abs(l0)
        assert torch.allclose(g0, g1, rtol=1e-5, atol=1e-6)
