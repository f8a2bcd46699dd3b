"""tcgen05 SGNS block kernel vs the fp32 PyTorch oracle (bf16 operands, fp32 accumulate)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

EMPTY = 0xFFFFFFFF


def _case(device, n_tok, U, sent_len, window, seed=0, with_invalid=True):
    from swiftsnails_b200.ops._util import u32_tensor

    rng = np.random.default_rng(seed)
    torch.manual_seed(seed)
    params = (torch.rand(U, 128, device=device) - 0.5) * 0.6
    inv_in = rng.integers(0, U, n_tok).astype(np.int64)
    inv_out = rng.integers(0, U, n_tok).astype(np.int64)
    if with_invalid:
        inv_in[::53] = EMPTY
        inv_out[::47] = EMPTY
    tiles = (n_tok + 127) // 128
    neg = rng.integers(0, U, tiles * 16).astype(np.int64)
    if with_invalid:
        neg[5] = EMPTY
    return params, inv_in, inv_out, neg, (u32_tensor(inv_in, device), u32_tensor(inv_out, device),
                                          u32_tensor(neg, device))


def _run(device, mode, n_tok, U, sent_len, window, box_rows=1):
    from swiftsnails_b200.ops import sgns
    from swiftsnails_b200.utils.reference import sgns_block_ref

    params, inv_in, inv_out, neg, (d_in, d_out, d_neg) = _case(device, n_tok, U, sent_len, window)
    params_bf16 = params.to(torch.bfloat16)
    params = params_bf16.float()          # the oracle sees the same bf16-rounded inputs, in fp32
    grads = torch.zeros_like(params)
    loss = torch.zeros(2, device=device)
    sgns.sgns_block_tc(params_bf16, grads, d_in, d_out, sent_len, window, 7, 123, d_neg, 5.0, loss, mode=mode,
                       box_rows=box_rows)
    torch.cuda.synchronize()
    gref, lref, npairs = sgns_block_ref(params, inv_in, inv_out, neg, sent_len, window, 7, 123, 5.0)
    assert float(loss[1]) == npairs
    assert abs(float(loss[0]) - lref) / lref < 2e-3, (float(loss[0]), lref)
    err = (grads - gref).abs().max().item()
    scale = gref.abs().max().item()
    assert err < 1e-2 * scale + 1e-4, (err, scale)


@pytest.mark.parametrize("n_tok,U,sent_len,window", [(128, 300, 1000, 5), (1000, 500, 37, 5), (5000, 64, 100, 8),
                                                      (300, 2000, 50, 1)])
def test_sgns_block_tc_vector_load(device, n_tok, U, sent_len, window):
    _run(device, 0, n_tok, U, sent_len, window)


@pytest.mark.parametrize("n_tok,U,sent_len,window", [(128, 300, 1000, 5), (1000, 500, 37, 5), (5000, 64, 100, 8)])
def test_sgns_block_tc_tma_gather4(device, n_tok, U, sent_len, window):
    _run(device, 1, n_tok, U, sent_len, window, box_rows=int(os.environ.get("SSB200_GATHER_BOX_ROWS", "1")))


@pytest.mark.parametrize("n_tok,U,sent_len,window", [(128, 300, 1000, 5), (1000, 500, 37, 5), (5000, 64, 100, 8), (128 * 148 * 2 + 5, 3000, 100, 5)])
def test_sgns_block_tc_cp_async_gather(device, n_tok, U, sent_len, window):
    """Mode bit 32: rows are gathered by four warps of 16-byte cp.async (manual 128-byte swizzle) instead of TMA."""
    _run(device, 33, n_tok, U, sent_len, window)


def test_fused_pull_direct_mode_matches_cache_mode(device):
    """tcgen05_direct: the TMA gathers operand rows straight from the table's bf16 row mirror; the
    gradients must equal those of the cache path (same bf16 inputs, same reductions)."""
    from swiftsnails_b200.ops import sgns
    from swiftsnails_b200.ops._util import u32_tensor
    from swiftsnails_b200.ops.dedup import Deduper
    from swiftsnails_b200.ops.table import ShardedTable

    rng = np.random.default_rng(3)
    n_tok, V = 2048, 700
    t = ShardedTable.local(4096, 128, device, with_state=True, with_shadow=True, init_scale=0.5)
    tokens = rng.integers(0, V, n_tok).astype(np.uint32)
    negs = rng.integers(0, V, (n_tok // 128) * 16).astype(np.uint32) + V
    keys = u32_tensor(np.concatenate([tokens, tokens + V, negs]), device)
    uniq, inv, nu = Deduper(keys.numel(), device)(keys)
    n = keys.numel()
    cache = torch.empty(n, 128, dtype=torch.bfloat16, device=device)
    gslot = torch.empty(n, dtype=torch.int32, device=device)
    t.pull(uniq, n_dev=nu, rows_bf16=cache, want_fp32=False, gslot_out=gslot, n=n)
    g1, g2 = torch.zeros(n, 128, device=device), torch.zeros(n, 128, device=device)
    l1, l2 = torch.zeros(2, device=device), torch.zeros(2, device=device)
    args = (inv[:n_tok], inv[n_tok:2 * n_tok], 100, 5, 9, 0, inv[2 * n_tok:], 5.0)
    sgns.sgns_block_tc(cache, g1, *args, l1, mode=1)
    sgns.sgns_block_tc_direct(t, gslot, g2, *args, l2, mode=1)
    torch.cuda.synchronize()
    assert float(l2[1]) == float(l1[1]) > 0
    assert abs(float(l1[0]) - float(l2[0])) < 1e-3 * float(l1[0])
    # duplicate rows of a tile fold into G with bf16 shared-memory reductions whose order is not fixed,
    # so two runs agree to bf16 rounding of those entries, not bit for bit
    err, scale = (g1 - g2).abs().max().item(), g1.abs().max().item()
    assert err < 1e-2 * scale, (err, scale)


def _serial_child():
    """Runs in a child process: the round-1 serial kernel (mode bit 8; bit 4 = its diagonal epilogue),
    kept for A/B timing, on the standard cases."""
    dev = torch.device("cuda", 0)
    for mode in (9, 13):
        for n_tok, U, sent_len, window in [(128, 300, 1000, 5), (1000, 500, 37, 5), (5000, 64, 100, 8), (300, 2000, 50, 1)]:
            _run(dev, mode, n_tok, U, sent_len, window)
    print("SERIAL_OK")


def test_sgns_block_tc_serial_kernel_kept_for_ab(device):
    """Mode bit 8 selects the serial (round-1) kernel: same gradients and loss as the oracle. Run in a child
    process so that a fault there cannot poison this process's CUDA context."""
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r); import test_gpu_sgns_tc as t; t._serial_child()" % (
        root, os.path.join(root, "tests"))
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=root)
    assert p.returncode == 0 and "SERIAL_OK" in p.stdout, p.stdout[-1500:] + p.stderr[-1500:]


@pytest.mark.parametrize("n_tok", [128 * 148 * 3 + 77, 128 * 5])
def test_sgns_block_tc_many_tiles_per_cta(device, n_tok):
    """Several tiles per persistent CTA: the cross-tile pipeline (S of tile t+1 under the epilogues of tile
    t, single-buffered TMEM accumulators, two shared-memory stages) must not mix tiles up."""
    _run(device, 1, n_tok, 4000, 211, 5)
