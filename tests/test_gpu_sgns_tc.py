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
