"""Bring-up diagnostics for the tcgen05 SGNS kernel: per-operand-group gradient error."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from swiftsnails_b200.ops import sgns
from swiftsnails_b200.ops._util import u32_tensor
from swiftsnails_b200.utils.reference import sgns_block_ref

dev = torch.device("cuda:0")
mode = int(sys.argv[1]) if len(sys.argv) > 1 else 0
box = int(sys.argv[2]) if len(sys.argv) > 2 else 1
n_tok, sent_len, window = 256, 1000, 5
rng = np.random.default_rng(0)
torch.manual_seed(0)
U1, U2, U3 = 1000, 2000, 2100
params_bf16 = ((torch.rand(U3, 128, device=dev) - 0.5) * 0.6).to(torch.bfloat16)
params = params_bf16.float()
inv_in = rng.permutation(U1)[:n_tok].astype(np.int64)
inv_out = (U1 + rng.permutation(U2 - U1)[:n_tok]).astype(np.int64)
tiles = (n_tok + 127) // 128
neg = (U2 + rng.permutation(U3 - U2)[:tiles * 16]).astype(np.int64)
grads = torch.zeros_like(params)
loss = torch.zeros(2, device=dev)
sgns.sgns_block_tc(params_bf16, grads, u32_tensor(inv_in, dev), u32_tensor(inv_out, dev), sent_len, window, 7, 123,
                   u32_tensor(neg, dev), 5.0, loss, mode=mode, box_rows=box)
torch.cuda.synchronize()
gref, lref, npairs = sgns_block_ref(params, inv_in, inv_out, neg, sent_len, window, 7, 123, 5.0)
print("mode", mode, "loss", float(loss[0]), lref, "pairs", float(loss[1]), npairs)
for name, lo, hi in (("in", 0, U1), ("ctx", U1, U2), ("neg", U2, U3)):
    g, r = grads[lo:hi], gref[lo:hi]
    print(name, "max|ref|=%.4f max|got|=%.4f maxerr=%.4f nnz_rows got=%d ref=%d" % (
        r.abs().max().item(), g.abs().max().item(), (g - r).abs().max().item(),
        int((g.abs().sum(1) > 0).sum()), int((r.abs().sum(1) > 0).sum())))
    # is got a column/row permutation? compare sorted values of one row
    nz = torch.nonzero(r.abs().sum(1) > 0).flatten()
    if len(nz):
        i = int(nz[0])
        print("   row", i, "ref[:8]", [round(x, 4) for x in r[i, :8].tolist()])
        print("   row", i, "got[:8]", [round(x, 4) for x in g[i, :8].tolist()])
        print("   sorted-equal:", torch.allclose(torch.sort(r[i]).values, torch.sort(g[i]).values, atol=2e-2),
              " sum ref %.4f got %.4f" % (r[i].sum().item(), g[i].sum().item()))

# ---- raw dump of tile 0: G (swizzled smem image), TMEM S | dIn | dON^T ----
dbg = torch.zeros(32768 + 128 * 416, device=dev)
grads.zero_(); loss.zero_()
sgns.sgns_block_tc(params_bf16, grads, u32_tensor(inv_in, dev), u32_tensor(inv_out, dev), sent_len, window, 7, 123,
                   u32_tensor(neg, dev), 5.0, loss, mode=mode, box_rows=box, debug=dbg)
torch.cuda.synchronize()
d = dbg.cpu().numpy()
Gs = d[:32768].reshape(4, 128, 64)   # [col block][row][64 bf16 swizzled in 16B chunks]
Gf = np.zeros((128, 256), np.float32)
for cb in range(4):
    for r in range(128):
        row = Gs[cb, r].reshape(8, 8)
        for ch in range(8):
            Gf[r, cb * 64 + ch * 8: cb * 64 + ch * 8 + 8] = row[ch ^ (r & 7)]
G = Gf[:, :160]
T = d[32768:].reshape(128, 416)
dOnB, dIn, dOnA = T[:, :128], T[:, 160:288], T[:, 288:416]
P = params.cpu().numpy()
A = P[inv_in[:128]]
ctx_tok = np.arange(-8, 136)
ON = np.zeros((160, 128), np.float32)
ok = (ctx_tok >= 0) & (ctx_tok < n_tok)
ON[:144][ok] = P[inv_out[ctx_tok[ok]]]
ON[144:] = P[neg[:16]]
print("G nnz", int((G != 0).sum()), "G absmax", np.abs(G).max())
dIn_ref = G @ ON
dOn_ref = G.T @ A
dOnT_ref = dOn_ref[:128]
dOnT = dOnA
print("dOnB err", np.abs(dOnB[:32] - dOn_ref[128:160]).max())
print("dIn  max|ref| %.4f max|got| %.4f err %.4f" % (np.abs(dIn_ref).max(), np.abs(dIn).max(), np.abs(dIn - dIn_ref).max()))
print("dOnT max|ref| %.4f max|got| %.4f err %.4f" % (np.abs(dOnT_ref).max(), np.abs(dOnT).max(), np.abs(dOnT - dOnT_ref).max()))
print("dIn got[0,:6]", dIn[0, :6], "ref", dIn_ref[0, :6])
print("dOnA got[9,:6]", dOnT[9, :6], "ref", dOnT_ref[9, :6])
