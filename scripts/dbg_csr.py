"""Debug helper: CSR build + SpMM on a small random graph, step by step with synchronisation."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gnn_b200 import ops
from oracle import pyg_shim

def sync(tag):
    torch.cuda.synchronize(); print("ok", tag, flush=True)

n, e = 97, 900
rng = np.random.default_rng(1)
ei = rng.integers(0, n, size=(2, e)); ei[:, :45] = rng.integers(0, n, size=45)
ew = rng.random(e).astype(np.float32)
ei_t, ew_t = torch.from_numpy(ei), torch.from_numpy(ew)
for asl in (True, False):
    g = ops.gcn_graph(ei_t.cuda(), ew_t.cuda(), n, asl); sync(f"csr asl={asl}")
    ei2, w2 = pyg_shim.gcn_norm(ei_t, ew_t, n, asl)
    rp = g.rowptr.cpu().numpy(); col = g.col.cpu().numpy(); val = g.val.cpu().numpy()
    src, dst, w = ei2[0].numpy(), ei2[1].numpy(), w2.numpy()
    bad = 0
    for i in range(n):
        want_c = list(src[dst == i]); got_c = list(col[rp[i]:rp[i+1]])
        want_v = w[dst == i]; got_v = val[rp[i]:rp[i+1]]
        if want_c != got_c or not np.allclose(want_v, got_v, rtol=1e-6):
            bad += 1
            if bad < 4: print("row", i, "want", want_c, want_v, "got", got_c, got_v)
    print("asl", asl, "bad rows", bad, "nnz", g.nnz, rp[-1])
    for F in (1, 3, 100):
        H = torch.randn(n, F, device="cuda")
        out = ops.spmm(g.rowptr, g.col, g.val, H, None); sync(f"spmm F={F}")
        A = torch.zeros(n, n, dtype=torch.float64); A.index_put_((ei2[1], ei2[0]), w2.double(), accumulate=True)
        ref = (A @ H.cpu().double())
        print("F", F, "max err", float((out.cpu().double() - ref).abs().max()))
        out_t = ops.spmm(g.rowptr_t, g.col_t, g.val_t, H, None); sync(f"spmm_t F={F}")
        print("F", F, "max err T", float((out_t.cpu().double() - A.t() @ H.cpu().double()).abs().max()))
