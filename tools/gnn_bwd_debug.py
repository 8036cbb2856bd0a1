import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from gsc_b200 import ops, synth
from gsc_b200.graph import GraphBatch
from oracle import gsc_oracle as O
from test_gpu_kernels import _embed_case, _flat
B, n_max, lo, hi, H, F, nl, ni = [int(v) for v in (sys.argv[1:9] if len(sys.argv) > 8 else "9 8 2 8 22 3 2 2".split())]
st, sd = _embed_case(B, n_max, lo, hi, H, F, nl, ni, seed=B + H + 1)
sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
ref = O.gnn_embedder(torch.from_numpy(st.x).double(), torch.from_numpy(st.edge_index), torch.from_numpy(st.batch), sd, nl, ni, prefix="", num_graphs=B)
gen = torch.Generator().manual_seed(1)
d_emb = torch.randn(B, H, generator=gen, dtype=torch.float64)
(ref * d_emb).sum().backward()
gb = GraphBatch.from_states(st).to("cuda")
rowptr, col, gptr = gb.csr()
flat = _flat({k: v.detach() for k, v in sd.items()}, nl)
emb, saved = ops.gnn_embed_fwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, nl, ni, save=True)
d = ops.gnn_embed_bwd(gb.x, rowptr, col, gptr, B, gb.max_nodes, flat, H, nl, ni, saved, d_emb.float().to("cuda"), pack_stats=gb.pack_stats())
got = d.cpu().numpy()
off = 0
names = ["encoder."] + [f"process.{i}." for i in range(nl - 1)]
for nme in names:
    for k in ("att", "bias", "lin_l.weight", "lin_l.bias", "lin_r.weight", "lin_r.bias"):
        w = sd[nme + k].grad.reshape(-1).numpy(); n = w.size
        g = got[off:off + n]
        print(f"{nme+k:28s} max|ref|={np.abs(w).max():.3e} max|got|={np.abs(g).max():.3e} err={np.abs(g-w).max()/max(np.abs(w).max(),1e-30):.3e}")
        if "lin_l.weight" in k and n <= 80: print("   ref", np.round(w[:12],5), "\n   got", np.round(g[:12],5))
        off += n
