import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gsc_b200 import synth, ops
from gsc_b200.graph import GraphBatch
from gsc_b200.models import Actor, AgentSpec
B, nm, lo, hi, links = 1024, 100, 100, 100, 150
st = synth.make_states(B, nm, lo, hi, seed=1, fixed_links=links, num_topologies=32)
gb = GraphBatch.from_states(st).to('cuda')
actor = Actor(AgentSpec(nm)).to('cuda')
gnn, _ = actor.views(actor.flat)
rp, col, gp = gb.csr()
emb, saved = ops.gnn_embed_fwd(gb.x, rp, col, gp, B, gb.max_nodes, gnn, 22, 2, 2, save=True)
d_emb = torch.randn_like(emb)
for it in range(3):
    ops.gnn_embed_bwd(gb.x, rp, col, gp, B, gb.max_nodes, gnn, 22, 2, 2, saved, d_emb, pack_stats=gb.pack_stats())
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for it in range(10):
    ops.gnn_embed_bwd(gb.x, rp, col, gp, B, gb.max_nodes, gnn, 22, 2, 2, saved, d_emb, pack_stats=gb.pack_stats())
e1.record(); torch.cuda.synchronize()
print("gnn_bwd us/call", e0.elapsed_time(e1) * 100)
e0.record()
for it in range(10):
    ops.gnn_embed_fwd(gb.x, rp, col, gp, B, gb.max_nodes, gnn, 22, 2, 2, save=True)
e1.record(); torch.cuda.synchronize()
print("gnn_fwd us/call", e0.elapsed_time(e1) * 100)
