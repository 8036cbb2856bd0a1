import ctypes, torch, numpy as np, sys
sys.path.insert(0,'/root/repo')
from gsc_b200 import synth, ops, _lib
from gsc_b200.graph import GraphBatch
from gsc_b200.models import Actor, AgentSpec
lib=_lib.load()
lib.gsc_debug_set_gnn_timers.argtypes=[ctypes.c_void_p]; lib.gsc_debug_set_gnn_timers.restype=None
for (B,nm,lo,hi,links) in [(4096,24,16,24,None),(1024,100,100,100,150)]:
    st=synth.make_states(B,nm,lo,hi,seed=1,fixed_links=links,num_topologies=32)
    gb=GraphBatch.from_states(st).to('cuda')
    actor=Actor(AgentSpec(nm)).to('cuda')
    gnn,_=actor.views(actor.flat)
    rp,col,gp=gb.csr()
    dbg=torch.zeros(16,device='cuda')
    for it in range(3):
        ops.gnn_embed_fwd(gb.x,rp,col,gp,B,gb.max_nodes,gnn,22,2,2)
    torch.cuda.synchronize()
    lib.gsc_debug_set_gnn_timers(dbg.data_ptr())
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record(); ops.gnn_embed_fwd(gb.x,rp,col,gp,B,gb.max_nodes,gnn,22,2,2); e1.record(); torch.cuda.synchronize()
    lib.gsc_debug_set_gnn_timers(None)
    print(B,nm,"fwd kernel ms",e0.elapsed_time(e1),"cycles {setup,xrows,mma,xl,phaseB,pool}:",dbg.tolist()[:6])
    emb,saved=ops.gnn_embed_fwd(gb.x,rp,col,gp,B,gb.max_nodes,gnn,22,2,2,save=True)
    d_emb=torch.randn_like(emb)
    for it in range(2): ops.gnn_embed_bwd(gb.x,rp,col,gp,B,gb.max_nodes,gnn,22,2,2,saved,d_emb,pack_stats=gb.pack_stats())
    torch.cuda.synchronize(); dbg.zero_()
    lib.gsc_debug_set_gnn_timers(dbg.data_ptr())
    e0.record(); ops.gnn_embed_bwd(gb.x,rp,col,gp,B,gb.max_nodes,gnn,22,2,2,saved,d_emb,pack_stats=gb.pack_stats()); e1.record(); torch.cuda.synchronize()
    lib.gsc_debug_set_gnn_timers(None)
    print(B,nm,"bwd kernel ms",e0.elapsed_time(e1),"cycles {chunk,hrows,R,rows,edgeS,W,split,H,edgeT}:",dbg.tolist()[:9])
