"""Critic layer 1 (M=1024, N=64, K=24+30000 with table epilogue; split-K) in a loop: for ncu captures."""
import sys, torch
sys.path.insert(0, '/root/repo')
from gsc_b200 import ops
ops.precision = "tf32"
M, N, A, nm = 1024, 64, 30000, 100
emb = torch.randn(M, 24, device='cuda'); act = torch.rand(M, A, device='cuda')
W = torch.randn(N, 24 + 2 * A, device='cuda') / 100; b = torch.randn(N, device='cuda')
n = torch.full((M,), nm, dtype=torch.int32, device='cuda')
table = ops.mask_table(W[:, 24:24 + A], nm, 3)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
def run():
    return ops.linear_fwd([emb, act], W, b, relu=True, col_offsets=[0, 24 + A], row_table=(table, n))
for it in range(4):
    flush.zero_(); run()
torch.cuda.synchronize()
ts = []
for it in range(10):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
print("layer1 us/launch (L2 flushed)", sorted(ts)[len(ts) // 2])
