"""Actor layer 2 (M=1024, N=30000, K=256, structured-mask epilogue) in a loop: for ncu captures."""
import sys, torch
sys.path.insert(0, '/root/repo')
from gsc_b200 import ops
ops.precision = "tf32"
M, N, K, nm = 1024, 30000, 256, 100
h = torch.randn(M, K, device='cuda'); W = torch.randn(N, K, device='cuda') / 16; b = torch.randn(N, device='cuda')
n = torch.full((M,), nm, dtype=torch.int32, device='cuda')
out = torch.empty(M, N, device='cuda')
for it in range(6):
    ops.linear_fwd([h], W, b, smask=(n, nm, 3), out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for it in range(10):
    ops.linear_fwd([h], W, b, smask=(n, nm, 3), out=out)
e1.record(); torch.cuda.synchronize()
print("layer2 us/launch", e0.elapsed_time(e1) * 100)
