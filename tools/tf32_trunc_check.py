"""Does tcgen05 kind::tf32 ignore the 13 low mantissa bits of fp32 operands?  Single-pass TF32 GEMM with
raw fp32 in shared memory (GSC_TF32_TMA_ROUND=0) on x and on x with the low bits cleared: bit-identical
results mean the hardware truncates."""
import os, sys, torch
os.environ["GSC_TF32_TMA_ROUND"] = "0"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gsc_b200 import ops
ops.precision = "tf32"
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(1024, 256, generator=g, device="cuda"); W = torch.randn(2048, 256, generator=g, device="cuda")
tr = lambda t: (t.view(torch.int32) & -8192).view(torch.float32)
a = ops.linear_fwd([x], W, None).clone()
b = ops.linear_fwd([tr(x)], tr(W), None).clone()
ref_t = (tr(x).double() @ tr(W).double().t())
ref = (x.double() @ W.double().t())
print("raw == pre-truncated:", torch.equal(a, b), " max|a-b|", float((a - b).abs().max()))
print("err vs truncated-operand fp64:", float((a.double() - ref_t).abs().max() / ref_t.abs().max()),
      " err vs exact fp64:", float((a.double() - ref).abs().max() / ref.abs().max()))
