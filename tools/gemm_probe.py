"""GEMM experiments on the c3 shapes: per-launch time of each dense-layer role in tf32 / tf32x3, and
(GSC_TC_DBG) ablations of actor layer 2 that switch off the MMAs, the TMA loads, the epilogue or the
hi/lo conversion to see which of them the main loop waits for."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gsc_b200 import ops

M, A, nm = 1024, 30000, 100
dev = "cuda"
h = torch.relu(torch.randn(M, 256, device=dev)); W2 = torch.randn(A, 256, device=dev) / 16; b2 = torch.randn(A, device=dev)
n = torch.full((M,), nm, dtype=torch.int32, device=dev)
emb = torch.randn(M, 24, device=dev); act = torch.rand(M, A, device=dev)
W1 = torch.randn(64, 24 + 2 * A, device=dev) / 100; b1 = torch.randn(64, device=dev)
dh = torch.randn(M, 64, device=dev); dy = torch.randn(M, A, device=dev)
gW2 = torch.empty(A, 256, device=dev); gW1 = torch.empty(64, A, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = torch.empty(M, A, device=dev)


def timeit(fn, reps=8):
    for _ in range(3):
        flush.zero_(); fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts)[len(ts) // 2]


cases = {
    "actor_l2 fwd [1024x30000x256] smask": lambda: ops.linear_fwd([h], W2, b2, smask=(n, nm, 3), out=out),
    "critic_l1 fwd [1024x64x(24+30000)] table": None,
    "dgrad action [1024x30000x64] smask": lambda: ops.linear_dgrad(dh, W1, 24 + A, A, smask=(n, nm, 3), out=out),
    "actor_l2 dgrad [1024x256x30000] gate": lambda: ops.linear_dgrad(dy, W2, 0, 256, gate=h),
    "actor_l2 wgrad [30000x256x1024]": lambda: ops.linear_wgrad(dy, [h], gW2),
    "critic_l1 wgrad [64x30000x1024]": lambda: ops.linear_wgrad(dh, [act], gW1),
}
for prec in ("tf32", "tf32x3", "fp32"):
    ops.precision = prec
    table = ops.mask_table(W1[:, 24:24 + A], nm, 3)
    cases["critic_l1 fwd [1024x64x(24+30000)] table"] = lambda: ops.linear_fwd(
        [emb, act], W1, b1, relu=True, col_offsets=[0, 24 + A], row_table=(table, n))
    for name, fn in cases.items():
        if prec == "fp32" and "actor_l2 fwd" not in name:
            continue
        print(f"{prec:7s} {name:45s} {timeit(fn):8.1f} us", flush=True)

names = {0: "full", 1: "no MMAs", 2: "no TMA loads", 4: "no epilogue", 5: "no MMAs, no epilogue", 6: "no loads, no epilogue",
         8: "no conversion", 12: "no conversion, no epilogue", 10: "no loads, no conversion", 14: "MMAs only"}
which = sys.argv[1:] or ["actor_l2 fwd"]
names[32] = "no TMA stores"
outp = torch.empty(M, A + 16, device=dev)[:, :A]
cases["dgrad action, ld 30016"] = lambda: ops.linear_dgrad(dh, W1, 24 + A, A, smask=(n, nm, 3), out=outp)
cases["actor_l2 fwd, ld 30016"] = lambda: ops.linear_fwd([h], W2, b2, smask=(n, nm, 3), out=outp)
print(f"fill_ [1024x30000] fp32 {timeit(lambda: out.fill_(1.0)):8.1f} us;  copy_ {timeit(lambda: out.copy_(dy)):8.1f} us", flush=True)
ops.precision = "tf32x3"
table = ops.mask_table(W1[:, 24:24 + A], nm, 3)
for name, fn in cases.items():
    if not any(w in name for w in which):
        continue
    for prec in ("tf32", "tf32x3"):
        ops.precision = prec
        for dbg, label in names.items():
            if prec == "tf32" and dbg & 8:
                continue
            os.environ["GSC_TC_DBG"] = str(dbg)
            print(f"ablate {prec:7s} {name[:28]:28s} dbg={dbg:2d} ({label:28s}) {timeit(fn):8.1f} us", flush=True)
os.environ["GSC_TC_DBG"] = "0"
