"""Can torch.distributed NCCL collectives be captured into a CUDA graph on this stack?  (2 ranks)"""
import os, sys, time, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
x = torch.full((1 << 20,), float(rank + 1), device=dev)
dist.all_reduce(x); torch.cuda.synchronize()
print(rank, "eager ok", float(x[0]), flush=True)
mode = sys.argv[1] if len(sys.argv) > 1 else "default"
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for _ in range(2):
        y = x * 0.5
        dist.all_reduce(y)
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
kw = {} if mode == "default" else {"capture_error_mode": mode}
t0 = time.time()
with torch.cuda.graph(g, **kw):
    y = x * 0.5
    dist.all_reduce(y)
    z = y + 1
print(rank, "captured", round(time.time() - t0, 2), flush=True)
for it in range(3):
    g.replay()
torch.cuda.synchronize()
print(rank, "replayed", float(z[0]), flush=True)
dist.barrier(); torch.cuda.synchronize()
print(rank, "barrier ok", flush=True)
dist.destroy_process_group()
