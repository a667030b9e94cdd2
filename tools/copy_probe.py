"""Bare host-to-device copy ceiling with N ranks copying at once (one process per GPU, pinned buffers, one stream per GPU, no
kernels): what the end-to-end number of bench.py can reach at most.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/copy_probe.py
prints one JSON line from rank 0: per-GPU and aggregate GB/s (max-over-ranks time)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from flink_jpmml_b200 import dist as D

rank, world, local = D.env_rank()
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist = D.init("nccl", dev)
nbytes = 1 << 30
h = torch.empty(nbytes, dtype=torch.uint8).pin_memory(); d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
o = torch.empty(nbytes // 64, dtype=torch.uint8).pin_memory()
st = torch.cuda.Stream(device=dev)
def run(reps, chunk):
    with torch.cuda.stream(st):
        for _ in range(reps):
            for lo in range(0, nbytes, chunk):
                d[lo:lo + chunk].copy_(h[lo:lo + chunk], non_blocking=True)
            o.copy_(d[:nbytes // 64], non_blocking=True)     # the result stream back (1/64 of the volume, like 9 B of 816)
    st.synchronize()
out = {}
for chunk in (24 << 20, 256 << 20):
    run(1, chunk)
    if dist: dist.barrier()
    torch.cuda.synchronize(dev)
    t = time.perf_counter(); run(4, chunk); dt = time.perf_counter() - t
    (dt_max,) = D.max_over_ranks(dist, [dt], dev)
    out[f"chunk_{chunk >> 20}MiB"] = {"per_gpu_GBs": 4 * nbytes / dt_max / 1e9, "aggregate_GBs": world * 4 * nbytes / dt_max / 1e9}
if rank == 0:
    print(json.dumps({"n_gpus": world, "h2d_pinned": out}), flush=True)
if dist:
    dist.barrier(); dist.destroy_process_group()
