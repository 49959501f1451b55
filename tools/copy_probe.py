"""Copy-only ceiling of the e2e leg: every rank moves one step's worth of pinned host traffic (78.6 MB host->device,
19.7 MB device->host for 256 frames of 640x480) with no kernels, all ranks at once.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/copy_probe.py
Prints one JSON line (rank 0): per-mode GB/s per rank (min over ranks) and the frames/s those rates would allow."""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
torch.cuda.set_device(local)
dev = torch.device(f"cuda:{local}")
B, W, H, CAP = 256, 640, 480, 1000
h_in = torch.zeros(B * W * H, dtype=torch.uint8).pin_memory()
h_out = torch.zeros(B * CAP * (28 + 32 + 17) + 4 * B, dtype=torch.uint8).pin_memory()
d_in, d_out = torch.zeros_like(h_in, device=dev), torch.zeros_like(h_out, device=dev)
s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
res = {}
for mode in ("h2d", "d2h", "both"):
    for timed in (False, True):
        reps = 30 if timed else 5
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s_in):
                    d_in.copy_(h_in, non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s_out):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    nbytes = (h_in.numel() if mode in ("h2d", "both") else 0) + (h_out.numel() if mode in ("d2h", "both") else 0)
    res[mode] = {"gbs_per_rank": nbytes * reps / dt / 1e9, "steps_per_s_per_rank": reps / dt, "frames_per_s_all_ranks": world * B * reps / dt}
if rank == 0:
    print(json.dumps({"n_gpus": world, "h2d_bytes_per_step": h_in.numel(), "d2h_bytes_per_step": h_out.numel(), "modes": res,
                      "note": "pinned host memory, one copy stream per direction per rank, no kernels; max over ranks of the wall time"}))
if world > 1:
    dist.destroy_process_group()
