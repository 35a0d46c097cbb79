"""Host<->device copy bandwidth of N concurrent ranks (torchrun), with the NUMA placement bench.py uses: what the
end-to-end number of a multi-GPU run can reach at best.  Each rank copies `mb` MB host->device and device->host from
pinned buffers, both directions at once on two streams, and reports its rates; rank 0 prints the table."""
import os
import subprocess
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speech_recognition_tools_b200.sharding import bind_host_to_gpu  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
bind = os.environ.get("PROBE_BIND", "1") == "1"
old = bind_host_to_gpu(local) if bind else None
aff = sorted(os.sched_getaffinity(0))
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 272
n = mb * 1000 * 1000
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_in.fill_(1)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_out.fill_(2)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.ones(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, reps=8):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    return n / dt / 1e9


for _ in range(2):
    run(True, True, 2)
res = torch.tensor([run(True, False), run(False, True), run(True, True), float(len(aff)), float(aff[0]), float(aff[-1]),
                    1.0 if old is not None else 0.0], dtype=torch.float64, device="cuda")
if world > 1:
    allr = [torch.zeros_like(res) for _ in range(world)]
    dist.all_gather(allr, res)
else:
    allr = [res]
if rank == 0:
    try:
        print(subprocess.run("lscpu | grep -i -E 'numa|socket|^CPU\\(s\\)'", shell=True, capture_output=True, text=True).stdout)
    except Exception:
        pass
    print("bind", bind, " MB per copy", mb)
    print("rank  h2d_only  d2h_only  both(each dir)  ncpu first last rebound   [GB/s]")
    for r, t in enumerate(allr):
        v = t.cpu().tolist()
        print("%4d  %8.1f  %8.1f  %8.1f   %4d %4d %4d %d" % (r, v[0], v[1], v[2], v[3], v[4], v[5], v[6]))
    tot = sum(t[2].item() for t in allr)
    print("aggregate with both directions active: %.0f GB/s each way" % tot)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
