#!/usr/bin/env python
"""Host-link yardstick: what this box's PCIe / host memory does with pinned buffers and plain
cudaMemcpyAsync, no kernels -- the ceiling of bench.py's end-to-end (`e2e`) leg.

  python tools/hostlink_yardstick.py                       one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
         --master-port 29511 tools/hostlink_yardstick.py   N GPUs at the same time

Per rank: H2D alone, D2H alone, both directions at once on two streams (the shape of the e2e leg:
8 B/row up, 12 B per kept row down), for a few chunk sizes.  Device-side timing (CUDA events), barrier
on both sides, the slowest rank counts.  One JSON line per measurement on rank 0.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--buf-mb", type=int, default=1024, help="pinned buffer per direction")
    ap.add_argument("--up-gb", type=float, default=8.0)
    ap.add_argument("--down-gb", type=float, default=18.0)
    ap.add_argument("--chunks-mb", default="16,64,256,1024")
    ap.add_argument("--numa", type=int, default=1, help="bind to the CPUs NVML names for the GPU before allocating")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bound = None
    if args.numa:
        import bench
        bound = bench.bind_to_gpu_numa_node(local)
    dist = None
    if world > 1:
        import torch.distributed as d
        d.init_process_group("nccl", device_id=dev)
        dist = d
    nbytes = args.buf_mb << 20
    h_up = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    h_down = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    h_up.fill_(1)
    d_up = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    d_down = torch.ones(nbytes, dtype=torch.uint8, device=dev)
    s_up, s_down = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(up_bytes, down_bytes, chunk):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        ev[0].record(s_up)
        ev[2].record(s_down)
        off = 0
        done_up = done_down = 0
        while done_up < up_bytes or done_down < down_bytes:       # interleave the enqueues like the streamed leg does
            n = min(chunk, nbytes - off)
            if done_up < up_bytes:
                with torch.cuda.stream(s_up):
                    d_up[off:off + n].copy_(h_up[off:off + n], non_blocking=True)
                done_up += n
            if done_down < down_bytes:
                with torch.cuda.stream(s_down):
                    h_down[off:off + n].copy_(d_down[off:off + n], non_blocking=True)
                done_down += n
            off = (off + n) % nbytes
        ev[1].record(s_up)
        ev[3].record(s_down)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        t_up, t_down = ev[0].elapsed_time(ev[1]), ev[2].elapsed_time(ev[3])
        t = torch.tensor([t_up, t_down, max(t_up, t_down)], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(x) for x in t.tolist()], done_up, done_down

    for chunk_mb in [int(x) for x in args.chunks_mb.split(",")]:
        chunk = chunk_mb << 20
        for mode, up, down in (("h2d", args.up_gb, 0.0), ("d2h", 0.0, args.down_gb), ("both", args.up_gb, args.down_gb),
                               ("both_sel_only", args.up_gb, args.down_gb / 3.0)):
            run(int(min(up, 1.0) * 2 ** 30), int(min(down, 1.0) * 2 ** 30), chunk)          # warm-up
            (t_up, t_down, t_all), bu, bd = run(int(up * 2 ** 30), int(down * 2 ** 30), chunk)
            if rank == 0:
                print(json.dumps({
                    "yardstick": "hostlink", "mode": mode, "n_gpus": world, "chunk_mb": chunk_mb, "cpus_bound_near_gpu": bound,
                    "h2d_gb_per_s_per_gpu": bu / (t_up * 1e-3) / 1e9 if bu else None,
                    "d2h_gb_per_s_per_gpu": bd / (t_down * 1e-3) / 1e9 if bd else None,
                    "ms": t_all, "h2d_bytes": bu, "d2h_bytes": bd,
                    "aggregate_gb_per_s": world * (bu + bd) / (t_all * 1e-3) / 1e9}), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
