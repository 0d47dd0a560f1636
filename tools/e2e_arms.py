#!/usr/bin/env python
"""is the e2e leg's 1.5x at 8 GPUs the box or the code?

Runs under the same launch as bench.py (python -m torch.distributed.run --nproc-per-node N ...).
Every rank moves the ENC step's e2e traffic (135 MB host->device + 135 MB device->host, both
directions at once on two streams, all ranks at the same time after a barrier) through three
staging arrangements and reports GB/s per rank and direction:
  (i)   as bench.py: torch pin_memory buffers, one cudaMemcpyAsync per tensor (4 + 4)
  (ii)  rank bound to its GPU's NUMA-local cores (NVML ideal CPU affinity) BEFORE the buffers are
        allocated, host->device source allocated write-combined (cudaHostAllocWriteCombined)
  (iii) (i) with ONE packed staging buffer per direction (1 + 1 copies)
No batched-memcpy API is used.  Rank 0 prints a markdown table (profiles/r2_e2e_<N>gpu.md)."""
import ctypes
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SIZES = {"value": 16 * 1500 * 8 * 64 * 4, "loc": 16 * 1500 * 8 * 4 * 4 * 2 * 4, "aw": 16 * 1500 * 8 * 16 * 4,
         "grad_out": 16 * 1500 * 512 * 4}          # = the four tensors of an ENC step, 135.2 MB


def host_alloc(nbytes, write_combined):
    rt = ctypes.CDLL("libcudart.so.12")
    p = ctypes.c_void_p()
    flags = 4 if write_combined else 0            # cudaHostAllocWriteCombined
    err = rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(nbytes), ctypes.c_uint(flags))
    assert err == 0, err
    buf = (ctypes.c_uint8 * nbytes).from_address(p.value)
    return torch.frombuffer(buf, dtype=torch.uint8)


def bind_numa_local(index):
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]
        if cpus:
            os.sched_setaffinity(0, cpus)
        return "%d cpus (%d-%d)" % (len(cpus), cpus[0], cpus[-1]) if cpus else "none"
    except Exception as exc:  # noqa: BLE001
        return "unavailable: %r" % (exc,)


def run_arm(dev, h2d_pairs, d2h_pairs, iters, sync):
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for it in range(2 + iters):
        if it == 2:
            sync()
            e[0].record(s_in)
            e[2].record(s_out)
        with torch.cuda.stream(s_in):
            for d, h in h2d_pairs:
                d.copy_(h, non_blocking=True)
        with torch.cuda.stream(s_out):
            for h, d in d2h_pairs:
                h.copy_(d, non_blocking=True)
    e[1].record(s_in)
    e[3].record(s_out)
    torch.cuda.synchronize()
    nb = sum(h.numel() for _, h in h2d_pairs)
    return nb * iters / (e[0].elapsed_time(e[1]) * 1e-3) / 1e9, nb * iters / (e[2].elapsed_time(e[3]) * 1e-3) / 1e9


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    total = sum(SIZES.values())
    d_in = {k: torch.empty(n, dtype=torch.uint8, device=dev) for k, n in SIZES.items()}
    d_out = {k: torch.empty(n, dtype=torch.uint8, device=dev) for k, n in SIZES.items()}
    res = {}
    # (i)
    h_in = {k: torch.empty(n, dtype=torch.uint8).pin_memory() for k, n in SIZES.items()}
    h_out = {k: torch.empty(n, dtype=torch.uint8).pin_memory() for k, n in SIZES.items()}
    res["(i) torch pinned, 4+4 copies"] = run_arm(dev, [(d_in[k], h_in[k]) for k in SIZES],
                                                   [(h_out[k], d_out[k]) for k in SIZES], 10, sync)
    # (iii)
    hp_in, hp_out = torch.empty(total, dtype=torch.uint8).pin_memory(), torch.empty(total, dtype=torch.uint8).pin_memory()
    dp_in, dp_out = torch.empty(total, dtype=torch.uint8, device=dev), torch.empty(total, dtype=torch.uint8, device=dev)
    res["(iii) one packed buffer per direction"] = run_arm(dev, [(dp_in, hp_in)], [(hp_out, dp_out)], 10, sync)
    del h_in, h_out, hp_in, hp_out
    # (ii)
    aff = bind_numa_local(local)
    w_in = host_alloc(total, True)
    w_out = host_alloc(total, False)
    w_in.fill_(1)
    res["(ii) NUMA-local binding (%s) + write-combined source, packed" % aff] = run_arm(
        dev, [(dp_in, w_in)], [(w_out, dp_out)], 10, sync)
    rows = []
    for name, (a, b) in res.items():
        t = torch.tensor([a, b, -a, -b], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        rows.append((name, float(t[0]), -float(t[2]), float(t[1]), -float(t[3])))
    if rank == 0:
        print("| arm (%d ranks at once, %.1f MB each way per step) | H2D GB/s per rank min / max | D2H GB/s per rank min / max |" % (world, total / 1e6))
        print("|---|---|---|")
        for name, a0, a1, b0, b1 in rows:
            print("| %s | %.1f / %.1f | %.1f / %.1f |" % (name, a0, a1, b0, b1))
        print(json.dumps({"n_gpus": world, "arms": {r[0]: {"h2d_min": r[1], "h2d_max": r[2], "d2h_min": r[3], "d2h_max": r[4]} for r in rows}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
