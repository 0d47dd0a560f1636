#!/usr/bin/env python
"""bench.py -- MSDeformAttn forward+backward throughput on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--workload ENC|DEC|CPU] [--dtype f32|bf16] [--dist model|uniform]

A "step" is one pass of the hot path (msda_forward_ex, zero-fill of grad_value on a side stream
under the forward, msda_backward_ex) over one batch of synthetic input of the named workload
(default ENC = BASELINE.json configs[1]: N=16, S=Lq=1500 tokens, M=8, D=64, L=4, P=4, fp32).
The timed region of `--steps` steps is repeated `--repeats` times; the line reports the median
region (`value`, `ms_per_step`) with min / max beside it.

  value     queries/s (N*Lq per step, summed over ranks) with inputs resident in HBM,
            kernels launched through the C ABI; CUDA events, max over ranks.
  e2e       same metric through the public API (MSDeformAttnFunction.apply + autograd
            backward) from pinned HOST buffers: H2D of value/loc/aw/grad_out and D2H
            of out + the three gradients inside the timed region.
  roofline  dominant kernel = the backward kernel; achieved = algorithmic bytes
             / its CUDA-event duration averaged over the timed steps.
  cpu_baseline  the oracle port of the reference's pure-PyTorch CPU path
            (oracle/core_pytorch.py, fwd + autograd bwd) on a bounded sample, rank 0, N=1 only.
  secondary (N=1) the other lines of SURVEY.md 8(d), each with kernel times, roofline fractions and
            the kernel family: DEC, bf16, uniform locations, a locality sweep (rows of noise on the
            offsets), permuted query order (default vs row windows vs per-tap), L2-warm inputs, the
            four calls of a multimodal encoder layer, and the kernel to beat (the reference's own
            CUDA op from oracle/_ref, test infrastructure, timed after the headline region).
  train_step  (any N; --no-train-step skips it) BASELINE.json configs[4]: the 6+6-layer deformable
            transformer step of tools/train_step.py with the NCCL gradient all-reduce, under this
            launch, with its own clocks record; measured last, behind a time limit, so that the
            line is printed whatever happens in it.

Multi-GPU (torchrun): batch-sharded, weak scaling (per-GPU batch fixed), no data-path
collective; barrier + synchronize around the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "msdeformattn_fwd_bwd_queries_per_sec"
UNIT = "queries/s"
CPU_SAMPLE_BATCH = 16     # batch items of the workload the CPU legs process per step
RING = 4                  # distinct input sets rotated so a step never finds its inputs in L2


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ENC", choices=["ENC", "DEC", "CPU"])
    ap.add_argument("--dtype", default="f32", choices=["f32", "bf16"])
    ap.add_argument("--dist", default="model", choices=["model", "uniform"])
    ap.add_argument("--padding", default="zeros", choices=["zeros", "border"])
    ap.add_argument("--qtile", type=int, default=0, help="tuning: queries-per-tile multiplier (0=auto)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--repeats", type=int, default=7, help="timed regions of --steps steps; the median is reported")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary records (N=1 only anyway)")
    ap.add_argument("--train-step", action="store_true", help="(default now; kept so that older command lines still parse)")
    ap.add_argument("--no-train-step", action="store_true",
                    help="skip the configs[4] training-step record (6+6-layer transformer step with the NCCL "
                         "gradient all-reduce; about half a minute)")
    ap.add_argument("--flags", type=int, default=0, help="extra MSDA_FLAG_* bits for the GPU arm (A/B aid)")
    return ap.parse_args()


TRAIN_STEP_LIMIT_S = 240   # the configs[4] leg takes ~30 s; see run_ours


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return None
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 - 0.05 <= t <= t1 + 0.15 and len(r) >= 6] or \
               [r for _, r in self.rows if len(r) >= 6]
        if not rows:
            return None
        sm = sorted(int(float(r[0])) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": int(float(rows[0][1])), "reasons": reasons,
                "samples": len(rows)}


# ----------------------------------------------------------------------------- CPU legs (oracle port)
def workload_string(workload, N, Lq, S, T, M=8, D=64, L=4, P=4):
    return "%s: MSDeformAttn fwd+bwd, per-GPU batch %d, Lq=%d, S=%d (T=%s), M=%d, D=%d, L=%d, P=%d" % (
        workload, N, Lq, S, T, M, D, L, P)


def cpu_leg(workload, steps, warmup, dist, padding="zeros"):
    """Times the oracle port of the reference's ms_deform_attn_core_pytorch (fwd + autograd
    bwd, fp32, all host threads) on a bounded sample: CPU_SAMPLE_BATCH batch items of the
    workload per step.  Returns (queries/s, ms_per_step, cores, sample description)."""
    import torch
    from mfl_b200 import synthetic
    from oracle.core_pytorch import ms_deform_attn_core_pytorch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    nb = min(CPU_SAMPLE_BATCH, synthetic.WORKLOADS[workload]["N"])
    inp = synthetic.make_workload(workload, torch.float32, "cpu", seed=1234, dist=dist, value="normal", N=nb)
    v, loc, aw = (inp[k].requires_grad_(True) for k in ("value", "loc", "aw"))
    Lq = inp["loc"].shape[1]

    def step():
        out = ms_deform_attn_core_pytorch(v, inp["shapes"], loc, aw, padding)
        out.backward(inp["grad_out"])
        v.grad = loc.grad = aw.grad = None

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    sample = "%s workload cut to batch %d (Lq=%d), fwd+autograd bwd, fp32, grid_sample padding=%s, %d threads, %d steps" % (
        workload, nb, Lq, padding, cores, steps)
    return nb * Lq / dt, dt * 1e3, cores, sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # every step is a bounded sample (CPU_SAMPLE_BATCH batch items, ~0.4 s at ENC); --steps / --warmup are
    # honoured up to a cap that keeps the run within minutes
    steps, warmup = min(args.steps, 50), max(1, min(args.warmup, 10))
    qps, ms, cores, sample = cpu_leg(args.workload, steps, warmup, args.dist, args.padding)
    from mfl_b200 import synthetic
    w = synthetic.WORKLOADS[args.workload]
    S = sum(w["T"])
    Lq = w["Lq"] if w["Lq"] is not None else sum(w.get("Tq", w["T"]))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": qps, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(args.workload, w["N"], Lq, S, w["T"]), "locations": args.dist,
                   "padding": args.padding,
                   "parallelism": "batch-sharded x%d, no data-path collective" % args.gpus,
                   "l2": "n/a (host memory)",
                   "impl_detail": "oracle port of the reference's ms_deform_attn_core_pytorch (grid_sample, padding=%s) "
                                  "on the host cores; each step = a sample of %d batch items" % (
                                      args.padding, min(CPU_SAMPLE_BATCH, w["N"]))},
        "cpu_baseline": {"value": qps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": qps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ----------------------------------------------------------------------------- GPU arm
class OpBuffers:
    """One input set + preallocated outputs; launches the hot path through the C ABI."""

    def __init__(self, inp, padding, qtile):
        import torch
        from mfl_b200 import _lib
        self.lib = _lib.load()
        self.inp = inp
        v, loc, aw = inp["value"], inp["loc"], inp["aw"]
        self.N, self.S, self.M, self.D = v.shape
        self.L, self.Lq, self.P = inp["shapes"].shape[0], loc.shape[1], loc.shape[4]
        self.dt = {torch.float32: 0, torch.float64: 1, torch.bfloat16: 2}[v.dtype]
        self.pad = _lib.PADDING_MODES[padding]
        self.flags = _lib.flag_qtile(qtile)
        self.skip = _lib.MSDA_FLAG_SKIP_ZERO_FILL
        self.out = torch.empty(self.N, self.Lq, self.M * self.D, dtype=v.dtype, device=v.device)
        self.gv, self.gl, self.ga = torch.empty_like(v), torch.empty_like(loc), torch.empty_like(aw)
        self.dims = (self.N, self.S, self.M, self.D, self.L, self.Lq, self.P)
        self.set_flags(self.flags)
        self.check = _lib.check

    def set_flags(self, flags):
        """(Re)size the workspaces for a flags word: the forward's plan (time order of the queries,
        reused by the backward) and the backward's accumulation image (bf16 only)."""
        import torch
        self.flags = flags
        dev = self.out.device
        self.plan_bytes = self.lib.msda_forward_workspace_bytes(*self.dims, self.dt, flags)
        self.plan = torch.empty(max(self.plan_bytes // 4, 1), dtype=torch.int32, device=dev)
        # the backward's REQUIRED workspace only (an fp32 image of grad_value for bf16); its optional
        # tail, room for the plan, is not needed: the plan comes from the forward
        self.ws_bytes = self.N * self.S * self.M * self.D * 4 if self.dt == 2 else 0
        self.ws = torch.empty(max(self.ws_bytes // 4, 1), dtype=torch.float32, device=dev)

    def fwd(self, stream):
        i = self.inp
        self.check(self.lib.msda_forward_ex(i["value"].data_ptr(), i["shapes"].data_ptr(), i["lsi"].data_ptr(),
                                            i["loc"].data_ptr(), i["aw"].data_ptr(), self.out.data_ptr(),
                                            *self.dims, self.dt, self.pad, self.flags,
                                            self.plan.data_ptr() if self.plan_bytes else None, self.plan_bytes,
                                            stream))

    def zero(self):
        (self.ws if self.ws_bytes else self.gv).zero_()

    def bwd(self, stream):
        i = self.inp
        self.check(self.lib.msda_backward_ex(
            i["value"].data_ptr(), i["shapes"].data_ptr(), i["lsi"].data_ptr(), i["loc"].data_ptr(),
            i["aw"].data_ptr(), i["grad_out"].data_ptr(), self.gv.data_ptr(), self.gl.data_ptr(),
            self.ga.data_ptr(), *self.dims, self.dt, self.pad, self.flags | self.skip,
            self.ws.data_ptr() if self.ws_bytes else None, self.ws_bytes,
            self.plan.data_ptr() if self.plan_bytes else None, stream))


def time_device(bufs, steps, warmup, dist_sync, overlap_zero=True):
    """Timed region: `steps` x (fwd, zero-fill of the accumulation target, bwd) rotating over the
    input ring.  overlap_zero: the zero-fill of step i is enqueued on a side stream when the step
    starts and joined before the backward -- it runs under the forward kernels, inside the timed
    step (what MSDeformAttnFunction does, functions/ms_deform_attn_func.py).
    Returns (total_ms, fwd_ms_avg, zero_ms_avg (0 when overlapped), bwd_ms_avg)."""
    import torch
    main = torch.cuda.current_stream()
    st = main.cuda_stream
    side = torch.cuda.Stream()
    ev_go = [torch.cuda.Event() for _ in range(steps + warmup)]
    ev_z = [torch.cuda.Event() for _ in range(steps + warmup)]

    def step(i, ev=None):
        b = bufs[i % len(bufs)]
        if ev:
            ev[0].record()
        if overlap_zero:
            ev_go[i].record(main)
            side.wait_event(ev_go[i])           # not before the step has started
            with torch.cuda.stream(side):
                b.zero()
                ev_z[i].record(side)
        b.fwd(st)
        if ev:
            ev[1].record()
        if overlap_zero:
            main.wait_event(ev_z[i])
        else:
            b.zero()
        if ev:
            ev[2].record()
        b.bwd(st)
        if ev:
            ev[3].record()

    for i in range(warmup):
        step(i)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(steps)]
    dist_sync()
    for i in range(steps):
        step(warmup + i, ev[i])
    torch.cuda.synchronize()
    dist_sync()
    total = ev[0][0].elapsed_time(ev[-1][3])
    seg = [sum(e[k].elapsed_time(e[k + 1]) for e in ev) / steps for k in range(3)]
    return total, seg[0], seg[1], seg[2]


def time_e2e(host, dev_meta, padding, steps, warmup, dist_sync):
    """Public-API path from pinned host buffers.  Every step copies its inputs host->device,
    runs MSDeformAttnFunction (apply + autograd backward) and copies out + the three gradients
    device->host; all of it inside the timed region.  The three stages run on three streams
    over two buffer slots, so step i's D2H overlaps step i+1's H2D (PCIe is full duplex) and
    both overlap the kernels -- the way a caller that streams batches would drive the op."""
    import torch
    from mfl_b200.functions.ms_deform_attn_func import ms_deform_attn
    dev = dev_meta["shapes"].device
    keys = ("value", "loc", "aw", "grad_out")
    NB = 2
    comp = torch.cuda.current_stream()
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    dev_in = [{k: torch.empty(host[k].shape, dtype=host[k].dtype, device=dev) for k in keys} for _ in range(NB)]
    res_host = [None] * NB
    ev_in = [torch.cuda.Event() for _ in range(NB)]        # inputs of the slot are on the device
    ev_comp = [torch.cuda.Event() for _ in range(NB)]      # kernels of the slot are done
    ev_out = [torch.cuda.Event() for _ in range(NB)]       # results of the slot are on the host
    pending = [None] * NB

    def h2d(i):
        slot = i % NB
        with torch.cuda.stream(s_in):
            if i >= NB:
                s_in.wait_event(ev_comp[slot])             # the slot's previous inputs were consumed
            for k in keys:
                dev_in[slot][k].copy_(host[k], non_blocking=True)
            ev_in[slot].record(s_in)

    def compute(i):
        slot = i % NB
        comp.wait_event(ev_in[slot])
        v, loc, aw = (dev_in[slot][k].detach().requires_grad_(True) for k in ("value", "loc", "aw"))
        out = ms_deform_attn(v, dev_meta["shapes"], dev_meta["lsi"], loc, aw, 64, padding)
        out.backward(dev_in[slot]["grad_out"])
        ev_comp[slot].record(comp)
        pending[slot] = (out.detach(), v.grad, loc.grad, aw.grad)

    def d2h(i):
        slot = i % NB
        res = pending[slot]
        if res_host[slot] is None:
            res_host[slot] = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in res]
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_comp[slot])
            for h, t in zip(res_host[slot], res):
                t.record_stream(s_out)
                h.copy_(t, non_blocking=True)
            ev_out[slot].record(s_out)

    def run(n):
        h2d(0)
        for i in range(n):
            if i + 1 < n:
                h2d(i + 1)
            compute(i)
            d2h(i)
        comp.wait_stream(s_out)
        comp.wait_stream(s_in)

    run(max(warmup, NB))
    torch.cuda.synchronize()
    dist_sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s_in.wait_stream(comp)
    s_out.wait_stream(comp)
    run(steps)
    e1.record()
    torch.cuda.synchronize()
    dist_sync()
    h2d_b = sum(host[k].numel() * host[k].element_size() for k in keys)
    d2h_b = sum(t.numel() * t.element_size() for t in res_host[0])
    return e0.elapsed_time(e1), h2d_b, d2h_b


def measure(ring, padding, flags, steps, warmup, repeats, dist_sync, hbm_peak, overlap_zero=True):
    """Kernel times of one workload: `repeats` timed regions of `steps` steps -> medians."""
    from mfl_b200 import synthetic
    bufs = [OpBuffers(i, padding, 0) for i in ring]
    for b in bufs:
        b.set_flags(b.flags | flags)
    b0 = bufs[0]
    esz = ring[0]["value"].element_size()
    bf, bb = synthetic.algorithmic_bytes(b0.N, b0.S, b0.Lq, b0.M, b0.D, b0.L, b0.P, esz)
    regs = [time_device(bufs, steps, warmup if r == 0 else 1, dist_sync, overlap_zero) for r in range(repeats)]
    med = lambda xs: sorted(xs)[len(xs) // 2]
    step = [r[0] / steps for r in regs]
    fwd, zero, bwd = (med([r[k] for r in regs]) for k in (1, 2, 3))
    fam = b0.lib.msda_kernel_choice(b0.N, b0.S, b0.M, b0.D, b0.L, b0.Lq, b0.P, b0.dt, b0.flags).decode()
    gb = lambda byts, ms: byts / (ms * 1e-3) / 1e9
    return {"family": fam, "fwd_ms": fwd, "bwd_ms": bwd, "zero_fill_ms": zero, "step_ms": med(step),
            "step_ms_min": min(step), "step_ms_max": max(step),
            "fwd_frac": gb(bf, fwd) / hbm_peak, "bwd_frac": gb(bb, bwd) / hbm_peak,
            "step_frac": gb(bf + bb, med(step)) / hbm_peak, "bytes_fwd": bf, "bytes_bwd": bb,
            "queries": b0.N * b0.Lq}


def time_kernel_to_beat(ring, steps, warmup):
    """The reference's own CUDA op (oracle/_ref, built from /root/reference by oracle/build_ref.py:
    test infrastructure) on the same ring of inputs: fwd / bwd ms through its pybind entry points
    (allocation and zero-fill inside, as the reference does)."""
    import torch
    from oracle import ref_cuda
    if not ref_cuda.available():
        return None
    ref = ref_cuda.load()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
    for i in range(warmup + steps):
        r = ring[i % len(ring)]
        a = (r["value"], r["shapes"], r["lsi"], r["loc"], r["aw"])
        k = i - warmup
        if k >= 0:
            ev[k][0].record()
        ref.ms_deform_attn_forward(*a, 64)
        if k >= 0:
            ev[k][1].record()
        ref.ms_deform_attn_backward(*a, r["grad_out"], 64)
        if k >= 0:
            ev[k][2].record()
    torch.cuda.synchronize()
    med = lambda xs: sorted(xs)[len(xs) // 2]
    return {"impl": "reference CUDA op (models/ops/src/cuda) recompiled for sm_100a, oracle/_ref",
            "fwd_ms": med([e[0].elapsed_time(e[1]) for e in ev]),
            "bwd_ms": med([e[1].elapsed_time(e[2]) for e in ev])}


def time_module_call(dev, steps, warmup, reps):
    """what sits between the module's Linear layers and its output projection
    (ms_deform_attn.py:98-122) at the ENC shape, through the C ABI: (a) front-end kernel + op forward,
    op backward + front-end backward (loc / aw through HBM); (b) msda_module_forward / backward
    (softmax and locations inside the tile kernels).  ms per forward + backward, CUDA events, median."""
    import torch
    from mfl_b200 import _lib, synthetic
    lib = _lib.load()
    w = synthetic.WORKLOADS["ENC"]
    N, M, D, P, T = w["N"], w["M"], w["D"], w["P"], w["T"]
    L, S = len(T), sum(T)
    Lq = S
    sets = []
    for r in range(2):
        g = torch.Generator(device=dev).manual_seed(777 + r)
        f = dict(device=dev, generator=g)
        ref = synthetic.encoder_reference_points(T, dev)[None, :, None, None].expand(N, -1, L, 1).contiguous()
        off = synthetic.head_bias(M, P, dev)[None, None, :, None, :] + 0.5 * torch.randn(N, Lq, M, L, P, **f)
        sets.append(dict(value=torch.randn(N, S, M, D, **f), off=off.contiguous(), logits=torch.randn(N, Lq, M, L * P, **f),
                         ref=ref, gout=torch.randn(N, Lq, M * D, **f)))
    Tt = torch.tensor(T, device=dev)
    shapes = torch.stack((torch.ones_like(Tt), Tt), -1).contiguous()
    lsi = torch.cat((Tt.new_zeros(1), Tt.cumsum(0)[:-1]))
    e = lambda *sh: torch.empty(*sh, device=dev)
    loc, aw, out = e(N, Lq, M, L, P, 2), e(N, Lq, M, L, P), e(N, Lq, M * D)
    gv, gloc, gaw = e(N, S, M, D), e(N, Lq, M, L, P, 2), e(N, Lq, M, L, P)
    goff, glog, gref = e(N, Lq, M, L, P), e(N, Lq, M, L * P), e(N, Lq, L, 1)
    dims = (N, S, M, D, L, Lq, P)
    nb = lib.msda_forward_workspace_bytes(*dims, 0, 0)
    plan = torch.empty(max(nb // 4, 1), dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    ck = _lib.check
    p = lambda t: t.data_ptr()

    def two_kernel(c):
        ck(lib.msda_frontend_forward(p(c["off"]), p(c["logits"]), p(c["ref"]), p(Tt), p(loc), p(aw), N, Lq, M, L, P, 1, 0, st))
        ck(lib.msda_forward_ex(p(c["value"]), p(shapes), p(lsi), p(loc), p(aw), p(out), *dims, 0, 0, 0, p(plan), nb, st))
        ck(lib.msda_backward_ex(p(c["value"]), p(shapes), p(lsi), p(loc), p(aw), p(c["gout"]), p(gv), p(gloc), p(gaw),
                                *dims, 0, 0, 0, None, 0, p(plan), st))
        ck(lib.msda_frontend_backward(p(c["off"]), p(aw), p(c["ref"]), p(Tt), p(gloc), p(gaw), p(goff), p(glog), p(gref),
                                      N, Lq, M, L, P, 1, 0, st))

    def fused(c):
        ck(lib.msda_module_forward(p(c["value"]), p(shapes), p(lsi), p(c["off"]), p(c["logits"]), p(c["ref"]), None, p(out),
                                   *dims, 1, 0, 0, 0, p(plan), nb, st))
        ck(lib.msda_module_backward(p(c["value"]), p(shapes), p(lsi), p(c["off"]), p(c["logits"]), p(c["ref"]), None,
                                    p(c["gout"]), p(gv), p(goff), p(glog), p(gref), *dims, 1, 0, 0, 0, None, 0, p(plan), st))

    res = {}
    for name, fn in (("two_kernel_path_ms", two_kernel), ("fused_module_op_ms", fused)):
        ts = []
        for r in range(reps):
            for i in range(warmup):
                fn(sets[i % 2])
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                fn(sets[i % 2])
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) / steps)
        res[name] = round(sorted(ts)[len(ts) // 2], 5)
    res["note"] = ("ENC f32, forward + backward of the module's core (front-end + op), zero-fill of grad_value "
                   "inside the backward (not overlapped), C ABI calls")
    return res


def secondary_records(args, dev, hbm_peak, dist_sync, ring_enc):
    """SURVEY.md 8(d) lines beside the headline; every entry is timed like the headline
    (CUDA events, inputs rotated over distinct sets unless named l2warm)."""
    import torch
    from mfl_b200 import _lib, synthetic
    out = {}
    steps, warmup, reps = 10, 3, 3

    def ring_of(wl, dtype=torch.float32, n=2, **kw):
        return [synthetic.make_workload(wl, dtype, dev, seed=4321 + r, value="normal", **kw) for r in range(n)]

    def rec(name, ring, flags=0, **kw):
        m = measure(ring, args.padding, flags, steps, warmup, reps, dist_sync, hbm_peak, **kw)
        out[name] = {k: (round(v, 5) if isinstance(v, float) else v) for k, v in m.items()
                     if k not in ("bytes_fwd", "bytes_bwd", "queries")}

    rec("ENC_f32_l2warm", ring_enc[:1])
    rec("ENC_f32_no_tile(row windows, round 1 default)", ring_enc, _lib.MSDA_FLAG_NO_TILE)
    rec("ENC_f32_per_tap", ring_enc, _lib.MSDA_FLAG_NO_WINDOW)
    rec("ENC_f32_plan_apply_split", ring_enc, _lib.MSDA_FLAG_PLAN)
    from mfl_b200 import MultiScaleDeformableAttention as MSDA
    if MSDA.saved_plans_supported():      # only in a -DMSDA_TILE_SAVED_PLANS=1 build (an experiment)
        rec("ENC_f32_saved_plan(backward starts from the forward's tile plans)", ring_enc, _lib.MSDA_FLAG_SAVE_PLAN)
    rec("ENC_f32_zero_fill_serial", ring_enc, 0, overlap_zero=False)
    out["kernel_to_beat"] = time_kernel_to_beat(ring_enc, 5, 2)
    del ring_enc
    torch.cuda.empty_cache()
    for nm, kw in (("ENC_f32_uniform", dict(dist="uniform")),
                   ("ENC_f32_sigma2", dict(dist="model", sigma=2.0)),
                   ("ENC_f32_sigma8", dict(dist="model", sigma=8.0)),
                   ("ENC_f32_sigma32", dict(dist="model", sigma=32.0))):
        rec(nm, ring_of("ENC", **kw))
    rp = ring_of("ENC", dist="model", permute=True)
    rec("ENC_f32_permuted", rp)
    rec("ENC_f32_permuted_no_tile", rp, _lib.MSDA_FLAG_NO_TILE)
    rec("ENC_f32_permuted_per_tap", rp, _lib.MSDA_FLAG_NO_WINDOW)
    del rp
    rec("ENC_bf16", ring_of("ENC", torch.bfloat16, dist="model"))
    rec("DEC_f32", ring_of("DEC", dist="model"))
    rec("DEC_bf16", ring_of("DEC", torch.bfloat16, dist="model"))
    for wl in ("MM_VV", "MM_AA", "MM_AV", "MM_VA"):
        rec(wl + "_f32", ring_of(wl, dist="model"))
    rec("CPU_shape_f32(configs[0] on the GPU)", ring_of("CPU", dist="model"))
    torch.cuda.empty_cache()
    out["ENC_f32_module_call"] = time_module_call(dev, steps, warmup, reps)
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from mfl_b200 import synthetic

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if not args.no_train_step:
            # the training-step record also replays the step as a CUDA graph with the all-reduce inside:
            # NCCL's watchdog polls events, which is illegal during a capture
            os.environ["TORCH_NCCL_ASYNC_ERROR_HANDLING"] = "0"
        dist.init_process_group("nccl", device_id=dev)

    def dist_sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    dtype = {"f32": torch.float32, "bf16": torch.bfloat16}[args.dtype]
    w = synthetic.WORKLOADS[args.workload]
    ring = [synthetic.make_workload(args.workload, dtype, dev, seed=1234 + 17 * rank + r, dist=args.dist,
                                    value="normal") for r in range(RING)]
    ring_bytes = sum(sum(t.numel() * t.element_size() for t in (i["value"], i["loc"], i["aw"], i["grad_out"]))
                     for i in ring)
    peaks, peak_src = None, "fallback"
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak_src = "measured"
    except Exception:
        pass
    hbm_peak = float(peaks["hbm_gbs"]) if peaks else 6650.0

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    t0 = time.time()
    m = measure(ring, args.padding, args.flags | _flag_qtile(args.qtile), args.steps, args.warmup,
                max(1, args.repeats), dist_sync, hbm_peak)
    t1 = time.time()

    from mfl_b200 import sharding

    def max_over_ranks(x):
        return sharding.max_over_ranks(x, dev)

    N, Lq, S = ring[0]["value"].shape[0], ring[0]["loc"].shape[1], ring[0]["value"].shape[1]
    M, D = ring[0]["value"].shape[2], ring[0]["value"].shape[3]
    L, P = ring[0]["loc"].shape[3], ring[0]["loc"].shape[4]
    ms_per_step = max_over_ranks(m["step_ms"])
    value = world * N * Lq / (ms_per_step * 1e-3)

    e2e = None
    if not args.no_e2e:
        host = {k: ring[0][k].detach().cpu().pin_memory() for k in ("value", "loc", "aw", "grad_out")}
        e_steps = max(3, min(args.steps, 20))
        e_ms, h2d, d2h = time_e2e(host, ring[0], args.padding, e_steps, 3, dist_sync)
        e_ms = max_over_ranks(e_ms) / e_steps
        e2e = {"value": world * N * Lq / (e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": e_ms, "steps": e_steps}
    clocks = sampler.stop(t0, time.time()) if rank == 0 else None

    gb = lambda byts, ms: byts / (ms * 1e-3) / 1e9
    kname = {"tile": "tile", "window": "win", "fast": "fast", "generic": "generic"}.get(m["family"], m["family"])
    roofline = {"bound": "hbm", "kernel": "msda_bwd_%s_kernel" % kname, "achieved": gb(m["bytes_bwd"], m["bwd_ms"]),
                "peak": hbm_peak,
                "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s",
                "unit": "GB/s", "frac": m["bwd_frac"], "traffic": None,
                "algorithmic_bytes_per_launch": m["bytes_bwd"], "kernel_ms": m["bwd_ms"],
                "fwd_kernel": {"name": "msda_sort_cluster_kernel + msda_fwd_%s_kernel" % kname if kname == "tile"
                               else "msda_fwd_%s_kernel" % kname,
                               "achieved": gb(m["bytes_fwd"], m["fwd_ms"]), "frac": m["fwd_frac"],
                               "kernel_ms": m["fwd_ms"], "algorithmic_bytes_per_launch": m["bytes_fwd"]},
                "zero_fill_ms": m["zero_fill_ms"],
                "zero_fill": "cudaMemsetAsync of grad_value on a side stream, enqueued at step start, joined "
                             "before the backward (inside the timed step)",
                "step": {"achieved": gb(m["bytes_fwd"] + m["bytes_bwd"], ms_per_step),
                         "frac": gb(m["bytes_fwd"] + m["bytes_bwd"], ms_per_step) / hbm_peak}}
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get("%s_%s_%s_bwd" % (args.workload, args.dtype, args.dist))
        except Exception:
            pass

    secondary = None
    if rank == 0 and world == 1 and not args.no_secondary and args.workload == "ENC" and args.dtype == "f32":
        secondary = secondary_records(args, dev, hbm_peak, dist_sync, ring)
    del ring
    torch.cuda.empty_cache()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        qps, ms, cores, sample = cpu_leg(args.workload, 3, 1, args.dist, args.padding)
        cpu = {"value": qps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "ms_per_step": ms}

    line = None
    if rank == 0:
        launches_per_step = {"tile": 3, "window": 2, "fast": 2, "generic": 2}.get(m["family"], 2)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": workload_string(args.workload, N, Lq, S, w["T"], M, D, L, P),
                       "locations": args.dist, "padding": args.padding,
                       "parallelism": "batch-sharded x%d, no data-path collective" % world,
                       "l2": "inputs rotate over %d distinct sets (%.0f MB) > 126 MB L2" % (RING, ring_bytes / 1e6)},
            "repeats": {"regions": max(1, args.repeats), "steps_per_region": args.steps,
                        "ms_per_step_median": m["step_ms"], "ms_per_step_min": m["step_ms_min"],
                        "ms_per_step_max": m["step_ms_max"]},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": launches_per_step * args.steps * max(1, args.repeats), "clocks": clocks,
            "secondary": secondary, "train_step": None,
        }

    # BASELINE.json configs[4]: the training step with its gradient all-reduce, last, on every rank.
    # Everything above is already measured: if this leg does not come back within
    # TRAIN_STEP_LIMIT_S (a rank lost inside a collective), rank 0 still prints the line, with the
    # reason in train_step, and every rank leaves.
    if not args.no_train_step:
        def give_up():
            if rank == 0:
                line["train_step"] = {"error": "no result within %d s" % TRAIN_STEP_LIMIT_S}
                print(json.dumps(line), flush=True)
            os._exit(0)
        guard = threading.Timer(TRAIN_STEP_LIMIT_S, give_up)
        guard.daemon = True
        guard.start()
        tsampler = ClockSampler(local)
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import train_step
            if rank == 0:
                tsampler.start()
            tt0 = time.time()
            train = train_step.bench_record(dev, world, rank)
            if rank == 0:
                train["clocks"] = tsampler.stop(tt0, time.time())
        except Exception as exc:  # the headline line must not depend on this leg
            train = {"error": "%s: %s" % (type(exc).__name__, exc)}
            if world > 1:
                # the other ranks may be inside a collective this rank will never join
                if rank == 0:
                    line["train_step"] = train
                    print(json.dumps(line), flush=True)
                os._exit(0)
        guard.cancel()
        if rank == 0:
            line["train_step"] = train

    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _flag_qtile(k):
    from mfl_b200 import _lib
    return _lib.flag_qtile(k)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
