#!/usr/bin/env python
"""BASELINE.json configs[4]: the data-parallel training step around the op.

A compact deformable encoder-decoder of the reference's size and layer recipe
(models/deformable/unimodal_deformable_transformer.py:189-250 encoder layer,
:298-374 decoder layer; 6 + 6 layers, d_model 512, ffn 2048, 8 heads, 4 levels,
4 points, 100 event queries, T = [800, 400, 200, 100]) whose deformable attention
is this repo's MSDeformAttn module, i.e. the sm_100a kernels behind the C-ABI.
It exists to time the op in the step it serves and the 1 -> N GPU scaling of that
step; it is a measurement harness, not a model zoo: no captioning heads, no
matcher, no dataset.  Per SURVEY.md section 8(d): synthetic features, per-GPU batch 16
(weak scaling), loss = mean of the decoder outputs squared, AdamW 1e-4, gradient clip
0.1, one NCCL gradient all-reduce per step: over a flat buffer all .grad tensors are views
of (default), or DistributedDataParallel (--sync ddp).  --cuda-graph records the whole
step once and replays it.

  python tools/train_step.py [--steps 20] [--warmup 5] [--batch 16] [--layers 6]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port P tools/train_step.py --steps 20

Prints one JSON line on rank 0: samples/s over all ranks (device-timed, max over
ranks) and the share of the step's GPU time spent inside the op's kernels (torch
profiler kernel records of one extra, untimed step).  fp32 like the reference (its op
dispatches fp32/fp64 only and the training script has no autocast); the modules'
`shapes.sum() == Len_in` assertion, a host sync per call in the reference
(ms_deform_attn.py:93), is switched off after the first step.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import mfl_b200  # noqa: E402,F401
from mfl_b200 import synthetic  # noqa: E402
from mfl_b200.modules import MSDeformAttn  # noqa: E402


class EncoderLayer(nn.Module):
    """post-norm: x = LN(x + attn(x + pos)); x = LN(x + ffn(x))  (…transformer.py:228-250)"""

    def __init__(self, C, ffn, M, L, P):
        super().__init__()
        self.self_attn = MSDeformAttn(C, L, M, P)
        self.norm1 = nn.LayerNorm(C)
        self.linear1 = nn.Linear(C, ffn)
        self.linear2 = nn.Linear(ffn, C)
        self.norm2 = nn.LayerNorm(C)

    def forward(self, x, pos, ref, T, lsi):
        x = self.norm1(x + self.self_attn(x + pos, ref, x, T, lsi))
        return self.norm2(x + self.linear2(F.relu(self.linear1(x))))


class DecoderLayer(nn.Module):
    """query self-attention (nn.MultiheadAttention), deformable cross-attention over the
    encoder memory, ffn; post-norm  (…transformer.py:342-374)"""

    def __init__(self, C, ffn, M, L, P):
        super().__init__()
        self.cross_attn = MSDeformAttn(C, L, M, P)
        self.norm1 = nn.LayerNorm(C)
        self.self_attn = nn.MultiheadAttention(C, M, batch_first=True)
        self.norm2 = nn.LayerNorm(C)
        self.linear1 = nn.Linear(C, ffn)
        self.linear2 = nn.Linear(ffn, C)
        self.norm3 = nn.LayerNorm(C)

    def forward(self, tgt, qpos, ref, mem, T, lsi):
        q = tgt + qpos
        tgt = self.norm2(tgt + self.self_attn(q, q, tgt, need_weights=False)[0])
        tgt = self.norm1(tgt + self.cross_attn(tgt + qpos, ref, mem, T, lsi))
        return self.norm3(tgt + self.linear2(F.relu(self.linear1(tgt))))


class _Stack(nn.Module):
    """`encoder.layers.N...` / `decoder.layers.N...`: the reference's state-dict paths."""

    def __init__(self, layers):
        super().__init__()
        self.layers = nn.ModuleList(layers)

    def __iter__(self):
        return iter(self.layers)


class DeformableTransformer(nn.Module):
    """Same parameters, names and shapes as the reference's DeformableTransformer
    (models/deformable/unimodal_deformable_transformer.py:13-60): its state dict loads into this
    one key for key (tests/test_train_step_keys.py, fixture tests/golden/ref_transformer_keys.json)."""

    def __init__(self, C=512, ffn=2048, M=8, L=4, P=4, layers=6):
        super().__init__()
        self.encoder = _Stack(EncoderLayer(C, ffn, M, L, P) for _ in range(layers))
        self.decoder = _Stack(DecoderLayer(C, ffn, M, L, P) for _ in range(layers))
        self.level_embed = nn.Parameter(torch.randn(L, C))
        self.reference_points = nn.Linear(C, 1)


class DeformableEncoderDecoder(nn.Module):
    """The transformer plus the event-query embedding the DVC model owns in the reference."""

    def __init__(self, C=512, ffn=2048, M=8, L=4, P=4, layers=6, queries=100):
        super().__init__()
        self.transformer = DeformableTransformer(C, ffn, M, L, P, layers)
        self.query_embed = nn.Embedding(queries, 2 * C)
        self.L = L

    @property
    def encoder(self):
        return self.transformer.encoder

    @property
    def decoder(self):
        return self.transformer.decoder

    @property
    def level_embed(self):
        return self.transformer.level_embed

    @property
    def reference_points(self):
        return self.transformer.reference_points

    def forward(self, src, pos, enc_ref, T, lsi):
        N = src.shape[0]
        sizes = [int(t) for t in T.tolist()] if not hasattr(self, "_sizes") else self._sizes
        self._sizes = sizes
        lvl = torch.cat([self.level_embed[l].expand(t, -1) for l, t in enumerate(sizes)])
        pos = pos + lvl[None]
        x = src
        for layer in self.encoder:
            x = layer(x, pos, enc_ref, T, lsi)
        qpos, tgt = self.query_embed.weight.chunk(2, -1)
        qpos = qpos[None].expand(N, -1, -1)
        tgt = tgt[None].expand(N, -1, -1)
        ref = self.reference_points(qpos).sigmoid()[:, :, None, :].expand(-1, -1, self.L, -1)
        for layer in self.decoder:
            tgt = layer(tgt, qpos, ref, x, T, lsi)
        return tgt


def op_share(step_fn):
    """Share of one step's GPU time spent in this repo's kernels (names msda_*), from the
    torch profiler's kernel records; None if the profiler is unavailable on the box."""
    try:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            step_fn()
            torch.cuda.synchronize()
        ours = total = 0.0
        launches = 0
        for e in prof.events():
            if e.device_type == torch.autograd.DeviceType.CUDA:
                t = float(e.device_time_total or 0.0)
                total += t
                if "msda_" in e.name:
                    ours += t
                    launches += 1
        return (ours / total if total > 0 else None), ours, total, launches
    except Exception as exc:  # noqa: BLE001
        print("profiler unavailable: %r" % (exc,), file=sys.stderr)
        return None, None, None, None


def bench_record(dev, world, rank, steps=10, warmup=3, batch=16, layers=6):
    """The step under bench.py's launch (python bench.py --train-step, any N; the process group is
    bench.py's): eager, one flat NCCL all-reduce (avg) of all gradients per step.  Returns the
    sub-record of the bench line: samples/s over all ranks (device-timed, max over ranks), the
    all-reduce alone (ms and bus bandwidth 2(N-1)/N * bytes / time) and the op's share."""
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
    w = synthetic.WORKLOADS["ENC"]
    T_list, C = w["T"], w["M"] * w["D"]
    torch.manual_seed(1234)
    model = DeformableEncoderDecoder(C, 2048, w["M"], len(T_list), w["P"], layers, 100).to(dev)
    n_params = sum(p.numel() for p in model.parameters())
    flat = torch.zeros(n_params, device=dev)
    o = 0
    for p_ in model.parameters():
        p_.grad = flat[o:o + p_.numel()].view_as(p_)
        o += p_.numel()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=1e-4, fused=True)
    gen = torch.Generator(device=dev).manual_seed(99 + rank)
    S = sum(T_list)
    src = torch.randn(batch, S, C, device=dev, generator=gen)
    pos = torch.randn(batch, S, C, device=dev, generator=gen) * 0.1
    T = torch.tensor(T_list, dtype=torch.long, device=dev)
    lsi = torch.cat((T.new_zeros(1), T.cumsum(0)[:-1]))
    enc_ref = synthetic.encoder_reference_points(T_list, dev)[None, :, None, None].expand(
        batch, -1, len(T_list), 1).contiguous()

    def step():
        flat.zero_()
        loss = model(src, pos, enc_ref, T, lsi).float().pow(2).mean()
        loss.backward()
        if world > 1:
            dist.all_reduce(flat, op=dist.ReduceOp.AVG)
        torch.nn.utils.clip_grad_norm_(model.parameters(), 0.1, foreach=True)
        opt.step()
        return loss

    step()
    for mod in model.modules():
        if isinstance(mod, MSDeformAttn):
            mod.check_input_lengths = False
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    ar_ms = None
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        a0.record()
        for _ in range(10):
            dist.all_reduce(flat, op=dist.ReduceOp.AVG)
        a1.record()
        torch.cuda.synchronize()
        t = torch.tensor([a0.elapsed_time(a1) / 10], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ar_ms = float(t)
    ms_step = float(ms) / steps
    # the eager loss keeps its autograd graph alive, AccumulateGrad nodes bound to the stream of the eager
    # steps included: they must be gone before a capture on another stream reuses the parameters
    loss_val = float(loss.detach())
    del loss

    # the same step replayed as ONE CUDA graph (forward + backward + all-reduce + clip + AdamW): the eager
    # step is bound by the host's launch rate (8 ranks share the box's cores), the replay is not
    graph_ms = None
    if os.environ.get("TORCH_NCCL_ASYNC_ERROR_HANDLING", "") == "0" or world == 1:
        opt_g = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=1e-4, fused=True, capturable=True)

        def body():
            flat.zero_()
            l_ = model(src, pos, enc_ref, T, lsi).float().pow(2).mean()
            l_.backward()
            if world > 1:
                dist.all_reduce(flat, op=dist.ReduceOp.AVG)
            torch.nn.utils.clip_grad_norm_(model.parameters(), 0.1, foreach=True)
            opt_g.step()
            return l_

        side = torch.cuda.Stream()
        torch.cuda.synchronize()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            body()
        for _ in range(warmup):
            graph.replay()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(steps):
            graph.replay()
        g1.record()
        torch.cuda.synchronize()
        gm = torch.tensor([g0.elapsed_time(g1)], device=dev)
        if world > 1:
            dist.all_reduce(gm, op=dist.ReduceOp.MAX)
        graph_ms = float(gm) / steps
        # the captured graph holds references into the communicator: drop it before the group goes away
        torch.cuda.synchronize()
        graph.reset()
        del graph
        if world > 1:
            dist.barrier()
    # (the profiled step comes last: a capture after a profiler session fails)
    share = op_share(step) if rank == 0 else (step(), None, None, None)[1:]
    rec = {"metric": "configs[4] training step (6+6-layer deformable transformer, 39.4 M parameters + query "
                     "embedding), eager, flat NCCL gradient all-reduce", "samples_per_s": batch * world / ms_step * 1e3,
           "ms_per_step": ms_step, "per_gpu_batch": batch, "n_gpus": world, "steps": steps, "params": n_params,
           "allreduce_bytes": n_params * 4, "allreduce_ms": ar_ms,
           "allreduce_busbw_GBps": (2.0 * (world - 1) / world * n_params * 4 / (ar_ms * 1e-3) / 1e9) if ar_ms else None,
           "collective": "ncclAllReduce (avg) of one flat fp32 buffer over NVLink 5 / NVSwitch" if world > 1 else None,
           "loss": loss_val,
           "cuda_graph_ms_per_step": graph_ms,
           "cuda_graph_samples_per_s": (batch * world / graph_ms * 1e3) if graph_ms else None}
    if rank == 0 and share and share[0] is not None:
        rec.update(op_share_of_gpu_time=share[0], op_us_per_step=share[1], gpu_us_per_step=share[2],
                   op_launches_per_step=share[3])
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=16, help="per-GPU batch (weak scaling)")
    ap.add_argument("--layers", type=int, default=6)
    ap.add_argument("--cuda-graph", action="store_true",
                    help="capture forward + backward + gradient all-reduce + clip + AdamW in one CUDA graph and "
                         "replay it per step (the op enqueues on the capturing stream, allocates nothing itself)")
    ap.add_argument("--torch-frontend", action="store_true",
                    help="MSDeformAttn.fused_frontend = False: the reference's softmax + elementwise torch kernels "
                         "instead of msda_frontend_forward/backward")
    ap.add_argument("--sync", default="flat", choices=["flat", "ddp"],
                    help="gradient exchange at N > 1: 'flat' = every .grad is a view of one 158 MB buffer, one NCCL "
                         "all-reduce (avg) after the backward, no autograd hooks; 'ddp' = DistributedDataParallel "
                         "(bucketed, overlapped with the backward, ~3 ms more host work per step)")
    ap.add_argument("--bucket-mb", type=int, default=25, help="DDP gradient bucket size")
    ap.add_argument("--static-graph", action="store_true", help="DDP static_graph=True")
    ap.add_argument("--bf16-grads", action="store_true", help="all-reduce gradients in bf16 (DDP comm hook)")
    ap.add_argument("--no-tf32", action="store_true",
                    help="full-fp32 GEMMs (default: TF32 matmuls, as in the reference's torch 1.11 image)")
    args = ap.parse_args()
    if args.cuda_graph and args.sync == "ddp":
        raise SystemExit("--cuda-graph needs --sync flat (DDP's reducer waits on the host inside the backward)")

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("tools/train_step.py needs a CUDA device: the op has no CPU path")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if args.cuda_graph:
            os.environ["TORCH_NCCL_ASYNC_ERROR_HANDLING"] = "0"   # its watchdog polls events: illegal in capture
        dist.init_process_group("nccl", device_id=dev)

    torch.backends.cuda.matmul.allow_tf32 = not args.no_tf32
    torch.backends.cudnn.allow_tf32 = not args.no_tf32
    w = synthetic.WORKLOADS["ENC"]
    T_list, C = w["T"], w["M"] * w["D"]
    torch.manual_seed(1234)                                  # same initial weights on every rank
    model = DeformableEncoderDecoder(C, 2048, w["M"], len(T_list), w["P"], args.layers, 100).to(dev)
    n_params = sum(p.numel() for p in model.parameters())
    for mod in model.modules():
        if isinstance(mod, MSDeformAttn):
            mod.fused_frontend = not args.torch_frontend
    net = model
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    flat = None
    if world > 1 and args.sync == "ddp":
        with torch.cuda.stream(side if args.cuda_graph else torch.cuda.current_stream()):
            net = nn.parallel.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True,
                                                      bucket_cap_mb=args.bucket_mb, static_graph=args.static_graph)
        if args.bf16_grads:
            from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
            net.register_comm_hook(None, default_hooks.bf16_compress_hook)
    elif world > 1:
        flat = torch.zeros(n_params, device=dev)
        o = 0
        for p_ in model.parameters():
            p_.grad = flat[o:o + p_.numel()].view_as(p_)
            o += p_.numel()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=1e-4, fused=True,
                            capturable=args.cuda_graph)

    gen = torch.Generator(device=dev).manual_seed(99 + rank)  # each rank its own shard of videos
    S = sum(T_list)
    src = torch.randn(args.batch, S, C, device=dev, generator=gen)
    pos = torch.randn(args.batch, S, C, device=dev, generator=gen) * 0.1
    T = torch.tensor(T_list, dtype=torch.long, device=dev)
    lsi = torch.cat((T.new_zeros(1), T.cumsum(0)[:-1]))
    enc_ref = synthetic.encoder_reference_points(T_list, dev)[None, :, None, None].expand(
        args.batch, -1, len(T_list), 1).contiguous()

    def eager_step():
        if flat is None:
            opt.zero_grad(set_to_none=True)
        else:
            flat.zero_()
        out = net(src, pos, enc_ref, T, lsi)
        loss = out.float().pow(2).mean()
        loss.backward()
        if flat is not None:
            dist.all_reduce(flat, op=dist.ReduceOp.AVG)
        torch.nn.utils.clip_grad_norm_(model.parameters(), 0.1, foreach=True)
        opt.step()
        return loss

    step = eager_step
    eager_step()                                             # checks the level sizes once
    for mod in model.modules():
        if isinstance(mod, MSDeformAttn):
            mod.check_input_lengths = False
    if args.cuda_graph:
        # warm up on a side stream (DDP wants 11 iterations before a capture), then record one step
        torch.cuda.synchronize()
        with torch.cuda.stream(side):
            for _ in range(11):
                eager_step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        if flat is None:
            opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(graph):
            if flat is not None:
                flat.zero_()
            out = net(src, pos, enc_ref, T, lsi)
            graph_loss = out.float().pow(2).mean()
            graph_loss.backward()
            if flat is not None:
                dist.all_reduce(flat, op=dist.ReduceOp.AVG)
            torch.nn.utils.clip_grad_norm_(model.parameters(), 0.1, foreach=True)
            opt.step()

        def step():
            graph.replay()
            return graph_loss
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = float(ms) / args.steps
    if rank == 0:
        share = op_share(step)
    else:
        step()          # the profiled step still all-reduces gradients: every rank takes part
    if rank == 0:
        print(json.dumps({
            "metric": "deformable encoder-decoder training step, samples/s", "value": args.batch * world / ms_step * 1e3,
            "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "scaling": "weak", "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[4]: %d+%d layers, d_model %d, ffn 2048, T=%s, 100 queries, "
                                   "batch %d per GPU, AdamW, clip 0.1, %s GEMMs%s%s" % (
                                       args.layers, args.layers, C, T_list, args.batch,
                                       ("fp32" if args.no_tf32 else "tf32") + (", torch front-end" if args.torch_frontend else ", fused front-end"),
                                       ", whole step replayed as one CUDA graph" if args.cuda_graph else "",
                                       ("" if world == 1 else ", one flat NCCL all-reduce of %.1f MB" % (n_params * 4 / 1e6)
                                        if flat is not None else
                                        ", DDP/NCCL bucket %d MB%s%s" % (args.bucket_mb, ", static graph" if args.static_graph else "",
                                                                         ", bf16 gradient all-reduce" if args.bf16_grads else ""))),
                       "params": n_params},
            "loss": float(loss.detach()),
            "op_share_of_gpu_time": share[0], "op_us_per_step": share[1], "gpu_us_per_step": share[2],
            "op_launches_per_step": share[3],
        }), flush=True)
    if world > 1:
        # a captured graph holds references into the communicator: drop the graph (and everything
        # captured with it) before the process group goes away, then tear down in order
        torch.cuda.synchronize()
        if args.cuda_graph:
            del step
            graph.reset()
            del graph
        dist.barrier()
        torch.cuda.synchronize()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
