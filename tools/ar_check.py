#!/usr/bin/env python
"""Parity + latency of the peer-memory all-reduce (csrc/allreduce.cu) against NCCL, one process per GPU:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/ar_check.py
Parity: bit-identical on all ranks, and equal to the fp32 sum of the gathered inputs (rank order)
rounded once to bf16; for N = 2 that is also bit-equal to NCCL.  Latency: CUDA-graph replays of 20
back-to-back all-reduces (the decode regime), CUDA events, max over ranks."""
import json
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200")):
    sys.path.insert(0, p)
os.environ.setdefault("B200ATTN_AR_MAX_BYTES", str(4 << 20))       # measure the whole range, not just the default cut-off
import torch
import torch.distributed as dist
from b200attn import ops

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
dev = torch.device("cuda")
ok = True
results = []
# first use builds the communicator (collective)
ops.tp_all_reduce_(torch.zeros(8, dtype=torch.bfloat16, device=dev))
ar = ops._TP_AR
assert ar is not False and ar is not None
sizes = [8, 1024, 4096, 8 * 4096, 64 * 4096, 256 * 4096, 512 * 4096]      # elements; 256 x 4096 = decode bs 256, 8B
for n in sizes:
    for two_shot in ([0, 1, 2, 3] if world > 2 else [0, 3]):
        for rep in range(3):                                               # repeats exercise slot alternation
            g = torch.Generator(device=dev).manual_seed(1000 * rank + n % 997 + rep)
            x = torch.randn(n, generator=g, device=dev, dtype=torch.bfloat16)
            allx = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(allx, x)
            acc = torch.zeros(n, dtype=torch.float32, device=dev)
            for t in allx:
                acc += t.float()
            want = acc.bfloat16()
            if two_shot == 3 and 2 * n * world > 2 * ar.max_bytes:
                continue
            got = ar.all_reduce_(x.clone(), two_shot=two_shot)
            same = [torch.empty_like(got) for _ in range(world)]
            dist.all_gather(same, got)
            if not all(torch.equal(same[0], s) for s in same):
                ok = False; print(f"rank {rank}: n={n} two_shot={two_shot}: ranks disagree", flush=True)
            if not torch.equal(got, want):
                ok = False
                print(f"rank {rank}: n={n} two_shot={two_shot}: mismatch vs fp32 sum, max {(got.float()-want.float()).abs().max().item()}", flush=True)
            if world == 2:
                nc = x.clone(); dist.all_reduce(nc)
                if not torch.equal(nc, got):
                    ok = False; print(f"rank {rank}: n={n}: differs from NCCL", flush=True)


def graph_time(fn, x, iters=20, reps=20):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn(x)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize(); dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn(x)
    g.replay(); torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (reps * iters)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


for n in sizes[1:]:
    x = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    rec = {"elems": n, "bytes": 2 * n, "nccl_us": graph_time(lambda t: dist.all_reduce(t), x),
           "one_shot_us": graph_time(lambda t: ar.all_reduce_(t, two_shot=0), x)}
    if 2 * n * world <= 2 * ar.max_bytes:
        rec["one_shot_push_us"] = graph_time(lambda t: ar.all_reduce_(t, two_shot=3), x)
    if world > 2:
        rec["two_shot_us"] = graph_time(lambda t: ar.all_reduce_(t, two_shot=1), x)
        rec["two_shot_push_us"] = graph_time(lambda t: ar.all_reduce_(t, two_shot=2), x)
    results.append(rec)
    if rank == 0:
        print(json.dumps(rec), flush=True)
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    out = {"world": world, "parity_ok": bool(flag.item()), "gpu": torch.cuda.get_device_name(0), "latency": results}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"ar_check_tp{world}.json"), "w"), indent=1)
    print("PARITY", "OK" if flag.item() else "FAILED", flush=True)
dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)
