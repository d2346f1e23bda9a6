#!/usr/bin/env python
"""Parity + latency of the NVLS launch that fuses the tensor-parallel all-reduce with the residual add + RMSNorm
(csrc/allreduce_nvls.cu) against dist.all_reduce followed by the RMSNorm kernel, one process per GPU:

  B200ATTN_AR_TIMEOUT_S=3 python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29512 tools/nvls_check.py [--out gpurun_out/nvls_tpN.json]

Parity per (tokens, hidden, form 0-3 of include/b200attn.h), in two links: (a) the sum (weight NULL: plain all-reduce)
is identical on all ranks and within the error model of an fp32-accumulating all-reduce (|err| <= 2^-8 |sum| + 2^-21
sum|addends|, checked against the fp64 sum; NCCL's ring, which rounds to bf16 at every hop, is reported against the same
model) -- the in-switch forms 0/1 sum in the switch's own order, the flag-in-data forms 2/3 must EQUAL the fp32 sum in
rank order rounded once; (b) the fused launch equals
add_rms_forward applied to THAT sum: residual bit for bit, normed output identical in >= 99.9 % of the elements and
within 2 ulps elsewhere (the sum of squares is accumulated in another order than in b200_rmsnorm; a last-bit difference
of rstd moves bf16(x * rstd) by an ulp, and bf16(that * w) can land two ulps apart), both identical on all ranks.  (Comparing the residual directly with the NCCL path
in ulps is meaningless where sum and residual cancel: one ulp of a sum of magnitude 2 is 2^-7, whatever is left after
the cancellation.)  Forms are interleaved and replayed from CUDA graphs (slot alternation, flag counters).  Latency: graph replays of 20 back-to-back calls, max over ranks."""
import argparse
import datetime
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200")):
    sys.path.insert(0, p)
os.environ.setdefault("B200ATTN_AR_TIMEOUT_S", "3")
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from b200attn import ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=None)
a = ap.parse_args()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=120))
dev = torch.device("cuda")
nv = ops.tp_nvls()
if not nv:
    if rank == 0:
        print(json.dumps({"nvls": "unavailable"}))
    dist.destroy_process_group()
    sys.exit(0)


def ulps(x, y):
    yf = y.float()
    ulp = torch.pow(2.0, torch.floor(torch.log2(yf.abs().clamp_min(2.0 ** -126))) - 7)
    return float(((x.float() - yf).abs() / ulp).max().item())


def same_everywhere(t):
    h = t.view(torch.int16).to(torch.int64).sum().reshape(1)
    lo, hi = h.clone(), h.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    return bool((lo == hi).item())


FORMS = (0, 1, 2, 3)
FORM_NAMES = {0: "nvls_two_shot_barrier", 1: "nvls_one_shot_barrier", 2: "nvls_ll_one_shot", 3: "nvls_ll_two_shot"}
ok, rows = True, []
for T, H in [(1, 4096), (8, 4096), (64, 4096), (256, 4096), (512, 4096), (256, 5120), (3, 1024)]:
    g = torch.Generator(device=dev).manual_seed(7 * T + H + 1000 * rank)
    gs = torch.Generator(device=dev).manual_seed(T + H)
    x = torch.randn(T, H, generator=g, device=dev).bfloat16()
    res = torch.randn(T, H, generator=gs, device=dev).bfloat16()
    w = (1 + 0.1 * torch.randn(H, generator=gs, device=dev)).bfloat16()
    ref = x.clone()
    dist.all_reduce(ref)
    gathered = [torch.empty_like(x) for _ in range(world)]
    dist.all_gather(gathered, x)
    once = torch.zeros(T, H, device=dev, dtype=torch.float32)
    mag = torch.zeros(T, H, device=dev, dtype=torch.float64)
    exact = torch.zeros(T, H, device=dev, dtype=torch.float64)
    for part in gathered:                                          # fp32 sum in rank order, rounded once
        once += part.float()
        mag += part.double().abs()
        exact += part.double()
    once = once.bfloat16()
    # error model of ANY correct all-reduce that accumulates in fp32 and rounds once: one bf16 rounding of the result
    # plus fp32 accumulation error relative to the addends (cancellation leaves results far smaller than the addends,
    # where "ulps of the result" means nothing; NCCL's ring rounds to bf16 at every hop and is only reported)
    tol = exact.abs() * 2.0 ** -8 + mag * 2.0 ** -21
    tol_switch = exact.abs() * 2.0 ** -7 + mag * 2.0 ** -21       # in-switch bf16 output: within one ulp, not correctly rounded

    def within_model(t, switch=False):
        return bool(((t.double() - exact).abs() <= (tol_switch if switch else tol)).all().item())

    def rounding_stats(t):
        """How the sum was rounded: fraction of elements equal to RN(fp32 sum), and the mean of |t| - |exact| in units of
        the result's bf16 ulp (0 for round-to-nearest, -0.5 for truncation)."""
        ulp = torch.pow(2.0, torch.floor(torch.log2(exact.abs().clamp_min(2.0 ** -100))) - 7)
        big = exact.abs() > 0.25                                   # away from cancellation: the rounding dominates
        bias = ((t.double().abs() - exact.abs()) / ulp)[big].mean().item()
        return dict(fraction_equal_to_rn_of_fp32_sum=float((t == once).float().mean().item()), mean_magnitude_bias_ulps=round(bias, 4))

    nccl_in_model = within_model(ref)
    for form in FORMS + (None,) + FORMS[::-1]:                     # interleaved forms: the flag counters stay in step
        if form is not None and not nv.form_fits(form, T * H * 2):
            continue
        plain = nv.all_reduce_add_rms_norm(x, None, None, 0.0, form=form)
        chosen = nv.choose_form(T * H * 2) if form is None else form
        rank_order = bool(torch.equal(plain, once))
        good = same_everywhere(plain) and within_model(plain, switch=chosen in (0, 1)) and (rank_order or chosen in (0, 1))
        want_y, want_r = ops.rms_norm(plain.clone(), w, 1e-6, res.clone())
        r2 = res.clone()
        y, _ = nv.all_reduce_add_rms_norm(x, r2, w, 1e-6, form=form)
        ident = same_everywhere(y) and same_everywhere(r2)
        res_equal, uy = bool(torch.equal(r2, want_r)), ulps(y, want_y)
        good = good and ident and res_equal and uy <= 2.0 and float((y == want_y).float().mean().item()) >= 0.999
        ok &= good
        rows.append(dict(tokens=T, hidden=H, form=form, chosen=chosen, sum_within_fp32_accumulation_model=within_model(plain, switch=chosen in (0, 1)), sum_rounding=rounding_stats(plain),
                         nccl_within_same_model=nccl_in_model, sum_max_abs_diff_vs_nccl=float((plain.float() - ref.float()).abs().max().item()),
                         sum_equals_fp32_rank_order_sum=rank_order, identical_on_all_ranks=ident, residual_bit_equal=res_equal,
                         normed_max_ulps=uy, normed_identical_fraction=float((y == want_y).float().mean().item()), ok=good))


def graph_us(fn, calls=20, reps=10):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(calls):
            fn()
    gr.replay()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / (reps * calls) * 1e3], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return round(t.item(), 2)


lat = []
for T in (1, 8, 32, 64, 128, 256, 512):
    H = 4096
    x = torch.randn(T, H, device=dev).bfloat16()
    res = torch.zeros(T, H, device=dev).bfloat16()
    w = torch.ones(H, device=dev).bfloat16()
    y = torch.empty_like(x)
    row = dict(tokens=T, bytes=T * H * 2,
               nccl_plus_rmsnorm=graph_us(lambda: (dist.all_reduce(x), ops.rms_norm(x, w, 1e-6, res))),
               peer_memory_plus_rmsnorm=graph_us(lambda: (ops.tp_all_reduce_(x), ops.rms_norm(x, w, 1e-6, res))),
               default_form=nv.choose_form(T * H * 2))
    for form in FORMS:
        if not nv.form_fits(form, T * H * 2):
            continue
        row[FORM_NAMES[form]] = graph_us(lambda: nv.all_reduce_add_rms_norm(x, res, w, 1e-6, form=form, out=y))
    lat.append(row)
status = dict(nvls=nv.status(), peer_memory=ops._TP_AR.status() if ops._TP_AR else None)
ok &= not any(status.values())
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    doc = dict(world=world, parity_ok=bool(flag.item()), status=status, parity=rows, us_per_call_graph_replay=lat)
    print(json.dumps(doc))
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(doc, f, indent=1)
ops.shutdown_collectives()
dist.destroy_process_group()
