#!/usr/bin/env python
"""bench.py — the ScaleProtoSeg prototype-head hot path on B200 (contract in DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--math f16x3|f16|bf16|bf16x3|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A step is one forward+backward of the stage-1 prototype head over one synthetic batch (BASELINE.json
configs[1]: B=16 maps of 256x65x65, P=760 prototypes in 4 scales, K=19 classes, fused mode: no
distance / activation map is written to HBM).  Rank 0 prints ONE JSON line.

  value        positions x prototypes per second, forward+backward, whole job (N independent replicas),
               inputs resident in HBM, CUDA-event timed, max over ranks
  e2e          the same metric through the public API (PrototypeHeadFn autograd) with the batch in pinned HOST
               memory: H2D of z, D2H of logits and of the prototype / last-layer gradients inside the timed region
  roofline     the dominant kernel phase against MEASURED_PEAKS.json
  cpu_baseline the CPU oracle (port of the reference head, torch CPU, all host threads) on a bounded sample
  push         secondary metric: patches/s of the push scan over 3000 Cityscapes-shaped maps sharded over N ranks
`--impl reference` times the CPU port of the reference head on the host cores (rank 0 only).
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

SEED = 20220227
C2 = dict(B=16, H=65, W=65, S=4, D=64, P=760, K=19)       # BASELINE.json configs[1]
PUSH = dict(n_maps=3000, H=129, W=257, P=760, K=19)         # configs[3], the reference's full-res push map
PUSH_COCO = dict(n_maps=512, H=65, W=65, P=6840, K=171)     # configs[4]'s push: COCO-Stuff head on 65x65 maps (bounded sample)
CPU_SAMPLE_B = 16                                           # maps per CPU step: the whole C2 batch (~0.3 s per step)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--math", default="f16x3", choices=["fp32", "bf16", "bf16x3", "f16", "f16x3"])
    ap.add_argument("--no-push", action="store_true", help="skip the secondary push measurement")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-c5", action="store_true", help="skip the secondary COCO-shape forward measurement")
    ap.add_argument("--push-maps", type=int, default=PUSH["n_maps"])
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=float(d["hbm_gbs"]), tensor_burst=float(d["bf16_tflops"]),
                    tensor_sustained=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, tensor_burst=1590.0, tensor_sustained=1400.0, source="fallback")


class ClockSampler:
    """SM clock / throttle reasons sampled while the GPU is under load (B200_PROFILING.md): `nvidia-smi -lms 20` in a
    child process (independent of this process's GIL), started before the warm-up and stopped after the end-to-end
    measurement, so the window covers the timed region and the phase / e2e loops around it."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,utilization.gpu"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index, self.proc = index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        rows = []
        if self.proc is not None:
            try:
                self.proc.terminate()
                out, _ = self.proc.communicate(timeout=10)
                rows = [[p.strip() for p in line.split(",")] for line in out.splitlines() if line.count(",") >= 6]
            except Exception:
                rows = []
        busy = [r for r in rows if r[6].isdigit() and int(r[6]) > 0] or rows     # samples taken under load
        sm = sorted(int(r[0]) for r in busy if r[0].isdigit())
        mx = [int(r[1]) for r in rows if r[1].isdigit()]
        reasons = [n for i, n in enumerate(self.NAMES) if any(r[2 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "source": "nvidia-smi -lms 20, samples with utilization > 0, window = warm-up .. e2e"}


MATH = {"fp32": 0, "bf16x3": 1, "bf16": 2, "f16": 3, "f16x3": 4}
N_SETS = 4   # rotating input sets: 4 x 69 MB of z > 126 MB of L2, so no step re-reads a cached batch


def make_head_inputs(dev, gen_seed):
    """SURVEY 8(d) synthetic inputs: z = sigmoid(randn), prototypes U[0,1), last layer +1 / -0.5."""
    from oracle import head_oracle as ho  # synthetic-input helpers only
    c = C2
    g = torch.Generator(device=dev).manual_seed(gen_seed)
    zs = [torch.sigmoid(torch.randn(c["B"], c["S"] * c["D"], c["H"], c["W"], device=dev, generator=g)) for _ in range(N_SETS)]
    protos = torch.rand(c["P"], c["D"], 1, 1, device=dev, generator=g)
    ident = ho.make_identity(c["P"], c["K"], c["S"])
    last_w = ho.make_last_layer(ident).contiguous().to(dev)
    n = c["B"] * c["H"] * c["W"]
    dls = [torch.randn(n, c["K"], device=dev, generator=g) / n for _ in range(N_SETS)]
    return zs, protos, last_w, dls, ident


def event_ms(fn, iters, stream=None):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def run_ours(args):
    import torch.distributed as dist
    import scaleprotoseg_b200 as sps
    from scaleprotoseg_b200 import _lib, push as spush
    from oracle import head_oracle as ho

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: there is no CPU fallback for the product path"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as ge
        ge.build()
    lib = _lib.load()
    c = C2
    math = MATH[args.math]
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(c["P"], c["S"]), c["K"], math=math)
    # backward: the tensor-core backward when the library has one for this mode, else the fp32 CUDA-core kernels
    cfg_bwd = cfg if getattr(sps, "HAS_TC_BACKWARD", False) else sps.HeadConfig.from_ranges(
        ho.uniform_ranges(c["P"], c["S"]), c["K"], math=sps.MATH_FP32)
    zs, protos, last_w, dls, ident = make_head_inputs(dev, SEED + rank)
    p2 = protos.view(c["P"], c["D"])
    units = c["B"] * c["H"] * c["W"] * c["P"]           # (position, prototype) pairs per step per GPU

    def fwd(i):
        return sps.head_forward_raw(zs[i % N_SETS], p2, last_w, cfg, False, False)[0]

    def bwd(i):
        return sps.head_backward_raw(zs[i % N_SETS], p2, last_w, cfg_bwd, dls[i % N_SETS], None, None, True, True, True)

    def step(i):
        fwd(i)
        bwd(i)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for i in range(max(args.warmup, 3)):
        step(i)
    barrier()
    l0 = int(lib.sps_launch_count())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    barrier()
    launches = int(lib.sps_launch_count()) - l0
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = ms.item() / args.steps
    value = world * units / (ms_per_step * 1e-3)

    # per-phase device time (rank 0 reports): which kernel dominates, and its roofline
    t_fwd = event_ms(fwd, args.steps)
    t_bwd = event_ms(bwd, args.steps)
    return dict(world=world, rank=rank, dev=dev, dist=dist, sps=sps, spush=spush, lib=lib, ho=ho, cfg=cfg, cfg_bwd=cfg_bwd,
                zs=zs, protos=protos, p2=p2, last_w=last_w, dls=dls, ident=ident, units=units, ms_per_step=ms_per_step,
                value=value, launches=launches, t_fwd=t_fwd, t_bwd=t_bwd, sampler=sampler, barrier=barrier)


def measure_nhwc16(ctx, args):
    """Second feature-map source (SpsHeadDesc.z_dtype / z_layout): the same C2 step on 16-bit channels-last maps, fed to
    the tensor cores by TMA (cp.async.bulk.tensor, no conversion work on the SM).  Algorithmic bytes halve (z read, dz
    written in 16 bit), flops stay."""
    sps, c, dev, pk = ctx["sps"], C2, ctx["dev"], peaks()
    n = c["B"] * c["H"] * c["W"]
    out = {"workload": "C2 shape, fp16 / bf16 channels-last feature maps (TMA-fed kernels), fused mode, dz returned in the same layout"}
    for name, dt, math in (("fp16 maps, f16x3", torch.float16, MATH["f16x3"]), ("fp16 maps, f16", torch.float16, MATH["f16"]),
                           ("bf16 maps, bf16", torch.bfloat16, MATH["bf16"])):
        zs = [z.to(dt).contiguous(memory_format=torch.channels_last) for z in ctx["zs"]]
        cfg = sps.HeadConfig.from_ranges(ctx["ho"].uniform_ranges(c["P"], c["S"]), c["K"], math=math)
        fwd = lambda i: sps.head_forward_raw(zs[i % N_SETS], ctx["p2"], ctx["last_w"], cfg, False, False)           # noqa: E731
        bwd = lambda i: sps.head_backward_raw(zs[i % N_SETS], ctx["p2"], ctx["last_w"], cfg, ctx["dls"][i % N_SETS], None, None, True, True, True)  # noqa: E731
        for i in range(3):
            fwd(i), bwd(i)
        it = max(5, min(args.steps, 100))
        tf, tb = event_ms(fwd, it), event_ms(bwd, it)
        bytes_b = n * (2 * c["S"] * c["D"] * 2 + 4 * c["K"])
        flops_b = ctx["units"] * (4 * c["D"] + 4 * c["K"])
        out[name] = {"fwd_ms": tf, "bwd_ms": tb, "step_ms": tf + tb, "value": ctx["units"] / ((tf + tb) * 1e-3),
                     "bwd_hbm_frac": bytes_b / (tb * 1e-3) / 1e9 / pk["hbm"], "bwd_tensor_frac": flops_b / (tb * 1e-3) / 1e12 / pk["tensor_sustained"],
                     "step_tensor_frac": ctx["units"] * (6 * c["D"] + 6 * c["K"]) / ((tf + tb) * 1e-3) / 1e12 / pk["tensor_sustained"]}
    return out


def measure_h2d_probe(ctx, args):
    """Copy-only probe behind the e2e number: the same bytes per step (z + dlogits from pinned host memory, no kernels).
    At N = 8 every GPU of the box hangs off the same host memory / PCIe fabric, which is what flattens the e2e curve."""
    dev, c = ctx["dev"], C2
    zh, dh = ctx["zs"][0].cpu().pin_memory(), ctx["dls"][0].cpu().pin_memory()
    zd, dd = torch.empty_like(ctx["zs"][0]), torch.empty_like(ctx["dls"][0])

    def cp(i):
        zd.copy_(zh, non_blocking=True)
        dd.copy_(dh, non_blocking=True)

    for i in range(3):
        cp(i)
    ctx["barrier"]()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    it = max(5, min(args.steps, 50))
    for i in range(it):
        cp(i)
    e1.record()
    ctx["barrier"]()
    ms = torch.tensor([e0.elapsed_time(e1) / it], device=dev)
    if ctx["world"] > 1:
        ctx["dist"].all_reduce(ms, op=ctx["dist"].ReduceOp.MAX)
    nbytes = zh.numel() * 4 + dh.numel() * 4
    return {"ms_per_step": ms.item(), "h2d_gbs_per_gpu": nbytes / (ms.item() * 1e-3) / 1e9, "h2d_gbs_all_gpus": ctx["world"] * nbytes / (ms.item() * 1e-3) / 1e9,
            "note": "pure pinned-host -> device copies of one step's inputs on every rank at once (max over ranks)"}


def measure_c1(args):
    """BASELINE.json configs[0] (C1): whole-model CPU forward on one 1x3x513x513 image, backbone vs prototype head.  The
    reference's DeepLabv2-ResNet101 is not on the GPU box, so the backbone here is torchvision's deeplabv3_resnet101 body
    (random init, output stride 8: the same ResNet-101 + ASPP cost class) and the head is the oracle port at 65x65;
    profiles/r02_c1_reference_cpu.json holds the same split measured with the UNMODIFIED reference in the build container."""
    from oracle import head_oracle as ho
    import torchvision
    torch.set_num_threads(os.cpu_count() or 1)
    net = torchvision.models.segmentation.deeplabv3_resnet101(weights=None, weights_backbone=None, num_classes=256,
                                                              aux_loss=None).eval()
    x = torch.randn(1, 3, 513, 513)
    c = C2
    protos = torch.rand(c["P"], c["D"], 1, 1)
    w = ho.make_last_layer(ho.make_identity(c["P"], c["K"], c["S"]))
    ranges = ho.uniform_ranges(c["P"], c["S"])
    with torch.no_grad():
        net.classifier(net.backbone(x)["out"])          # warm-up (thread pool, allocator)
        t0 = time.perf_counter()
        f = net.classifier(net.backbone(x)["out"])
        t1 = time.perf_counter()
        z = torch.sigmoid(torch.nn.functional.interpolate(f, size=(c["H"], c["W"]), mode="bilinear", align_corners=False))
        t2 = time.perf_counter()
        ho.head_stage1(z, protos, ranges, w)
        t3 = time.perf_counter()
    return {"workload": "C1: whole-model CPU forward, 1x3x513x513, random init; backbone = torchvision deeplabv3_resnet101 stand-in, head = oracle port (P=760, K=19, 65x65)",
            "cores": os.cpu_count(), "backbone_ms": (t1 - t0) * 1e3, "head_ms": (t3 - t2) * 1e3,
            "head_share": (t3 - t2) / max((t1 - t0) + (t3 - t2), 1e-9), "single_run": True,
            "reference_run": "profiles/r02_c1_reference_cpu.json (unmodified reference, build container)"}


def measure_e2e(ctx, args):
    """fwd+bwd through the public autograd API with HOST buffers: every step uploads its batch (z, dlogits) from pinned
    host memory and downloads logits + parameter gradients.  The upload of step i+1 runs on a copy stream while step i
    computes (the usual input-prefetch pattern); every step's copies are inside the timed region."""
    sps, dev, c = ctx["sps"], ctx["dev"], C2
    n = c["B"] * c["H"] * c["W"]
    NB = 3                                                   # device-side input ring
    z_host = [z.cpu().pin_memory() for z in ctx["zs"][:2]]
    dl_host = [d.cpu().pin_memory() for d in ctx["dls"][:2]]
    z_dev = [torch.empty_like(ctx["zs"][0]) for _ in range(NB)]
    dl_dev = [torch.empty_like(ctx["dls"][0]) for _ in range(NB)]
    logits_host = torch.empty(n, c["K"]).pin_memory()
    dp_host = torch.empty(c["P"], c["D"], 1, 1).pin_memory()
    dw_host = torch.empty(c["K"], c["P"]).pin_memory()
    protos = ctx["protos"].clone().requires_grad_(True)
    last_w = ctx["last_w"].clone().requires_grad_(True)
    main = torch.cuda.current_stream()
    copy = torch.cuda.Stream(device=dev)
    ready = [torch.cuda.Event() for _ in range(NB)]          # upload of slot finished
    freed = [torch.cuda.Event() for _ in range(NB)]          # compute on slot finished

    def upload(i):
        k = i % NB
        with torch.cuda.stream(copy):
            copy.wait_event(freed[k])
            z_dev[k].copy_(z_host[i % 2], non_blocking=True)
            dl_dev[k].copy_(dl_host[i % 2], non_blocking=True)
            ready[k].record(copy)

    def step(i, last):
        k = i % NB
        if not last:
            upload(i + 1)
        main.wait_event(ready[k])
        z = z_dev[k].detach().requires_grad_(True)           # dz is produced too (it goes to the backbone)
        protos.grad = last_w.grad = None
        logits, _, _ = sps.PrototypeHeadFn.apply(z, protos, last_w, ctx["cfg"], False, False)
        logits.backward(dl_dev[k])
        freed[k].record(main)
        logits_host.copy_(logits.detach(), non_blocking=True)
        dp_host.copy_(protos.grad, non_blocking=True)
        dw_host.copy_(last_w.grad, non_blocking=True)

    def run(count):
        upload(0)
        for i in range(count):
            step(i, i == count - 1)

    for k in range(NB):
        freed[k].record(main)
    run(3)
    ctx["barrier"]()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(args.steps)                                          # includes the un-overlapped upload of its first batch
    e1.record()
    ctx["barrier"]()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if ctx["world"] > 1:
        ctx["dist"].all_reduce(ms, op=ctx["dist"].ReduceOp.MAX)
    per = ms.item() / args.steps
    h2d = z_host[0].numel() * 4 + dl_host[0].numel() * 4
    d2h = (logits_host.numel() + dp_host.numel() + dw_host.numel()) * 4
    return {"value": ctx["world"] * ctx["units"] / (per * 1e-3), "unit": "positions*prototypes/s", "ms_per_step": per,
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "note": "public autograd API, pinned host buffers; upload of batch i+1 overlaps compute of batch i"}


def measure_push(ctx, args, shape=None, zdtype=None):
    """Push projection (configs[3]): n_maps Cityscapes-shaped maps sharded over the ranks.  The timed region is the WHOLE
    projection: scan of the shard, dataset arg-min, min all-reduce of the packed keys, gather of the winners on their owner
    ranks and the sum all-reduce that hands the (P, D) block to everybody.  Per-phase device times from CUDA events."""
    sps, spush, ho, dev, world, rank = ctx["sps"], ctx["spush"], ctx["ho"], ctx["dev"], ctx["world"], ctx["rank"]
    p = dict(PUSH if shape is None else shape)
    n_maps = args.push_maps if shape is None else p["n_maps"]
    per_rank = (n_maps + world - 1) // world
    lo = rank * per_rank
    n_local = max(0, min(n_maps, lo + per_rank) - lo)
    pool = min(n_local, 64)            # distinct resident maps (64 x 34 MB = 2.2 GB >> L2), cycled to cover the shard
    g = torch.Generator(device=dev).manual_seed(SEED + 17 * rank)
    z = torch.sigmoid(torch.randn(pool, 256, p["H"], p["W"], device=dev, generator=g))
    if zdtype is not None:
        z = z.to(zdtype).contiguous(memory_format=torch.channels_last)
    labels = torch.stack([ho.make_labels(lo + i, p["H"], p["W"], p["K"], absent_class=3) for i in range(pool)]).to(torch.int16).to(dev)
    math = ctx["cfg"].math if PUSH_TC(sps) else sps.MATH_FP32
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(p["P"], 4), p["K"], math=math)
    ident = ho.make_identity(p["P"], p["K"], 4)
    protos = torch.rand(p["P"], 64, device=dev, generator=g)
    pc = spush.proto_class_of(ident).to(dev)
    ps = spush.proto_scale_reference(p["P"], 4).to(dev)
    keys = torch.empty(n_local, p["P"], device=dev, dtype=torch.int64)
    batch = 64                         # maps per scan launch (the whole resident pool)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]

    def project(timed):
        if timed:
            ev[0].record()
        for i in range(0, n_local, batch):
            b = min(batch, n_local - i)
            j = i % pool
            if j + b > pool:
                j = 0
            spush.scan_maps(z[j:j + b], labels[j:j + b], protos, pc, cfg, keys_out=keys[i:i + b])
        if timed:
            ev[1].record()
        best = spush.reduce_keys(keys, lo)
        if timed:
            ev[2].record()
        if world > 1:
            ctx["dist"].all_reduce(best, op=ctx["dist"].ReduceOp.MIN)
        if timed:
            ev[3].record()
        # winners' vectors from the rank that owns the winning map (the resident pool stands for the shard: map m sits in
        # slot m % pool; which vector is copied is the tests' business, the cost of the gather is what is timed here)
        out = torch.zeros(p["P"], 64, device=dev)
        local = (best & 0xFFFFFFFF) - lo
        mine = (local >= 0) & (local < n_local)
        slot_best = torch.where(mine, (best & ~0xFFFFFFFF) | (local % max(pool, 1)), torch.full_like(best, 0xFFFFFFF))
        if pool > 0:
            spush.gather_winners(z, keys[:pool], slot_best, 0, ps, cfg, out=out)
        if timed:
            ev[4].record()
        if world > 1:
            ctx["dist"].all_reduce(out, op=ctx["dist"].ReduceOp.SUM)
        if timed:
            ev[5].record()
        return best, out

    project(False)
    ctx["barrier"]()
    project(True)
    ctx["barrier"]()
    ph = [ev[i].elapsed_time(ev[i + 1]) for i in range(5)]
    ms = torch.tensor([ev[0].elapsed_time(ev[5])] + ph, device=dev)
    if world > 1:
        ctx["dist"].all_reduce(ms, op=ctx["dist"].ReduceOp.MAX)
    ms = ms.tolist()
    patches = n_maps * p["H"] * p["W"]
    split = cfg.math in (sps.MATH_F16X3, sps.MATH_BF16X3)
    terms = 0 if cfg.math == sps.MATH_FP32 else (1 if not split else (3 if zdtype is None else 2))
    bytes_per_patch = 256 * (2 if zdtype is not None else 4) + 2
    return {"value": patches / (ms[0] * 1e-3), "unit": "patches/s", "ms": ms[0], "scaling": "strong",
            "phases_ms": {"scan": ms[1], "dataset_argmin": ms[2], "allreduce_min_keys": ms[3], "gather_winners": ms[4], "allreduce_sum_winners": ms[5]},
            "scan_hbm_frac_per_gpu": (patches / world) * bytes_per_patch / (ms[1] * 1e-3) / 1e9 / peaks()["hbm"],
            # distance GEMM as issued: 2*D*P flops per patch and product term (3 terms for fp32 maps in a 3-term mode, 2 for
            # 16-bit maps, 1 in the single-term modes) against the sustained bf16 tensor peak
            "scan_tensor_frac_per_gpu": (patches / world) * 2 * 64 * p["P"] * terms / (ms[1] * 1e-3) / 1e12 / peaks()["tensor_sustained"],
            "product_terms": terms,
            "workload": f"{n_maps} maps 256x{p['H']}x{p['W']} {'fp32 NCHW' if zdtype is None else str(zdtype).replace('torch.', '') + ' NHWC'}, "
                        f"P={p['P']}, K={p['K']}, sharded {world}-way, {pool} resident maps per rank cycled; whole projection timed",
            "math": "fp32" if cfg.math == 0 else args.math}


def measure_dense(ctx, args):
    """Secondary: the step as the UNCHANGED reference trainer drives it (API-exact 'dense' mode): forward returns logits
    AND the dense (B,P,H,W) distance map, backward receives dlogits AND a gradient through the distance map (KLD loss)."""
    sps, c, dev = ctx["sps"], C2, ctx["dev"]
    n = c["B"] * c["H"] * c["W"]
    g = torch.Generator(device=dev).manual_seed(SEED + 9)
    dd = torch.randn(c["B"], c["P"], c["H"], c["W"], device=dev, generator=g) * (1e-3 / n)
    zs, p2, w, dls, cfg = ctx["zs"], ctx["p2"], ctx["last_w"], ctx["dls"], ctx["cfg"]

    def step(i):
        sps.head_forward_raw(zs[i % N_SETS], p2, w, cfg, True, False)
        sps.head_backward_raw(zs[i % N_SETS], p2, w, cfg, dls[i % N_SETS], dd, None, True, True, True)

    iters = max(3, min(args.steps, 50))
    for i in range(3):
        step(i)
    ms = event_ms(step, iters)
    bytes_alg = n * (3 * c["S"] * c["D"] * 4 + 8 * c["K"]) + 2 * n * c["P"] * 4      # + distance map out, ddist in
    return {"workload": "same shape, dense mode: distance map written by the forward (205 MB) and its gradient read by the backward",
            "ms_per_step": ms, "value": ctx["units"] / (ms * 1e-3), "unit": "positions*prototypes/s",
            "hbm_frac": bytes_alg / (ms * 1e-3) / 1e9 / peaks()["hbm"]}


def measure_stage2(ctx, args):
    """Secondary: BASELINE.json configs[2], the stage-2 grouping finetune step on the same batch AS THE TRAINER CALLS IT
    (module_multiscale_group_train.py:224,242,260): the head forward returns BOTH dense maps (distances and activations),
    the group projection + exp + last_layer_group produce logits and the group activations, and the backward receives
    dlogits (cross-entropy) AND a gradient on the group activations (KLDLossGroup / entropy losses); prototypes and
    features frozen, as finetune_wandb_group trains."""
    sps, ho, dev, c = ctx["sps"], ctx["ho"], ctx["dev"], C2
    from scaleprotoseg_b200 import head as sh
    tables = sh.build_group_tables(ctx["ident"], 3, dev)
    g = torch.Generator(device=dev).manual_seed(SEED + 3)
    Gp = torch.rand(3 * c["P"], device=dev, generator=g).requires_grad_(True)
    Wg = torch.randn(c["K"], 3 * tables.Kc, device=dev, generator=g).requires_grad_(True)
    zs, p2, dls, cfg = ctx["zs"], ctx["p2"], ctx["dls"], ctx["cfg"]
    n = c["B"] * c["H"] * c["W"]
    dgr = torch.randn(n, 3 * tables.Kc, device=dev, generator=g) / n

    def step(i):
        _, dist, act = sps.head_forward_raw(zs[i % N_SETS], p2, None, cfg, True, True)
        groups, logits = sh.GroupHeadFn.apply(act, Gp, Wg, cfg, tables, (c["B"], c["H"], c["W"]))
        Gp.grad = Wg.grad = None
        torch.autograd.backward([logits, groups], [dls[i % N_SETS], dgr])

    iters = max(3, min(args.steps, 50))
    for i in range(3):
        step(i)
    ms = event_ms(step, iters)
    gt = 3 * tables.Kc
    # z read; distance + activation maps written (API outputs); activations read twice (projection, dG); groups / logits
    bytes_alg = n * c["S"] * c["D"] * 4 + 4 * n * c["P"] * 4 + 4 * n * gt * 4
    return {"workload": "stage-2 grouping head as the trainer calls it: dense distance + activation maps out, group projection (3 per class) + "
                        "last_layer_group, backward from dlogits and dgroups to the projections and the group last layer, same batch",
            "ms_per_step": ms, "value": ctx["units"] / (ms * 1e-3), "unit": "positions*prototypes/s",
            "hbm_frac": bytes_alg / (ms * 1e-3) / 1e9 / peaks()["hbm"],
            "traffic_model": "z + 2 dense maps written + activation map read twice = 4.3 passes of 205 MB (was 5.3: the backward no longer re-reads it for the groups)"}


def measure_kld(ctx, args):
    """Secondary: the KLD loss between same-class prototypes on the dense distance map of the same
    batch — sps_kld_fwd + sps_kld_bwd through the autograd Function, next to the CPU oracle's restatement of the
    reference loop (segmentation/model/loss.py:66-146) on one image."""
    sps, ho, dev, c = ctx["sps"], ctx["ho"], ctx["dev"], C2
    from scaleprotoseg_b200.kld import KldTables, kld_loss
    ranges = ho.uniform_ranges(c["P"], c["S"])
    tables = KldTables(ctx["ident"], ranges, dev)
    labels = torch.stack([ho.make_labels(i, c["H"], c["W"], c["K"]) for i in range(c["B"])])
    lab_dev = labels.to(dev)
    dist = sps.head_forward_raw(ctx["zs"][0], ctx["p2"], None, ctx["cfg"], True, False)[1].requires_grad_(True)

    def fwd(i):
        return kld_loss(dist, lab_dev, tables)

    def step(i):
        dist.grad = None
        kld_loss(dist, lab_dev, tables).backward()

    iters = max(3, min(args.steps, 50))
    for i in range(3):
        step(i)
    ms_fwd, ms = event_ms(fwd, iters), event_ms(step, iters)
    # feature mode: the same loss from z and the prototypes, no (B,P,H,W) map; and the whole stage-1 style step on it
    # (fused head fwd + bwd on dlogits, plus this loss's own dz / dprototypes)
    z0 = ctx["zs"][0].detach().requires_grad_(True)
    pv = ctx["protos"].detach().requires_grad_(True)
    cfg32 = sps.HeadConfig.from_ranges(ranges, c["K"], math=sps.MATH_FP32)

    def feat_step(i):
        z0.grad = pv.grad = None
        sps.kld_loss_from_features(z0, pv, lab_dev, tables, cfg32).backward()

    def train_step(i):
        z0.grad = pv.grad = None
        logits = sps.PrototypeHeadFn.apply(z0, pv, ctx["last_w"], ctx["cfg"], False, False)[0]
        loss = sps.kld_loss_from_features(z0, pv, lab_dev, tables, cfg32)
        torch.autograd.backward([logits, loss], [ctx["dls"][0].reshape(logits.shape), torch.ones_like(loss)])

    def train_step_dense(i):
        z0.grad = pv.grad = None
        logits, dmap, _ = sps.PrototypeHeadFn.apply(z0, pv, ctx["last_w"], ctx["cfg"], True, False)
        loss = kld_loss(dmap, lab_dev, tables)
        torch.autograd.backward([logits, loss], [ctx["dls"][0].reshape(logits.shape), torch.ones_like(loss)])

    for i in range(3):
        feat_step(i), train_step(i), train_step_dense(i)
    ms_feat, ms_train, ms_train_dense = event_ms(feat_step, iters), event_ms(train_step, iters), event_ms(train_step_dense, iters)
    n = c["B"] * c["H"] * c["W"]
    per_cs = c["P"] // c["K"] // c["S"]
    touched = n * c["S"] * per_cs                      # (pixel, prototype) pairs whose classes agree
    out = {"workload": "KLD loss between same-class prototypes on the dense distance map of the same batch "
                       "(B=%d, P=%d: %d pairs per image, %d class-matched (pixel, prototype) pairs)" % (c["B"], c["P"], tables.Q, touched),
           "ms_fwd": ms_fwd, "ms_fwd_bwd": ms, "loss": float(fwd(0).detach()),
           "note": "backward includes the zero fill of the 205 MB gradient map the head backward then reads",
           "feature_mode_ms_fwd_bwd": ms_feat,
           "train_step_ms": {"fused head + feature-mode KLD (no distance map)": ms_train,
                             "dense head + KLD on the map + head backward through ddist": ms_train_dense},
           "train_step_note": "autograd step incl. Python dispatch: head forward, loss, head backward, gradients accumulated"}
    if not args.no_cpu:
        import time
        d1 = dist.detach()[:1].cpu()
        t0 = time.perf_counter()
        d1r = d1.clone().requires_grad_(True)
        ho.kld_loss(d1r, labels[:1], ctx["ident"].cpu(), ranges).backward()
        cpu_ms = (time.perf_counter() - t0) * 1e3
        out["cpu_port_ms_per_image_fwd_bwd"] = cpu_ms
        out["cpu_port_note"] = "oracle port (one matrix per image x class x scale, already far fewer ops than the reference's per-pair loop), 1 image, all host threads"
    return out


def measure_eval_topk(ctx, args):
    """Secondaries for two more SURVEY 8f rows.  (i) Eval epilogue of one Cityscapes validation image
    (eval_valid_multiscale.py:229-234): distances (760,129,257) and logits (19,129,257) upsampled to 1024x2048 and
    reduced to nearest-prototype / predicted-class maps; next to it the reference's own two lines (F.interpolate +
    argmin / argmax) in stock PyTorch on the same GPU.  (ii) k = 5 nearest maps per prototype over 3 000 maps' keys."""
    import torch.nn.functional as F
    sps, spush, dev = ctx["sps"], ctx["spush"], ctx["dev"]
    g = torch.Generator(device=dev).manual_seed(SEED + 21)
    P, K, h, w, H, W = C2["P"], C2["K"], 129, 257, 1024, 2048
    dist = torch.rand(1, P, h, w, device=dev, generator=g) * 8.0
    logits = torch.randn(1, K, h, w, device=dev, generator=g)

    def ours(i):
        sps.upsample_argmin(dist, (H, W))
        sps.upsample_argmax(logits, (H, W))

    def stock(i):
        torch.argmin(F.interpolate(dist, size=(H, W), mode="bilinear", align_corners=False)[0], dim=0)
        torch.argmax(F.interpolate(logits, size=(H, W), mode="bilinear", align_corners=False)[0], dim=0)

    for i in range(2):
        ours(i), stock(i)
    ms_ours, ms_stock = event_ms(ours, 10), event_ms(stock, 5)
    n_maps, k = PUSH["n_maps"], 5
    keys = torch.randint(0, 2 ** 62, (n_maps, P), device=dev, dtype=torch.int64, generator=g)
    spush.topk_keys(keys, k)
    ms_topk = event_ms(lambda i: spush.topk_keys(keys, k), 10)
    return {"eval_epilogue": {"workload": "one image: distances (760,129,257) + logits (19,129,257) -> argmin / argmax at 1024x2048",
                              "ms": ms_ours, "torch_gpu_ms": ms_stock,
                              "note": "stock PyTorch materialises the 6.4 GB upsampled distance map; ours writes two index maps (2 x 8 MB)",
                              "out_pixels_x_channels_per_s": (H * W * (P + K)) / (ms_ours * 1e-3)},
            "topk": {"workload": "k=5 nearest maps per prototype from %d x %d packed keys (sps_push_topk)" % (n_maps, P), "ms": ms_topk}}


def measure_c5(ctx, args):
    """Secondary: the COCO-Stuff head shape of BASELINE.json configs[4] (K=171, P=6840, B=32, 65x65) per GPU, fused
    forward + backward in the package's default math (3-term distances; the wide last layer runs the logits GEMM on one
    16-bit term and the one-pass wide backward, see head_bwd_wide.cu).  At N > 1 every rank runs a replica."""
    sps, ho, dev, world = ctx["sps"], ctx["ho"], ctx["dev"], ctx["world"]
    B, H, W, P, K, S, D = 32, 65, 65, 6840, 171, 4, 64
    g = torch.Generator(device=dev).manual_seed(SEED + 5 + ctx["rank"])
    z = torch.sigmoid(torch.randn(B, S * D, H, W, device=dev, generator=g))
    protos = torch.rand(P, D, device=dev, generator=g)
    last_w = ho.make_last_layer(ho.make_identity(P, K, S)).contiguous().to(dev)
    dl = torch.randn(B * H * W, K, device=dev, generator=g) / (B * H * W)
    cfg = sps.HeadConfig.from_ranges(ho.uniform_ranges(P, S), K, math=MATH[args.math])
    r0 = int(ctx["lib"].sps_fp32_reroute_count())
    fwd = lambda i: sps.head_forward_raw(z, protos, last_w, cfg, False, False)                       # noqa: E731  (138 MB of z per pass > L2)
    bwd = lambda i: sps.head_backward_raw(z, protos, last_w, cfg, dl, None, None, True, True, True)  # noqa: E731
    fwd(0), bwd(0)
    iters = max(3, min(args.steps, 10))
    ctx["barrier"]()
    t = torch.tensor([event_ms(fwd, iters), event_ms(bwd, iters)], device=dev)
    if world > 1:
        ctx["dist"].all_reduce(t, op=ctx["dist"].ReduceOp.MAX)
    ms, ms_b = t.tolist()
    units = B * H * W * P
    pk = peaks()
    flops_f, flops_b = units * (2 * D + 2 * K), units * (4 * D + 4 * K)
    z16 = z.half().contiguous(memory_format=torch.channels_last)
    fwd16 = lambda i: sps.head_forward_raw(z16, protos, last_w, cfg, False, False)                       # noqa: E731
    bwd16 = lambda i: sps.head_backward_raw(z16, protos, last_w, cfg, dl, None, None, True, True, True)  # noqa: E731
    fwd16(0), bwd16(0)
    t16 = [event_ms(fwd16, iters), event_ms(bwd16, iters)]
    return {"workload": f"COCO-Stuff head (configs[4] shape) per GPU: B={B} maps 256x{H}x{W} fp32 NCHW, P={P}, K={K}, fused mode, {world} replica(s)",
            "math": args.math, "forward_ms": ms, "backward_ms": ms_b, "value": world * units / ((ms + ms_b) * 1e-3),
            "unit": "positions*prototypes/s (fwd+bwd, all replicas)",
            "forward_tensor_frac": flops_f / (ms * 1e-3) / 1e12 / pk["tensor_sustained"],
            "fwd_bwd_tflops_algorithmic_per_gpu": (flops_f + flops_b) / ((ms + ms_b) * 1e-3) / 1e12,
            "fwd_bwd_tensor_frac": (flops_f + flops_b) / ((ms + ms_b) * 1e-3) / 1e12 / pk["tensor_sustained"],
            "fp16_nhwc_maps": {"forward_ms": t16[0], "backward_ms": t16[1],
                               "fwd_bwd_tensor_frac": (flops_f + flops_b) / ((t16[0] + t16[1]) * 1e-3) / 1e12 / pk["tensor_sustained"]},
            "fp32_reroutes": int(ctx["lib"].sps_fp32_reroute_count()) - r0,
            "backward_note": "one pass: tile-stationary dz kernel + chunk-stationary dp/dW kernel (no class slabs, no dz atomics); "
                             "the fp32 map is re-laid once per call as a 16-bit channels-last array for the TMA-fed kernels"}


def PUSH_TC(sps):
    return bool(getattr(sps, "HAS_TC_PUSH", False))


def cpu_head_step(ho, z, protos, last_w, ranges, dl):
    zr, pr, wr = z.clone().requires_grad_(True), protos.clone().requires_grad_(True), last_w.clone().requires_grad_(True)
    logits, _, _ = ho.head_stage1(zr, pr, ranges, wr)
    logits.backward(dl)
    return logits


def measure_cpu(steps, warmup):
    """The CPU port of the reference head (oracle/head_oracle.py: same op sequence as model_multiscale.py:256-388),
    fwd+bwd on CPU_SAMPLE_B maps per step, all host threads."""
    from oracle import head_oracle as ho
    c = C2
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    g = torch.Generator().manual_seed(SEED)
    z = torch.sigmoid(torch.randn(CPU_SAMPLE_B, c["S"] * c["D"], c["H"], c["W"], generator=g))
    protos = torch.rand(c["P"], c["D"], 1, 1, generator=g)
    last_w = ho.make_last_layer(ho.make_identity(c["P"], c["K"], c["S"])).contiguous()
    ranges = ho.uniform_ranges(c["P"], c["S"])
    n = CPU_SAMPLE_B * c["H"] * c["W"]
    dl = (torch.randn(n, c["K"], generator=g) / n).view(CPU_SAMPLE_B, c["H"], c["W"], c["K"])
    for _ in range(warmup):
        cpu_head_step(ho, z, protos, last_w, ranges, dl)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_head_step(ho, z, protos, last_w, ranges, dl)
    dt = (time.perf_counter() - t0) / steps
    units = n * c["P"]
    return {"value": units / dt, "unit": "positions*prototypes/s", "cores": cores, "kind": "port",
            "sample": f"{CPU_SAMPLE_B} map(s) of 256x{c['H']}x{c['W']} per step, P={c['P']}, K={c['K']}, fp32 torch CPU, "
                      f"{steps} steps after {warmup} warm-up", "ms_per_step": dt * 1e3}


def measure_torch_gpu(ctx, args):
    """The same op sequence as the reference head (oracle port: model_multiscale.py:256-388 restated in stock PyTorch
    ops, fp32, autograd backward) run on this GPU: the GPU baseline of SURVEY 8(d).  A baseline leg like cpu_baseline:
    the oracle is the thing compared against, never the thing shipped."""
    from oracle import head_oracle as ho
    dev, c = ctx["dev"], C2
    g = torch.Generator(device=dev).manual_seed(SEED + 5)
    z = torch.sigmoid(torch.randn(c["B"], c["S"] * c["D"], c["H"], c["W"], device=dev, generator=g))
    protos = torch.rand(c["P"], c["D"], 1, 1, device=dev, generator=g)
    last_w = ho.make_last_layer(ho.make_identity(c["P"], c["K"], c["S"])).contiguous().to(dev)
    ranges = ho.uniform_ranges(c["P"], c["S"])
    n = c["B"] * c["H"] * c["W"]
    dl = (torch.randn(n, c["K"], device=dev, generator=g) / n).view(c["B"], c["H"], c["W"], c["K"])
    for _ in range(3):
        cpu_head_step(ho, z, protos, last_w, ranges, dl)
    ms = event_ms(lambda i: cpu_head_step(ho, z, protos, last_w, ranges, dl), 10)
    return {"ms_per_step": ms, "value": n * c["P"] / (ms * 1e-3), "unit": "positions*prototypes/s",
            "note": "oracle port of the reference head in stock PyTorch ops (2 convs + elementwise chain + linear, autograd "
                    "backward), fp32, on the same GPU; 10 steps after warm-up"}


def config_dict(args, world):
    c = C2
    return {"workload": f"ScaleProtoSeg Cityscapes head fwd+bwd (BASELINE.json configs[1]): B={c['B']} maps 256x{c['H']}x{c['W']} "
                        f"fp32 NCHW, P={c['P']} (4 scales x 190), K={c['K']}, D={c['D']}; fused mode (logits + dz, dprototypes, dW; "
                        f"no distance map in HBM)",
            "replicas": world, "l2": f"{N_SETS} rotating input sets ({N_SETS} x 69 MB z > 126 MB L2)", "math": args.math}


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner at
    communicator creation), so fd 1 is pointed at stderr for the whole run and the JSON line goes to the saved fd."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(saved, "w")


def main():
    args = parse_args()
    json_out = _claim_stdout()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    base = {"metric": "feature-positions x prototypes per second, prototype head fwd+bwd", "unit": "positions*prototypes/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "data": "synthetic", "config": config_dict(args, world)}
    if args.impl == "reference":
        if rank != 0:
            return
        steps = max(1, min(args.steps, 5))
        cpu = measure_cpu(steps, 1)
        out = dict(base)
        out.update({"impl": "reference", "value": cpu["value"], "ms_per_step": cpu["ms_per_step"], "dtype": "f32",
                    "steps": steps, "cpu_baseline": cpu, "gpu_launches": 0,
                    "e2e": {"value": cpu["value"], "unit": cpu["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        print(json.dumps(out), file=json_out, flush=True)
        return

    ctx = run_ours(args)
    e2e = measure_e2e(ctx, args)
    ctx["clocks"] = ctx["sampler"].stop() if rank == 0 else None
    h2d = measure_h2d_probe(ctx, args)
    push = None if args.no_push else measure_push(ctx, args)
    push16 = None if args.no_push else measure_push(ctx, args, zdtype=torch.float16)
    push_coco = None if (args.no_push or args.no_c5) else measure_push(ctx, args, shape=PUSH_COCO)
    c5 = None if args.no_c5 else measure_c5(ctx, args)
    single = not (args.no_c5 or ctx["world"] > 1)
    nhwc16 = measure_nhwc16(ctx, args) if single else None
    dense = measure_dense(ctx, args) if single else None
    stage2 = measure_stage2(ctx, args) if single else None
    torch_gpu = measure_torch_gpu(ctx, args) if single else None
    kld = measure_kld(ctx, args) if single else None
    extra = measure_eval_topk(ctx, args) if single else None
    if ctx["world"] > 1:
        ctx["barrier"]()
    if rank != 0:
        if ctx["world"] > 1:
            ctx["dist"].destroy_process_group()
        return
    c, pk = C2, peaks()
    n = c["B"] * c["H"] * c["W"]
    # dominant phase and its roofline (algorithmic counts: SURVEY 8(d), restated in DESIGN.md)
    phases = {
        "forward (head_fwd kernels)": dict(ms=ctx["t_fwd"], flops=ctx["units"] * (2 * c["D"] + 2 * c["K"]),
                                             bytes=n * (c["S"] * c["D"] * 4 + 4 * c["K"])),
        "backward (head_bwd kernels)": dict(ms=ctx["t_bwd"], flops=ctx["units"] * (4 * c["D"] + 4 * c["K"]),
                                            bytes=n * (2 * c["S"] * c["D"] * 4 + 4 * c["K"])),
    }
    name = max(phases, key=lambda k: phases[k]["ms"])
    ph = phases[name]
    t = ph["ms"] * 1e-3
    tensor_frac = ph["flops"] / t / 1e12 / pk["tensor_sustained"]
    hbm_frac = ph["bytes"] / t / 1e9 / pk["hbm"]
    if tensor_frac >= hbm_frac:
        roof = {"bound": "tensor", "achieved": ph["flops"] / t / 1e12, "peak": pk["tensor_sustained"], "unit": "TFLOP/s", "frac": tensor_frac}
    else:
        roof = {"bound": "hbm", "achieved": ph["bytes"] / t / 1e9, "peak": pk["hbm"], "unit": "GB/s", "frac": hbm_frac}
    traffic = None   # DRAM bytes per launch of the dominant kernel, from the committed ncu --set full capture
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r02_traffic.json")) as f:
            traffic = json.load(f).get("backward" if "backward" in name else "forward")
    except (OSError, ValueError):
        pass
    roof.update({"kernel": name, "ms_per_launch": ph["ms"], "peak_source": pk["source"] + " (sustained bf16 / copy)", "traffic": traffic,
                 "phases_ms": {k: v["ms"] for k, v in phases.items()}, "tensor_frac": tensor_frac, "hbm_frac": hbm_frac})
    out = dict(base)
    out.update({"value": ctx["value"], "ms_per_step": ctx["ms_per_step"], "dtype": args.math + " operands, f32 accumulate"})
    if push is not None:     # early in the line: the driver keeps the head of a long stdout line
        out["push"] = push
    out.update({"clocks": ctx["clocks"], "e2e": e2e, "gpu_launches": ctx["launches"], "roofline": roof,
                "fp32_reroutes": int(ctx["lib"].sps_fp32_reroute_count()), "h2d_probe": h2d})
    for k, v in (("push_fp16_nhwc", push16), ("push_coco", push_coco), ("coco", c5), ("nhwc16", nhwc16), ("dense_mode", dense),
                 ("stage2", stage2), ("kld", kld)):
        if v is not None:
            out[k] = v
    if extra is not None:
        out.update(extra)
    if torch_gpu is not None:
        out["torch_gpu_baseline"] = torch_gpu
    if not args.no_cpu and ctx["world"] == 1:
        out["cpu_baseline"] = measure_cpu(5, 1)
        out["c1"] = measure_c1(args)
    if ctx["world"] > 1:
        ctx["dist"].destroy_process_group()
    print(json.dumps(out), file=json_out, flush=True)


if __name__ == "__main__":
    main()
