"""bench.py — SANA-1.6B 1024px training-step throughput on B200 (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the per-step training hot path over one bucket batch: pad/mask of the text
embeddings, fused noising, SANA-1.6B forward (tcgen05 linear attention inside), fused loss, backward
(gradient checkpointing only if the activations would not fit in HBM), DDP gradient all-reduce (N>1), grad-norm clip,
AdamW.  Workload = BASELINE.json configs[1]: 32x32x32 DC-AE latents, batch 16 per GPU, bf16.

Own arm prints ONE JSON line on rank 0 with
  value     img/s, whole job, inputs already resident in HBM, CUDA-event timed, max over ranks
  e2e       img/s through the plugin call (SanaModel.train_step on a host ``Batch`` in pinned memory:
            H2D of latents + packed embeddings and D2H of the loss inside the timed region)
  roofline  our dominant kernel (by share of the step), timed in situ with CUDA events on the
            launching stream; `kernels` lists every hot-path kernel, in situ and stand-alone
  cpu_baseline  the oracle's full reference step (oracle/train_step.py) on the host cores, bounded
            batch (rank 0, N=1 only)
`--impl reference` times the same oracle step with all host threads (the reference is 100 % Python
on diffusers, which is not installed: the oracle port is the reference CPU path, kind "port").
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "SANA-1.6B 1024px train img/s"
UNIT = "img/s"
WORKLOAD = "SANA-1.6B 1024px bf16 full finetune, 32x32x32 DC-AE latents, batch 16/GPU, text [16,512,2304]"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="sana_1600m_1024", help="debug only: sana_tiny runs the same code small")
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--debug-linattn", default=None, choices=["simt"],
                    help="debug only: run linear attention on the CUDA-core restatement (flagged in config)")
    ap.add_argument("--ckpt", default="auto", choices=["auto", "on", "off"],
                    help="gradient checkpointing: auto = only if the activations do not fit in HBM (default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernels-only", action="store_true",
                    help="debug only: print the stand-alone kernel roofline table and exit (not a bench line)")
    ap.add_argument("--cpu-budget-s", type=float, default=30.0)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU every 200 ms while the timed region runs."""
    REASONS = {0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
               0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
               0x100: "display_clock_setting"}

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=2)
        return False

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# CPU leg: the oracle's full reference step (the only place bench.py executes oracle/)
# ------------------------------------------------------------------------------------------------
def cpu_reference_steps(model_name, images_per_step, steps, warmup, budget_s):
    """Times oracle.train_step.train_step (noise -> SANA forward -> fp32 MSE -> backward -> clip ->
    AdamW) on the host cores.  The per-step batch is a bounded sample of the workload
    (`images_per_step` images of the named shapes); steps are cut down to fit `budget_s` of wall
    clock and the counts actually run are returned."""
    import psutil
    import torch
    from oracle import train_step as ots
    from train_sana import SANA_CONFIGS
    from yat_b200.sana_transformer import PatchedSanaTransformer2DModel   # meta device: shapes only

    with torch.device("meta"):
        meta = PatchedSanaTransformer2DModel(**SANA_CONFIGS[model_name])
    cfg = dict(vars(meta.config))
    n_params = sum(p.numel() for p in meta.parameters())
    need = n_params * 4 * 4 * 1.15 + 6e9          # fp32 weights + grads + 2 Adam moments + activations of one block
    avail = psutil.virtual_memory().available
    if avail < need:
        return {"unavailable": f"host has {avail / 1e9:.0f} GB free, the fp32 oracle step needs ~{need / 1e9:.0f} GB"}
    # all the host threads the run can use: torchrun exports OMP_NUM_THREADS=1 to every rank, which would time the
    # reference arm on ONE core at N > 1 (measured: 4 min for 3 steps instead of 46 s for 4)
    usable = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    want = int(os.environ.get("YAT_CPU_THREADS", "0")) or max(torch.get_num_threads(), min(16, usable))
    if want != torch.get_num_threads():
        torch.set_num_threads(want)
    threads = torch.get_num_threads()
    sd = ots.random_state_dict(meta, dtype=torch.float32)
    opt = torch.optim.AdamW(list(sd.values()), lr=1e-5, weight_decay=0.0)
    s = cfg["sample_size"]
    g = torch.Generator().manual_seed(1234)
    lat = torch.randn((images_per_step, cfg["in_channels"], s, s), generator=g).to(torch.bfloat16)
    embs = [(0.5 * torch.randn((int(torch.randint(20, 301, (1,), generator=g)), cfg["caption_channels"]),
                               generator=g)).to(torch.bfloat16) for _ in range(images_per_step)]
    def one():
        t0 = time.perf_counter()
        ots.train_step(sd, cfg, opt, lat, embs, compute_dtype=torch.float32)
        return time.perf_counter() - t0

    t_begin = time.perf_counter()
    done_warm, last = 0, 0.0
    while done_warm < warmup and (time.perf_counter() - t_begin) + 2 * last < budget_s / 2:
        last = one()
        done_warm += 1
    times = []
    while len(times) < steps:
        if times and (time.perf_counter() - t_begin) + max(times) > budget_s:
            break
        times.append(one())
    sec = statistics.mean(times)
    return {"value": images_per_step / sec, "sec_per_step": sec, "steps": len(times), "warmup": done_warm,
            "cores": threads, "host_cpus": os.cpu_count(), "images_per_step": images_per_step,
            "sample": (f"{images_per_step} image(s)/step of the workload shapes through the oracle's full reference step "
                       f"(noise, SANA fwd with gradient checkpointing, fp32 MSE, bwd, clip, AdamW), fp32 compute on "
                       f"{threads} threads, {len(times)} timed step(s) after {done_warm} warm-up")}


def run_reference_arm(args):
    """`--impl reference`: the reference's CPU path = the oracle port (diffusers/accelerate are not
    installed, so the upstream scripts cannot be imported).  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    r = cpu_reference_steps(args.model, 1, max(1, args.steps), max(0, min(args.warmup, 1)), budget_s=240.0)
    line = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if args.model == "sana_1600m_1024" else f"DEBUG {args.model}",
                       "global_batch": 16 * args.gpus, "parallelism": f"dp{args.gpus}",
                       "sample": "1 image per step on the host cores (bounded sample of the batch-16 workload)"}}
    if "unavailable" in r:
        line["unavailable"] = r["unavailable"]
    else:
        line.update({"value": r["value"], "steps": r["steps"], "warmup": r["warmup"],
                     "ms_per_step": 1e3 * r["sec_per_step"],
                     "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                                      "sample": r["sample"]},
                     "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                     "gpu_launches": 0, "torch": torch.__version__})
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def _device_batch(trainer, batch, device):
    """Inputs of one step made resident in HBM ahead of the timed region: latents on the device and
    the text embeddings already padded + masked (the optimize hook accepts that tuple)."""
    import torch
    from yat_b200.bucket_sampler import Batch
    b = Batch()
    b.ratio = batch.ratio
    b.vae_features = batch.vae_features.to(device, dtype=torch.bfloat16)
    b.embeddings = trainer.pad_text(batch.embeddings)
    b.repa_features = None if batch.repa_features is None else batch.repa_features.to(device)
    return b


def _batch_bytes(batch):
    n = batch.vae_features.numel() * batch.vae_features.element_size()
    n += sum(e.numel() * e.element_size() for e in batch.embeddings)
    n += 4 * (len(batch.embeddings) + 1)     # int32 offsets
    return n


def _timed_steps(fn, steps, dist, device):
    """barrier + synchronize, K steps between two CUDA events on the current stream, synchronize +
    barrier; returns this rank's milliseconds."""
    import torch
    if dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize(device)
    if dist.is_initialized():
        dist.barrier()
    return e0.elapsed_time(e1)


def _max_over_ranks(ms, dist, device):
    import torch
    if not dist.is_initialized():
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# algorithmic HBM bytes per launch: per latent element / token
def _algo_bytes(name, B, chw, N, C, P=0):
    elems = B * chw
    return {
        "grad_clip_coef": 2 * P,                 # every bf16 gradient read once
        "adamw_step": 14 * P,                    # p, g, m, v in + p, m, v out, bf16
        "noise_flowmatch": 4 * elems,            # generate mode: read x0 2 B + write x_t 2 B
        "loss_fused": 6 * elems,                 # read pred 2 + x0 2, write dpred 2 (eps regenerated)
        "linattn_fwd_sm100": 8 * B * N * C,      # q,k,v in + o out, bf16
        "linattn_bwd_sm100": 14 * B * N * C,     # q,k,v,do in + dq,dk,dv out, bf16 (o is recomputed, not read)
        "ln_modulate_fwd": 4 * B * N * C,        # x in, y out
        "ln_modulate_bwd": 6 * B * N * C,        # x, dy in, dx out (one pass: dx and the column sums from the same registers)
        "gate_residual_fwd": 6 * B * N * C,      # res, y in, out
        "gate_residual_bwd": 6 * B * N * C,      # dout, y in, dy out
        "glumb_mid_fwd": 25 * B * N * C,         # u in (10 C) + mid out (5 C) + c out (10 C, kept for the backward)
        "glumb_mid_bwd": 35 * B * N * C,         # u, c (10 C each), dmid (5 C) in; du (10 C) out
    }.get(name)


def _standalone_kernels(device, pk):
    """Every hot-path kernel alone, CUDA events on the current stream, on shapes whose traffic exceeds
    the 126 MB L2 (bulk slab for the elementwise kernels, the K2 layer shape for linear attention)."""
    import torch
    from yat_b200.linear_attention import sana_linear_attention
    from yat_b200.loss import fused_denoise_loss
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    out = {}

    def timeit(fn, iters=20, warm=3):
        """Average device time per call: the launches are queued back to back between two events, so
        host-side allocation in the wrappers does not leak into the figure."""
        for _ in range(warm):
            fn()
        torch.cuda.synchronize(device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(device)
        return e0.elapsed_time(e1) / iters

    def entry(name, nbytes, ms, shape, flops=None):
        e = {"bound": "hbm", "achieved": nbytes / (ms * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
             "us": ms * 1e3, "bytes": nbytes, "shape": shape, "peak_source": pk["source"], "timing": "stand-alone"}
        e["frac"] = e["achieved"] / e["peak"]
        if flops:
            e["tflops"] = flops / (ms * 1e-3) / 1e12
        out[name] = e

    sch = FlowMatchEulerDiscreteScheduler()
    Bb, chw = 4096, 32 * 32 * 32                      # 256 MiB of latents per tensor (> L2)
    x0 = torch.randn((Bb, 32, 32, 32), device=device, dtype=torch.bfloat16)
    ms = timeit(lambda: flowmatch_noise(x0, sch, seed=1, offset=2, want_target=False))
    entry("noise_flowmatch", 4 * Bb * chw, ms, [Bb, 32, 32, 32])
    pred = torch.randn_like(x0).requires_grad_(True)
    ms = timeit(lambda: fused_denoise_loss(pred, x0, None, seed=1, offset=2))
    entry("loss_fused", 6 * Bb * chw, ms, [Bb, 32, 32, 32])
    del x0, pred
    B, N, H = 16, 1024, 70
    C = H * 32
    q, k, v, do = [torch.randn((B, N, C), device=device, dtype=torch.bfloat16) for _ in range(4)]
    from yat_b200.block_ops import gate_residual, ln_modulate
    mod = torch.randn((B, 6, C), device=device, dtype=torch.bfloat16).requires_grad_(True)
    sh, sc, gt = mod.chunk(6, dim=1)[:3]
    ms = timeit(lambda: ln_modulate(q, sc.detach(), sh.detach()))
    entry("ln_modulate_fwd", 4 * B * N * C, ms, [B, N, C])
    ms = timeit(lambda: gate_residual(q, k, gt.detach()))
    entry("gate_residual_fwd", 6 * B * N * C, ms, [B, N, C])
    # (the backward block ops are reported in situ only: stand-alone their ~50 us launches are shorter than the
    #  host cost of torch.autograd.grad, so a back-to-back loop would time the CPU)
    from yat_b200.glumb import glumb_mid
    Hd = int(2.5 * C)
    u = (torch.randn((B, N, 2 * Hd), device=device) * 1.2).to(torch.bfloat16).requires_grad_(True)
    wdw = (torch.randn((2 * Hd, 1, 3, 3), device=device) * 0.3).to(torch.bfloat16).requires_grad_(True)
    bdw = (torch.randn((2 * Hd,), device=device) * 0.2).to(torch.bfloat16).requires_grad_(True)
    dmid = torch.randn((B, N, Hd), device=device).to(torch.bfloat16)
    ms = timeit(lambda: glumb_mid(u, wdw, bdw, 32, 32))
    entry("glumb_mid_fwd", 25 * B * N * C, ms, [B, 32, 32, 2 * Hd])
    mid = glumb_mid(u, wdw, bdw, 32, 32)
    ms = timeit(lambda: torch.autograd.grad(mid, (u, wdw, bdw), dmid, retain_graph=True))
    entry("glumb_mid_bwd", 35 * B * N * C, ms, [B, 32, 32, 2 * Hd])
    del u, wdw, bdw, dmid, mid
    del mod
    ms = timeit(lambda: sana_linear_attention(q, k, v, H))
    entry("linattn_fwd_sm100", 8 * B * N * C, ms, [B, N, C], flops=4 * 32 * 33 * N * H * B)
    qg, kg, vg = [t.clone().requires_grad_(True) for t in (q, k, v)]
    o = sana_linear_attention(qg, kg, vg, H)

    def bwd():
        torch.autograd.grad(o, (qg, kg, vg), do, retain_graph=True)
    try:
        ms = timeit(bwd)
        entry("linattn_bwd_sm100", 14 * B * N * C, ms, [B, N, C], flops=int(2.5 * 4 * 32 * 33 * N * H * B))
    except Exception as e:  # noqa: BLE001 — reported, not hidden
        out["linattn_bwd_sm100"] = {"error": str(e)[:200]}
    return out


def run_own_arm(args):
    import torch
    import torch.distributed as dist
    from yat_b200 import _lib, build as _build
    from train_sana import SanaModel
    from yat_b200.training_parameters_reader import TrainingParameters

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device visible (yat_b200 has no CPU fallback)")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank == 0:
        _build.build()
    pk = peaks()
    params = TrainingParameters()
    params.batch_size = args.batch
    params.model_config = args.model
    params.learning_rate = 1e-5
    params.weight_decay = 0.0
    params.steps = args.steps
    params.gradient_checkpointing = {"auto": "auto", "on": True, "off": False}[args.ckpt]
    cwd = os.getcwd()
    import tempfile
    tmp = tempfile.mkdtemp(prefix="yat_bench_")
    os.chdir(tmp)                      # initialize() creates ./models
    try:
        trainer = SanaModel(params)    # Accelerator inside: init_process_group(nccl) when WORLD_SIZE > 1
        trainer.initialize()
    finally:
        os.chdir(cwd)
    device = trainer.accelerator.device
    _lib.lib()
    if args.debug_linattn:
        for proc in trainer.pipe.transformer.attn_processors.values():
            if hasattr(proc, "impl"):
                proc.impl = args.debug_linattn
    it = iter(trainer.sampler)
    host_batches = [next(it) for _ in range(4)]
    dev_batches = [_device_batch(trainer, b, device) for b in host_batches]
    B = args.batch
    cfg = trainer.pipe.transformer.config
    chw = cfg.in_channels * cfg.sample_size * cfg.sample_size
    N = (cfg.sample_size // cfg.patch_size) ** 2
    C = cfg.num_attention_heads * cfg.attention_head_dim

    n_params = sum(p.numel() for p in trainer.model.parameters() if p.requires_grad)

    losses = []

    def step_resident(i):
        losses.append(trainer.train_step(dev_batches[i % len(dev_batches)]))

    def step_e2e(i):
        # the call a user makes: host Batch in pinned memory -> loss value back on the host
        losses.append(trainer.train_step(host_batches[i % len(host_batches)]).item())

    # ---- warm-up (both paths), then `value`: K steps, inputs resident in HBM
    for i in range(args.warmup):
        step_resident(i)
    sink = []
    _lib.timing_sink = sink
    n0 = _lib.total_launches()
    with ClockSampler(trainer.accelerator.local_process_index) as clk:
        ms_local = _timed_steps(step_resident, args.steps, dist, device)
    launches = _lib.total_launches() - n0
    _lib.timing_sink = None
    ms = _max_over_ranks(ms_local, dist, device)
    value = world * B * args.steps / (ms * 1e-3)

    # ---- in-situ kernel durations from the events recorded inside the timed region
    insitu = {}
    for name, e0, e1 in sink:
        insitu.setdefault(name, []).append(e0.elapsed_time(e1))
    kernels = {}
    step_ms = ms_local / args.steps
    for name, ts in insitu.items():
        nbytes = _algo_bytes(name, B, chw, N, C, n_params)
        avg = statistics.mean(ts)
        k = {"timing": "in-situ", "launches_per_step": len(ts) / args.steps, "us": avg * 1e3,
             "share_of_step": (sum(ts) / args.steps) / step_ms}
        if nbytes:
            k.update({"bound": "hbm", "achieved": nbytes / (avg * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                      "bytes": nbytes})
            k["frac"] = k["achieved"] / k["peak"]
        kernels[name] = k

    # ---- e2e: same K steps through the plugin call with host buffers
    for i in range(max(1, args.warmup // 2)):
        step_e2e(i)
    ms_e2e = _max_over_ranks(_timed_steps(step_e2e, args.steps, dist, device), dist, device)
    e2e_value = world * B * args.steps / (ms_e2e * 1e-3)
    h2d = _batch_bytes(host_batches[0])

    if rank != 0:
        return
    # ---- rank 0 extras: stand-alone kernel roofline, CPU baseline
    standalone = {}
    try:
        del dev_batches
        torch.cuda.empty_cache()
        standalone = _standalone_kernels(device, pk)
    except Exception as e:  # noqa: BLE001
        standalone = {"error": str(e)[:300]}
    dominant = max((n for n in kernels if "frac" in kernels[n]), key=lambda n: kernels[n]["share_of_step"], default=None)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if dominant and os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(dominant, {}).get("dram_bytes_per_launch")
    roofline = None
    if dominant:
        d = kernels[dominant]
        roofline = {"kernel": dominant, "bound": "hbm", "achieved": d["achieved"], "peak": d["peak"], "unit": "GB/s",
                    "frac": d["frac"], "traffic": traffic, "peak_source": pk["source"] + " (MEASURED_PEAKS.json hbm_gbs)",
                    "timing": "in-situ CUDA events over the timed region", "us": d["us"],
                    "standalone": standalone.get(dominant)}
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            r = cpu_reference_steps(args.model, 1, 2, 0, budget_s=args.cpu_budget_s)
            cpu = ({"unavailable": r["unavailable"]} if "unavailable" in r else
                   {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"],
                    "sec_per_step": r["sec_per_step"]})
        except Exception as e:  # noqa: BLE001
            cpu = {"unavailable": str(e)[:300]}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": (WORKLOAD if args.model == "sana_1600m_1024" else f"DEBUG {args.model}")
                   + (f" DEBUG linattn={args.debug_linattn}" if args.debug_linattn else ""),
                   "global_batch": world * B, "parallelism": f"dp{world}", "optimizer": "AdamW bf16, clip + update as 2 own kernels",
                   "gradient_checkpointing": bool(trainer.pipe.transformer.gradient_checkpointing),
                   "l2": "no explicit flush: one step streams ~3.2 GB of weights + >10 GB of activations (>> 126 MB L2)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches,
        "clocks": clk.summary(),
        "roofline": roofline,
        "kernels": {"in_situ": kernels, "standalone": standalone},
        "cpu_baseline": cpu,
        "final_loss": float(losses[-1]) if losses else None,
        "torch": torch.__version__,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.kernels_only:
        import torch
        from yat_b200 import build as _build
        _build.build()
        for name, e in _standalone_kernels(torch.device("cuda", 0), peaks()).items():
            print(json.dumps({"kernel": name, **e}), flush=True)
        return
    if args.impl == "reference":
        run_reference_arm(args)
        return
    run_own_arm(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
    except Exception:
        pass


if __name__ == "__main__":
    main()
