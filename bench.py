"""bench.py — SANA-1.6B 1024px training-step throughput on B200 (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

`--workload k3|k4|k5` runs another BASELINE config alone (same JSON schema); the default k2 line also carries a short
run of each under `other_workloads`, and `account` = where the step's time goes (profiler pass: own / cuBLAS / cuDNN /
ATen / NCCL kernel time per step, GPU-busy fraction).

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
# the product pads the text batch to its longest caption rounded up to 64 tokens (<= 320 with the synthetic
# U{20..300} lengths) instead of always 512: masked keys weigh exactly 0, so the result is the reference's, but the
# K/V projection and cross-attention work measured is that of the trimmed batch (config.text_padding says so)
TEXT_PADDING = "trim64 (reference: fixed 512; masked keys weigh 0, tests/test_step_gpu.py::test_trimmed_text_padding_matches_fixed_512)"

# BASELINE.json configs[1..4].  k2 is the metric's configuration (the bench line); k3 / k4 / k5 are the other
# configs, timed briefly after the main run (`other_workloads`) or alone with --workload.
WORKLOADS = {
    "k2": {"workload": WORKLOAD, "batch": 16},
    "k3": {"workload": "PixArt-Sigma XL 600M 1024px, 4-ch 128x128 latents, aspect-ratio buckets, bf16 (full finetune, batch 16/GPU, "
                       "square bucket, text [16,300,4096])", "batch": 16},
    "k4": {"workload": "SD3.5-medium MMDiT with logit-normal timestep weighting, LoRA rank 32, batch 8/GPU "
                       "(16-ch 128x128 latents, text [8,333,4096])", "batch": 8},
    "k5": {"workload": "SANA-1.6B with REPA alignment loss (DINOv2-shaped synthetic features) + use_preservation, LoKr rank 16 "
                       "(batch 16/GPU, 32x32x32 latents)", "batch": 16},
}


def make_trainer(workload, batch, ckpt="auto", model=None):
    """(trainer class, TrainingParameters) of a BASELINE config; scripts and tests build the same objects."""
    from yat_b200.training_parameters_reader import TrainingParameters
    p = TrainingParameters()
    p.batch_size = batch
    p.learning_rate = 1e-5
    p.weight_decay = 0.0
    p.gradient_checkpointing = {"auto": "auto", "on": True, "off": False}[ckpt]
    if workload == "k2":
        from train_sana import SanaModel as T
        p.model_config = model or "sana_1600m_1024"
    elif workload == "k3":
        from train_pixart_sigma import PixartSigmaTrainer as T
        p.model_config = model or "pixart_sigma_xl_1024"
    elif workload == "k4":
        from train_sd35 import SD35Trainer as T
        p.model_config = model or "sd35_medium"
        p.lora_rank, p.lora_alpha, p.lora_algo = 32, 32, "lora"
        p.lora_target_modules = "to_q to_k to_v to_out.0 add_q_proj add_k_proj add_v_proj to_add_out"
        p.learning_rate = 1e-4
    elif workload == "k5":
        from train_sana import SanaModel as T
        p.model_config = model or "sana_1600m_1024"
        p.use_repa, p.use_preservation, p.preservation_ratio = True, True, 0.5
        p.lora_rank, p.lora_alpha, p.lora_algo = 16, 16, "lokr"
        p.lora_target_modules = "to_q to_k to_v to_out.0 conv_inverted conv_point linear_1 linear_2 proj"  # README.md:78
        p.learning_rate = 1e-4
    else:
        raise ValueError(workload)
    return T, p


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default=None, help="debug only: sana_tiny runs the same code small")
    ap.add_argument("--workload", default="k2", choices=sorted(WORKLOADS),
                    help="k2 = the metric's configuration (default); k3 / k4 / k5 = the other BASELINE configs")
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--no-other-workloads", action="store_true", help="skip the short k3 / k4 / k5 runs after the k2 line")
    ap.add_argument("--no-account", action="store_true", help="skip the profiler pass (GPU-busy / by-owner split)")
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
    installed, so the upstream scripts cannot be imported).  Rank 0 only.  `--warmup W` means
    the same as in the own arm (W untimed steps first) but is capped at 1: a CPU step of this size takes ~10 s and
    has no clock / cache state to settle; the cap is stated in the line (`warmup`, `warmup_cap`)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    if args.workload != "k2":
        print(json.dumps({"impl": "reference", "unavailable": "the CPU reference arm exists for the k2 metric only"}), flush=True)
        return
    model = args.model or "sana_1600m_1024"
    r = cpu_reference_steps(model, 1, max(1, args.steps), max(0, min(args.warmup, 1)), budget_s=240.0)
    line = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if model == "sana_1600m_1024" else f"DEBUG {model}",
                       "global_batch": 16 * args.gpus, "parallelism": f"dp{args.gpus}",
                       "text_padding": "fixed 512 (the reference's own padding, train_sana.py:167-179)",
                       "sample": "1 image per step on the host cores (bounded sample of the batch-16 workload)"}}
    if "unavailable" in r:
        line["unavailable"] = r["unavailable"]
    else:
        line.update({"value": r["value"], "steps": r["steps"], "warmup": r["warmup"], "warmup_cap": 1,
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
    the text embeddings already padded + masked (the optimize hooks accept that tuple)."""
    import torch
    from yat_b200.bucket_sampler import Batch
    b = Batch()
    b.ratio = batch.ratio
    b.vae_features = batch.vae_features.to(device, dtype=torch.bfloat16)
    if isinstance(batch.embeddings, tuple):                       # SD3.5: (joint text [B,333,4096], pooled [B,2048])
        b.embeddings = tuple(t.to(device) for t in batch.embeddings)
    elif hasattr(trainer, "pad_text"):
        b.embeddings = trainer.pad_text(batch.embeddings)
    else:
        from yat_b200.embeddings import pad_embeddings
        b.embeddings = pad_embeddings(batch.embeddings, trainer.max_text_tokens, device)
    b.repa_features = None if batch.repa_features is None else batch.repa_features.to(device)
    return b


def _batch_bytes(batch):
    n = batch.vae_features.numel() * batch.vae_features.element_size()
    n += sum(e.numel() * e.element_size() for e in batch.embeddings)
    if not isinstance(batch.embeddings, tuple):
        n += 4 * (len(batch.embeddings) + 1)     # int32 offsets
    if batch.repa_features is not None:
        n += batch.repa_features.numel() * batch.repa_features.element_size()
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
    # the fused q|k|v projection of one SANA-1.6B layer: own tcgen05 GEMM vs the library call it replaces (A/B)
    try:
        from yat_b200.gemm import gemm_bf16_tn
        xa = q.view(B * N, C)
        wq = (torch.randn((3 * C, C), device=device) / C ** 0.5).to(torch.bfloat16)
        fl = 2.0 * B * N * C * 3 * C
        for name, fn in (("gemm_qkv_own_tcgen05", lambda: gemm_bf16_tn(xa, wq)),
                         ("gemm_qkv_cublas", lambda: torch.nn.functional.linear(xa, wq))):
            ms = timeit(fn, iters=30, warm=5)
            out[name] = {"bound": "tensor", "achieved": fl / (ms * 1e-3) / 1e12, "peak": pk["bf16_tflops"],
                         "unit": "TFLOP/s", "us": ms * 1e3, "flops": fl, "shape": [B * N, 3 * C, C],
                         "peak_source": pk["source"] + " (bf16_tflops, burst: kernel timed alone)", "timing": "stand-alone"}
            out[name]["frac"] = out[name]["achieved"] / out[name]["peak"]
        out["gemm_qkv_own_tcgen05"]["vs_cublas_time"] = out["gemm_qkv_cublas"]["us"] / out["gemm_qkv_own_tcgen05"]["us"]
        # projection + linear attention forward, fused (phase 1 = the GEMM's epilogue; the attention streams only q)
        # vs the path the step runs today (library GEMM -> yat_linattn_fwd_ld).  FLOP = projection + attention.
        from yat_b200 import _lib as _L
        from yat_b200.linear_attention import sana_linear_attention_proj, sana_linear_attention_qkv
        x3 = q.view(B, N, C)
        fl_att = 4.0 * 32 * 33 * N * H * B
        with torch.no_grad():
            ms_f = timeit(lambda: sana_linear_attention_proj(x3, wq, H), iters=30, warm=5)
            ms_u = timeit(lambda: sana_linear_attention_qkv(torch.nn.functional.linear(x3, wq), H), iters=30, warm=5)
            L = _L.lib()
            qkv_buf = torch.empty((B, N, 3 * C), dtype=torch.bfloat16, device=device)
            st_buf = torch.empty(L.yat_qkv_linattn_state_bytes(B, N, H) // 4, dtype=torch.float32, device=device)
            o_buf = torch.empty((B, N, C), dtype=torch.bfloat16, device=device)
            cs = torch.cuda.current_stream().cuda_stream
            ms_g = timeit(lambda: L.yat_qkv_linattn_fwd(x3.data_ptr(), wq.data_ptr(), B, N, H, 32, C, qkv_buf.data_ptr(),
                                                        st_buf.data_ptr(), cs), iters=30, warm=5)
            ms_s = timeit(lambda: L.yat_linattn_fwd_state(qkv_buf.data_ptr(), 3 * C, st_buf.data_ptr(), B, N, H, 32,
                                                          o_buf.data_ptr(), cs), iters=30, warm=5)
        out["qkv_linattn_fused_gemm"] = {
            "bound": "tensor", "achieved": (fl + fl_att / 2) / (ms_g * 1e-3) / 1e12, "peak": pk["bf16_tflops"],
            "unit": "TFLOP/s", "us": ms_g * 1e3, "flops": fl + fl_att / 2, "shape": [B, N, C, 3 * C],
            "peak_source": pk["source"] + " (bf16_tflops, burst: kernel timed alone)", "timing": "stand-alone",
            "what": "yat_qkv_linattn_fwd: q|k|v projection + relu(k)^T v, sum relu(k) in its epilogue"}
        out["qkv_linattn_fused_gemm"]["frac"] = out["qkv_linattn_fused_gemm"]["achieved"] / pk["bf16_tflops"]
        entry("linattn_fwd_state", 4 * B * N * C, ms_s, [B, N, C], flops=fl_att / 2)
        out["qkv_linattn_fwd_path"] = {"fused_us": ms_f * 1e3, "unfused_us": ms_u * 1e3, "fused_over_unfused_time": ms_f / ms_u,
                                       "what": "projection + attention forward: (own fused GEMM + state kernel) vs "
                                               "(library GEMM + yat_linattn_fwd_ld)"}
        del wq, qkv_buf, st_buf, o_buf
    except Exception as e:  # noqa: BLE001 — reported, not hidden
        out["gemm_qkv_own_tcgen05"] = {"error": str(e)[:200]}
    # cross-attention forward of one SANA-1.6B layer (attn2: 20 heads x 112, 320 keys): own tcgen05 kernel vs the path the
    # library path (yat_head_pad x 2 + cuDNN SDPA with the key bias folded into an extra head column, YAT_XATTN_OWN_BACKWARD=0);
    # the step runs attn2 on the own forward + backward.
    try:
        from yat_b200 import _lib as _L
        from yat_b200.attention import _sdpa_key_bias_folded
        xh, xd, xl = 20, 112, 320
        xq = torch.randn((B, N, xh * xd), device=device).to(torch.bfloat16)
        xk = torch.randn((B, xl, xh * xd), device=device).to(torch.bfloat16)
        xv = torch.randn((B, xl, xh * xd), device=device).to(torch.bfloat16)
        xb = torch.zeros((B, xl), device=device)
        for b_, n_ in enumerate(torch.randint(20, 301, (B,), generator=torch.Generator().manual_seed(1)).tolist()):
            xb[b_, n_:] = -10000.0
        xo = torch.empty_like(xq)
        xlse = torch.empty((B, xh, N), dtype=torch.float32, device=device)
        cs = torch.cuda.current_stream().cuda_stream
        L = _L.lib()
        xb4 = xb.view(B, 1, 1, xl).to(torch.bfloat16)
        own = lambda: L.yat_xattn_fwd(xq.data_ptr(), xk.data_ptr(), xv.data_ptr(), xb.data_ptr(), B, xh, N, xl, xd,  # noqa: E731
                                      xh * xd, xh * xd, xd ** -0.5, xo.data_ptr(), xlse.data_ptr(), cs)
        ms_o = ms_l = float("inf")
        with torch.no_grad():   # both arms are clock-sensitive and the GEMM A/B above leaves the board at its power cap:
            for _ in range(3):  # alternate the arms and keep each one's best
                ms_l = min(ms_l, timeit(lambda: _sdpa_key_bias_folded(xq, xk, xv, xb4, xh, xd), iters=30, warm=5))
                ms_o = min(ms_o, timeit(own, iters=30, warm=5))
        fl_x = 4.0 * B * xh * N * xl * xd
        out["xattn_fwd_own_tcgen05"] = {"bound": "tensor", "achieved": fl_x / (ms_o * 1e-3) / 1e12, "peak": pk["bf16_tflops"],
                                        "unit": "TFLOP/s", "us": ms_o * 1e3, "flops": fl_x, "shape": [B, xh, N, xl, xd],
                                        "library_path_us": ms_l * 1e3, "vs_library_time": ms_l / ms_o,
                                        "timing": "stand-alone", "what": "yat_xattn_fwd vs head_pad + cuDNN SDPA (folded key bias)"}
        out["xattn_fwd_own_tcgen05"]["frac"] = out["xattn_fwd_own_tcgen05"]["achieved"] / pk["bf16_tflops"]
        # backward: own (prep + dq + dkv kernels) against the library path's backward ((fwd + bwd) - fwd of the folded SDPA)
        xdo = torch.randn_like(xq)
        xdq, xdk, xdv = torch.empty_like(xq), torch.empty_like(xk), torch.empty_like(xv)
        xws = torch.empty(L.yat_xattn_bwd_workspace_bytes(B, xh, N), dtype=torch.uint8, device=device)
        own()
        own_b = lambda: L.yat_xattn_bwd(xq.data_ptr(), xk.data_ptr(), xv.data_ptr(), xb.data_ptr(), xo.data_ptr(),  # noqa: E731
                                        xlse.data_ptr(), xdo.data_ptr(), B, xh, N, xl, xd, xh * xd, xh * xd, xd ** -0.5,
                                        xdq.data_ptr(), xdk.data_ptr(), xdv.data_ptr(), xws.data_ptr(), cs)
        qg_, kg_, vg_ = [t.clone().requires_grad_(True) for t in (xq, xk, xv)]

        def lib_fb():
            torch.autograd.grad(_sdpa_key_bias_folded(qg_, kg_, vg_, xb4, xh, xd), (qg_, kg_, vg_), xdo)

        ms_ob = ms_lfb = float("inf")
        for _ in range(3):
            ms_lfb = min(ms_lfb, timeit(lib_fb, iters=20, warm=3))
            ms_ob = min(ms_ob, timeit(own_b, iters=20, warm=3))
        fl_b = 2.5 * fl_x
        out["xattn_bwd_own_tcgen05"] = {"bound": "tensor", "achieved": fl_b / (ms_ob * 1e-3) / 1e12, "peak": pk["bf16_tflops"],
                                        "unit": "TFLOP/s", "us": ms_ob * 1e3, "flops": fl_b, "shape": [B, xh, N, xl, xd],
                                        "library_fwd_bwd_us": ms_lfb * 1e3, "library_bwd_us": (ms_lfb - ms_l) * 1e3,
                                        "vs_library_time": (ms_lfb - ms_l) / ms_ob, "timing": "stand-alone",
                                        "what": "yat_xattn_bwd (prep + dq + dkv) vs the backward of head_pad + cuDNN SDPA"}
        out["xattn_bwd_own_tcgen05"]["frac"] = out["xattn_bwd_own_tcgen05"]["achieved"] / pk["bf16_tflops"]
        del xq, xk, xv, xo, xlse, xdo, xdq, xdk, xdv, xws, qg_, kg_, vg_
    except Exception as e:  # noqa: BLE001 — reported, not hidden
        out["xattn_fwd_own_tcgen05"] = {"error": str(e)[:200]}
    from yat_b200.linear_attention import _launch_fwd
    ms = timeit(lambda: _launch_fwd(q, k, v, H, "sm100"))      # the C-ABI call (as for the backward below)
    entry("linattn_fwd_sm100", 8 * B * N * C, ms, [B, N, C], flops=4 * 32 * 33 * N * H * B)
    from yat_b200.linear_attention import _launch_bwd

    def bwd():     # the C-ABI call itself: at ~150 us the kernel is shorter than the host cost of torch.autograd.grad
        _launch_bwd(q, k, v, do, H, "sm100")
    try:
        ms = timeit(bwd)
        entry("linattn_bwd_sm100", 14 * B * N * C, ms, [B, N, C], flops=int(2.5 * 4 * 32 * 33 * N * H * B))
    except Exception as e:  # noqa: BLE001 — reported, not hidden
        out["linattn_bwd_sm100"] = {"error": str(e)[:200]}
    return out


def _build_trainer(workload, batch, ckpt, model, steps):
    """Trainer of a BASELINE config, initialised in a scratch directory (initialize() creates ./models)."""
    import tempfile
    T, params = make_trainer(workload, batch, ckpt, model)
    params.steps = steps
    cwd = os.getcwd()
    tmp = tempfile.mkdtemp(prefix="yat_bench_")
    os.chdir(tmp)
    try:
        trainer = T(params)            # Accelerator inside: init_process_group(nccl) when WORLD_SIZE > 1
        trainer.initialize()
    finally:
        os.chdir(cwd)
    return trainer


OWNERS = (("own", ("yat::",)), ("nccl", ("nccl",)),
          ("cublas", ("nvjet", "cublas", "gemm", "cutlass", "sm100_xmma", "sm90_xmma", "gemv", "splitK")),
          ("cudnn", ("cudnn", "fmha", "flash", "sdpa")))


def _owner(name):
    low = name.lower()
    for owner, keys in OWNERS:
        if any(k.lower() in low for k in keys):
            return owner
    if low.startswith("memcpy") or low.startswith("memset"):
        return "memcpy_memset"
    return "aten"


def _account(step_fn, n_steps, device, ms_per_step):
    """Where one step's time goes: a torch.profiler (CUPTI) capture of `n_steps` resident steps.  Kernel
    durations are summed by owner (own yat:: kernels / cuBLAS / cuDNN / ATen / NCCL / copies), `busy` is the union of
    all device intervals over the span from the first to the last of them, `idle` the rest.  Durations under the
    profiler are per-kernel device times (not cold-cache serialised like an ncu pass); launch gaps are slightly
    wider than unprofiled, so `idle_frac_profiled` is an upper bound - `sum_ms_over_unprofiled_step` divides the same
    kernel-time sum by the CUDA-event step time measured without the profiler."""
    import torch
    from torch.profiler import ProfilerActivity, profile
    torch.cuda.synchronize(device)
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for i in range(n_steps):
            step_fn(i)
        torch.cuda.synchronize(device)
    iv, by_owner, by_name = [], {}, {}
    for ev in prof.events():
        if getattr(ev, "device_type", None) is None or "CUDA" not in str(ev.device_type).upper():
            continue
        if getattr(ev, "is_user_annotation", False) or ev.name.startswith(("Optimizer.", "ProfilerStep", "## ")):
            continue                                                     # record_function ranges mirrored on the device
        t0, t1 = ev.time_range.start, ev.time_range.end
        if t1 <= t0:
            continue
        iv.append((t0, t1))
        dur = (t1 - t0) / 1e3 / n_steps                                  # ms per step
        o = _owner(ev.name)
        by_owner[o] = by_owner.get(o, 0.0) + dur
        e = by_name.setdefault(ev.name[:96], [0.0, 0, o])
        e[0] += dur
        e[1] += 1
    if not iv:
        return {"unavailable": "the profiler returned no device events"}
    iv.sort()
    busy, cur0, cur1 = 0.0, iv[0][0], iv[0][1]
    for t0, t1 in iv[1:]:
        if t0 > cur1:
            busy += cur1 - cur0
            cur0, cur1 = t0, t1
        else:
            cur1 = max(cur1, t1)
    busy += cur1 - cur0
    span = iv[-1][1] - iv[0][0]
    total = sum(by_owner.values())
    top = sorted(by_name.items(), key=lambda kv: -kv[1][0])[:14]
    return {"steps_profiled": n_steps, "launches_per_step": len(iv) / n_steps,
            "span_ms_per_step": span / 1e3 / n_steps, "busy_ms_per_step": busy / 1e3 / n_steps,
            "idle_frac_profiled": 1.0 - busy / span,
            "sum_kernel_ms_per_step": total, "sum_ms_over_unprofiled_step": total / ms_per_step,
            "by_owner_ms_per_step": {k: round(v, 3) for k, v in sorted(by_owner.items(), key=lambda kv: -kv[1])},
            "by_owner_frac_of_kernel_time": {k: round(v / total, 4) for k, v in by_owner.items()},
            "top_kernels": [{"name": n, "ms_per_step": round(v[0], 3), "launches_per_step": v[1] / n_steps,
                             "owner": v[2]} for n, v in top]}


def _measure(workload, args, steps, warmup, full):
    """Timed resident steps (`value`) and timed plugin-call steps with host batches (`e2e`) of one workload.
    `full` adds the in-situ kernel table (k2 line)."""
    import torch
    import torch.distributed as dist
    from yat_b200 import _lib
    batch = args.batch or WORKLOADS[workload]["batch"]
    trainer = _build_trainer(workload, batch, args.ckpt, args.model if workload == args.workload else None, steps)
    device = trainer.accelerator.device
    world = trainer.accelerator.num_processes
    if args.debug_linattn:
        for proc in getattr(trainer.pipe.transformer, "attn_processors", {}).values():
            if hasattr(proc, "impl"):
                proc.impl = args.debug_linattn
    it = iter(trainer.sampler)
    host_batches = [next(it) for _ in range(4)]
    dev_batches = [_device_batch(trainer, b, device) for b in host_batches]
    losses = []

    def step_resident(i):
        losses.append(trainer.train_step(dev_batches[i % len(dev_batches)]))

    def step_e2e(i):
        # the call a user makes: host Batch in pinned memory -> loss value back on the host
        losses.append(trainer.train_step(host_batches[i % len(host_batches)]).item())

    for i in range(warmup):
        step_resident(i)
    sink = [] if full else None
    _lib.timing_sink = sink
    n0 = _lib.total_launches()
    with ClockSampler(trainer.accelerator.local_process_index) as clk:
        ms_local = _timed_steps(step_resident, steps, dist, device)
    launches = _lib.total_launches() - n0
    _lib.timing_sink = None
    ms = _max_over_ranks(ms_local, dist, device)
    for i in range(max(1, warmup // 2)):
        step_e2e(i)
    ms_e2e = _max_over_ranks(_timed_steps(step_e2e, steps, dist, device), dist, device)
    res = {"trainer": trainer, "device": device, "world": world, "batch": batch, "ms": ms, "ms_local": ms_local,
           "ms_e2e": ms_e2e, "launches": launches, "clocks": clk.summary(), "sink": sink, "losses": losses,
           "h2d": _batch_bytes(host_batches[0]), "step_resident": step_resident,
           "value": world * batch * steps / (ms * 1e-3), "e2e_value": world * batch * steps / (ms_e2e * 1e-3),
           "mem_gib": torch.cuda.max_memory_allocated(device) / 2 ** 30}
    return res


def _other_workload(name, args):
    """Short run of another BASELINE config after the k2 line (3 timed steps each way after 2 warm-ups)."""
    import gc
    import torch
    try:
        torch.cuda.reset_peak_memory_stats()
        r = _measure(name, args, steps=3, warmup=2, full=False)
        out = {"workload": WORKLOADS[name]["workload"], "batch": r["batch"], "steps": 3, "warmup": 2,
               "ms_per_step": r["ms"] / 3, "value": r["value"], "unit": UNIT,
               "e2e": {"value": r["e2e_value"], "ms_per_step": r["ms_e2e"] / 3, "h2d_bytes_per_step": r["h2d"],
                       "d2h_bytes_per_step": 4},
               "gpu_launches": r["launches"], "final_loss": float(r["losses"][-1]), "peak_mem_gib": round(r["mem_gib"], 1)}
    except Exception as e:  # noqa: BLE001 — reported in the line, never hidden
        out = {"workload": WORKLOADS[name]["workload"], "error": f"{type(e).__name__}: {str(e)[:300]}"}
    r = None
    gc.collect()
    torch.cuda.empty_cache()
    return out


def run_own_arm(args):
    import gc
    import torch
    import torch.distributed as dist
    from yat_b200 import _lib, build as _build

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device visible (yat_b200 has no CPU fallback)")
    rank = int(os.environ.get("RANK", "0"))
    if rank == 0:
        _build.build()
    pk = peaks()
    _lib.lib()
    wl = args.workload
    r = _measure(wl, args, args.steps, args.warmup, full=True)
    trainer, device, world, B = r["trainer"], r["device"], r["world"], r["batch"]
    ms, ms_local, ms_e2e = r["ms"], r["ms_local"], r["ms_e2e"]
    cfg = trainer.pipe.transformer.config
    n_params = sum(p.numel() for p in trainer.model.parameters() if p.requires_grad)

    # ---- in-situ kernel durations from the events recorded inside the timed region
    kernels = {}
    if wl == "k2":
        chw = cfg.in_channels * cfg.sample_size * cfg.sample_size
        N = (cfg.sample_size // cfg.patch_size) ** 2
        C = cfg.num_attention_heads * cfg.attention_head_dim
        insitu = {}
        for name, e0, e1 in r["sink"]:
            insitu.setdefault(name, []).append(e0.elapsed_time(e1))
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

    # ---- where the step's time goes (profiler pass, not timed)
    account = None
    if not args.no_account:            # every rank runs the profiled steps (they contain the DDP all-reduce); rank 0 reports
        try:
            account = _account(r["step_resident"], 2, device, ms / args.steps)
        except Exception as e:  # noqa: BLE001
            account = {"unavailable": f"{type(e).__name__}: {str(e)[:200]}"}
    if dist.is_initialized():
        dist.barrier()
    final_loss = float(r["losses"][-1]) if r["losses"] else None
    ckpt_on = bool(getattr(trainer.pipe.transformer, "gradient_checkpointing", False))
    clocks, launches, h2d = r["clocks"], r["launches"], r["h2d"]
    value, e2e_value = r["value"], r["e2e_value"]
    if rank != 0:
        return
    # ---- rank 0 extras: stand-alone kernel roofline, CPU baseline, the other BASELINE configs
    r = trainer = None
    gc.collect()
    torch.cuda.empty_cache()
    standalone = {}
    if wl == "k2":
        try:
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
    if world == 1 and wl == "k2" and not args.no_cpu_baseline:
        try:
            c = cpu_reference_steps(args.model or "sana_1600m_1024", 1, 2, 0, budget_s=args.cpu_budget_s)
            cpu = ({"unavailable": c["unavailable"]} if "unavailable" in c else
                   {"value": c["value"], "unit": UNIT, "cores": c["cores"], "kind": "port", "sample": c["sample"],
                    "sec_per_step": c["sec_per_step"]})
        except Exception as e:  # noqa: BLE001
            cpu = {"unavailable": str(e)[:300]}
    others = None
    if world == 1 and wl == "k2" and not args.no_other_workloads and not args.model:
        others = {name: _other_workload(name, args) for name in ("k3", "k4", "k5")}
    debug = (f"DEBUG {args.model} " if args.model else "") + (f"DEBUG linattn={args.debug_linattn}" if args.debug_linattn else "")
    line = {
        "metric": METRIC if wl == "k2" else f"{wl} train img/s", "value": value, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": (debug + WORKLOADS[wl]["workload"]).strip(),
                   "global_batch": world * B, "parallelism": f"dp{world}", "optimizer": "AdamW bf16, clip + update as 2 own kernels",
                   "text_padding": TEXT_PADDING if wl in ("k2", "k5") else "fixed (hook default)",
                   "gradient_checkpointing": ckpt_on,
                   "l2": "no explicit flush: one step streams ~3.2 GB of weights + >10 GB of activations (>> 126 MB L2)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
        "account": account,
        "kernels": {"in_situ": kernels, "standalone": standalone},
        "cpu_baseline": cpu,
        "other_workloads": others,
        "final_loss": final_loss,
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
