"""Turn one GPU round's ncu output (gpurun_out/) into the tracked summaries under profiles/.

    python scripts/summarize_profiles.py r01a [--steps-in-list 4]

Writes profiles/<tag>_launches.md   per-kernel totals of the `ncu --metrics gpu__time_duration.sum` launch
                                    list of `bench.py --steps 1 --warmup 1` (cold-cache, serialised:
                                    the SHARES are what counts, not the absolute times)
       profiles/<tag>_kernels.md    the `ncu --set full` capture of our kernels: duration, DRAM bytes,
                                    DRAM / SM throughput, pipe utilisation, registers, occupancy
       profiles/traffic.json        dram bytes per launch per kernel (bench.py reads it for roofline.traffic)
       profiles/<tag>_bench.json    the bench line of the same run
"""
import collections
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

SHORT = [
    (r"yat::.*?(\w+_kernel)", lambda m: "yat::" + m.group(1)),
    (r"nvjet_\w+", lambda m: "cuBLAS " + m.group(0)),
    (r"cutlass3x_\w+?(dgrad|wgrad|fprop)", lambda m: "cuDNN implicit-gemm " + m.group(1)),
    (r"cudnn_generated_\w*?sdpa_sm100_flash_(fprop|bprop)", lambda m: "cuDNN SDPA " + m.group(1)),
    (r"(dgrad2d|wgrad2d|conv2d)_c1_k1_nhwc\w*", lambda m: "cuDNN depthwise " + m.group(0)),
    (r"multi_tensor_apply_kernel.*?(\w+Functor)", lambda m: "at::multi_tensor " + m.group(1)),
    (r"at::(?:native::)?(?:<unnamed>::)?(\w+_kernel).*?(CUDAFunctor_add|MulFunctor|direct_copy\w*|FillFunctor|"
     r"sum_functor|MeanOps|NormTwoOps|bfloat16_copy\w*|GeluCUDA\w*|silu\w*|sigmoid\w*)?", None),
]


def short_name(n):
    for pat, fn in SHORT[:-1]:
        m = re.search(pat, n)
        if m:
            return fn(m)
    m = re.search(r"at::(?:native::)?(?:<unnamed>::)?(\w+)", n)
    if m:
        extra = re.findall(r"(CUDAFunctor_add|CUDAFunctorOnSelf_add|MulFunctor|direct_copy_kernel_cuda|FillFunctor|"
                           r"sum_functor|MeanOps|NormTwoOps|bfloat16_copy_kernel_cuda|Gelu\w+|silu\w*|sigmoid\w*|"
                           r"CatArrayBatchedCopy\w*|layer_norm\w*)", n)
        return "at::" + m.group(1) + ((" " + extra[0]) if extra else "")
    return re.sub(r"<.*", "", n)[:60]


def launches(tag, steps_in_list):
    path = os.path.join(OUT, "launches.csv")
    if not os.path.exists(path):
        return
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    total = 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(row["Metric Unit"], 1.0)
        k = short_name(row["Kernel Name"])
        agg[k][0] += 1
        agg[k][1] += v
        total += v
    ours = sum(t for k, (c, t) in agg.items() if k.startswith("yat::"))
    with open(os.path.join(PROF, f"{tag}_launches.md"), "w") as f:
        f.write(f"# {tag}: launch list of `bench.py --steps 1 --warmup 1 --no-cpu-baseline` under "
                "`ncu --metrics gpu__time_duration.sum --clock-control none`\n\n")
        f.write(f"{sum(c for c, _ in agg.values())} launches, {total / 1e3:.1f} ms of kernel time over "
                f"{steps_in_list} train steps (1 warm-up + 1 timed + the e2e pair) + the stand-alone kernel loop; "
                f"cold-cache serialised times — read the shares.  Our kernels: {100 * ours / total:.1f} % of the total.\n\n")
        f.write("| kernel | launches | total ms | share | avg us |\n|---|---:|---:|---:|---:|\n")
        for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:45]:
            f.write(f"| `{k}` | {c} | {t / 1e3:.2f} | {100 * t / total:.1f} % | {t / c:.1f} |\n")


METRICS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram thr %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm thr %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu pipe %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma pipe %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu pipe %"),
    ("smsp__inst_executed.sum", "warp insts"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
]


def to_bytes(v, unit):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def kernels(tag):
    reps = sorted(p for p in os.listdir(OUT) if p.endswith(".ncu-rep"))
    if not reps:
        return
    traffic_path = os.path.join(PROF, "traffic.json")
    traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
    with open(os.path.join(PROF, f"{tag}_kernels.md"), "w") as f:
        f.write(f"# {tag}: `ncu --set full --clock-control none --import-source on` of scripts/prof_kernels.py\n\n"
                "Shapes: noise/loss `[4096,32,32,32]` bf16 (bulk slab > L2), REPA z `[64,1024,768]`, linear "
                "attention / block ops `[16,1024,2240]`, GLUMB `[16,32,32,2x5600]` (SANA-1.6B layer), optimiser 321 M bf16 params.  Numbers taken under the profiler are evidence "
                "for WHY, never bench values.\n\n")
        for rep in reps:
            r = subprocess.run(["ncu", "-i", os.path.join(OUT, rep), "--page", "raw", "--csv"], capture_output=True,
                               text=True)
            rows = list(csv.reader(r.stdout.splitlines()))
            if len(rows) < 3:
                continue
            hdr, units = rows[0], rows[1]
            col = {h: i for i, h in enumerate(hdr)}
            f.write(f"## {rep}\n\n| kernel | " + " | ".join(n for _, n in METRICS) + " |\n|---|" + "---:|" * len(METRICS) + "\n")
            for row in rows[2:]:
                name = re.sub(r".*?(\w+_kernel).*", r"\1", row[col["Kernel Name"]])
                cells = []
                for m, _ in METRICS:
                    if m not in col:
                        cells.append("-")
                        continue
                    v, u = row[col[m]], units[col[m]]
                    try:
                        x = float(v.replace(",", ""))
                        cells.append(f"{x:.1f} {u}" if u not in ("", "inst") and not u.startswith("register") else f"{x:.0f}")
                    except ValueError:
                        cells.append(v)
                f.write(f"| `{name}` | " + " | ".join(cells) + " |\n")
                try:
                    rd = to_bytes(row[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
                    wr = to_bytes(row[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
                    key = {"noise_bf16_kernel": "noise_flowmatch", "loss_fused_kernel": "loss_fused",
                           "linattn_fwd_sm100_kernel": "linattn_fwd_sm100", "linattn_bwd_sm100_kernel": "linattn_bwd_sm100",
                           "repa_pool_cos_kernel": "repa_cos", "glumb_fwd_kernel": "glumb_mid_fwd",
                           "glumb_bwd_kernel": "glumb_mid_bwd", "adamw_kernel": "adamw_step",
                           "grad_sumsq_kernel": "grad_clip_coef", "ln_modulate_fwd_kernel": "ln_modulate_fwd",
                           "gate_residual_fwd_kernel": "gate_residual_fwd"}.get(name, name)
                    traffic[key] = {"dram_bytes_per_launch": rd + wr, "read": rd, "write": wr, "capture": f"{tag}/{rep}"}
                except (KeyError, ValueError):
                    pass
            f.write("\n")
    with open(traffic_path, "w") as f:
        json.dump(traffic, f, indent=1, sort_keys=True)


def bench(tag):
    p = os.path.join(OUT, "bench.log")
    if os.path.exists(p):
        for line in open(p):
            if line.startswith("{"):
                with open(os.path.join(PROF, f"{tag}_bench.json"), "w") as f:
                    json.dump(json.loads(line), f, indent=1)
                    f.write("\n")


if __name__ == "__main__":
    tag = sys.argv[1]
    steps = 4
    if "--steps-in-list" in sys.argv:
        steps = int(sys.argv[sys.argv.index("--steps-in-list") + 1])
    os.makedirs(PROF, exist_ok=True)
    launches(tag, steps)
    kernels(tag)
    bench(tag)
    print(sorted(os.listdir(PROF)))
