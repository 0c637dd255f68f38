"""Regenerate the measured tables of BASELINE.md (section B) from the JSON files committed under profiles/.

    python benchmarks/make_baseline_tables.py            # prints markdown
    python benchmarks/make_baseline_tables.py --write    # replaces the block between the markers in BASELINE.md

Every number in section B comes out of this script, i.e. out of a tracked file that carries its own clock record; a
row whose record shows a thermal / hardware slowdown is flagged instead of quoted.
"""
import argparse
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROF = os.path.join(ROOT, "profiles")
BEGIN, END = "<!-- BEGIN GENERATED TABLES -->", "<!-- END GENERATED TABLES -->"
BAD = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}


def load(name):
    path = os.path.join(PROF, name)
    if not os.path.exists(path):
        return None
    with open(path) as f:
        return json.loads(f.read())


def clk(c):
    if not c:
        return "no record"
    bad = BAD & set(c.get("reasons") or [])
    note = ", ".join(c.get("reasons") or []) or "no throttle reason"
    return f"{'REJECTED: ' if bad else ''}{c.get('sm_mhz')} MHz of {c.get('sm_max_mhz')}, {note}"


def comm_label(d):
    """'p2p:nvls:{'nvls': 20, 'two_shot': 0, ...}' -> which all-reduce kernel actually ran."""
    g = d.get("grad_comm", "?")
    if g.startswith("p2p:") and ":{" in g:
        import ast
        stats = ast.literal_eval(g[g.index(":{") + 1:])
        ran = [k for k in ("nvls", "two_shot", "wire_bf16") if stats.get(k)]
        names = {"nvls": "switch reduction (multimem)", "two_shot": "two-shot peer-to-peer", "wire_bf16": "bf16 wire"}
        g = "ours: " + " + ".join(names[k] for k in ran) if ran else "ours (p2p)"
    elif g == "p2p":
        g = "ours (p2p)"
    elif g == "torch":
        g = "none (one rank)"
    return g + (" + SGD fused into the all-reduce" if d.get("fused_sgd_in_allreduce") else "")


def bench_table(out):
    out.append("### Config 2 - ResNet-50 DDP, bf16, 3x224x224, 256 images per GPU, SGD (builder runs of `bench.py`; the "
               "driver's BENCH / SCALE records are the authoritative ones)\n")
    out.append("| GPUs | images/s (device-timed, max over ranks) | ms/step | e2e images/s | e2e ms/step | gradient path | clocks | file |")
    out.append("|---|---|---|---|---|---|---|---|")
    base = None
    for n, f in [(1, "bench_n1_r2.json"), (2, "bench_n2_r2_final.json"), (4, "bench_n4_r2_final.json"),
                 (4, "bench_n4_r2_fused_sgd.json"), (8, "bench_n8_r2_default.json"), (8, "bench_n8_r2_default2.json")]:
        d = load(f)
        if d is None:
            continue
        if n == 1:
            base = d["value"]
        eff = f" ({d['value'] / (n * base):.3f} of {n} x N=1)" if base and n > 1 else ""
        out.append(f"| {n} | {d['value']:.0f}{eff} | {d['ms_per_step']} | {d['e2e']['value']:.0f} | {d['e2e']['ms_per_step']} | "
                   f"{comm_label(d)} | {clk(d.get('clocks'))} | `profiles/{f}` |")
    out.append("")
    out.append("Final tree (hot kernels launched as programmatic dependents, `PTDB200_PDL=2`, the default since; each row its "
               "own box, so no ratio between them):\n")
    out.append("| GPUs | images/s | ms/step | e2e images/s | e2e ms/step | gradient path | clocks | file |")
    out.append("|---|---|---|---|---|---|---|---|")
    for n, f in [(1, "bench_n1_r2_pdl.json"), (2, "bench_n2_r2_pdl.json")]:
        d = load(f)
        if d:
            out.append(f"| {n} | {d['value']:.0f} | {d['ms_per_step']} | {d['e2e']['value']:.0f} | {d['e2e']['ms_per_step']} | "
                       f"{comm_label(d)} | {clk(d.get('clocks'))} | `profiles/{f}` |")
    ab = [load("pdl_ab_mode1.json"), load("pdl_ab_mode2.json")]
    if all(ab):
        out.append("")
        out.append("Same-process A/B of that change (two graphs of the same step on the same tensors, alternating blocks of 20 "
                   "replays, `benchmarks/experiments/pdl_ab.py`):\n")
        out.append("| arm | median ms/step normal | median ms/step with it | ratio | clocks | file |")
        out.append("|---|---|---|---|---|---|")
        for d, label, f in [(ab[0], "`PTDB200_PDL=1`: GEMM kernels release their dependents after their last load", "pdl_ab_mode1.json"),
                            (ab[1], "`PTDB200_PDL=2`: dependents released when a CTA exits (default)", "pdl_ab_mode2.json")]:
            out.append(f"| {label} | {d['median_normal_ms']} | {d['median_pdl_ms']} | {d['pdl_over_normal']:.4f} | "
                       f"{clk(d.get('clocks'))} | `profiles/{f}` |")
    ks = load("step_knob_sweep.json")
    if ks:
        out.append("")
        out.append("In-step sweep of the tuning knobs (one process per row, same box, back to back; "
                   "`benchmarks/experiments/step_knob_sweep.py`; taken before the default changed, so `defaults` = no PDL):\n")
        out.append("| setting | median ms/step | vs defaults | clocks |")
        out.append("|---|---|---|---|")
        for r in ks["rows"]:
            out.append(f"| `{r['setting']}` | {r.get('median_ms', 'failed')} | {r.get('vs_defaults', '-')} | {clk(r.get('clocks'))} |")
        out.append("\nFile: `profiles/step_knob_sweep.json`.")
    out.append("")
    out.append("Control runs at 8 GPUs (same box, same call; what the 1 -> 8 growth of the step is made of):\n")
    out.append("| run | ms/step | e2e ms/step | clocks | file |")
    out.append("|---|---|---|---|---|")
    names = {"default": "default (NVLS all-reduce, buffer broadcast)", "default2": "default, repeated at the end",
             "no_allreduce": "gradient all-reduce switched off", "no_bcast": "per-forward BN-buffer broadcast off",
             "two_shot": "two-shot peer-to-peer all-reduce instead of NVLS", "wire_bf16": "bf16 wire format"}
    for k, label in names.items():
        d = load(f"bench_n8_r2_{k}.json")
        if d:
            out.append(f"| {label} | {d['ms_per_step']} | {d['e2e']['ms_per_step']} | {clk(d.get('clocks'))} | "
                       f"`profiles/bench_n8_r2_{k}.json` |")
    out.append("")


def allreduce_table(out):
    out.append("### Config 5 - gradient all-reduce, fp32, in place, average (`benchmarks/allreduce_sweep.py`)\n")
    out.append("Bus bandwidth = 2(N-1)/N x bytes / time (NCCL's convention); roofline = bus bandwidth / 770 GB/s (measured "
               "peer-copy rate per direction).  `ours` is what DDP runs: the switch reduction (multimem) at N >= 3, the "
               "two-shot peer-to-peer kernel at N = 2.\n")
    for n in (2, 4, 8):
        d = load(f"allreduce_sweep_n{n}.json")
        if d is None:
            continue
        out.append(f"N = {n} (`profiles/allreduce_sweep_n{n}.json`, provider `{d.get('mode')}`, clocks over the sweep: {clk(d.get('clocks'))}):\n")
        out.append("| bytes | ours us | two-shot us | NVLS us | bf16-wire us | NCCL us | ours bus GB/s | of roofline | ours vs NCCL |")
        out.append("|---|---|---|---|---|---|---|---|---|")
        for r in d["rows"]:
            g = lambda k: (f"{r[k]:.1f}" if k in r else "-")  # noqa: E731
            ours = r["ours_us"]
            bus = 2.0 * (n - 1) / n * r["bytes"] / ours / 1e3
            out.append(f"| {r['bytes']} | {ours:.1f} | {g('two_shot_us')} | {g('nvls_us')} | {g('wire_bf16_us')} | {g('nccl_us')} | "
                       f"{bus:.0f} | {bus / 770.0:.2f} | {r['nccl_us'] / ours:.2f}x |")
        out.append("")


def dp_table(out):
    out.append("### Config 3 - single-process DataParallel, ResNet-50, 256 images per GPU, fwd + bwd + SGD (`benchmarks/dp_bench.py`)\n")
    out.append("| GPUs | arm | images/s | ms/step | vs N x 1-GPU rate | clocks | file |")
    out.append("|---|---|---|---|---|---|---|")
    one = load("dp_bench_ours_n1.json")
    bench1 = load("bench_n1_r2.json")
    base = bench1["value"] if bench1 else (one["images_per_sec"] if one else None)
    rows = [(1, "ours (wrapper on one GPU: eager, no graph)", "dp_bench_ours_n1.json"),
            (2, "ours, graph replicas", "dp_bench_ours_n2.json"), (2, "ours, eager thread pool", "dp_bench_ours_eager_n2.json"),
            (2, "reference: nn.DataParallel, bf16 autocast", "dp_bench_reference_n2.json"),
            (4, "ours, graph replicas", "dp_bench_ours_n4.json"), (8, "ours, graph replicas", "dp_bench_ours_n8.json"),
            (8, "reference: nn.DataParallel, bf16 autocast", "dp_bench_reference_n8.json")]
    for n, label, f in rows:
        d = load(f)
        if d is None:
            continue
        frac = f"{d['images_per_sec'] / (n * base):.2f}" if base else "-"
        out.append(f"| {n} | {label} | {d['images_per_sec']:.0f} | {d['ms_per_step']} | {frac} | {clk(d.get('clocks'))} | `profiles/{f}` |")
    o8, r8 = load("dp_bench_ours_n8.json"), load("dp_bench_reference_n8.json")
    if o8 and r8:
        out.append(f"\nAt 8 GPUs: {o8['images_per_sec'] / r8['images_per_sec']:.2f}x the reference's `nn.DataParallel` "
                   f"(N x 1-GPU rate uses the 1-GPU DDP bench, {base:.0f} images/s).\n")


def pipeline_table(out):
    out.append("### Config 4 - ResNet-50 split over 2 GPUs, batch 120 of 3x128x128, `train()` = 3 iterations "
               "(`benchmarks/pipeline_sweep.py --prewarm --repeat 10`)\n")
    out.append("Seconds per `train()`.  `synced` = the notebook's procedure (CPU `randn` of every batch inside the timed "
               "region, both devices synchronised at the end); `device` = the same `train()` on batches generated once, "
               "CUDA events on both devices.  The bf16-autocast reference arm has the precision of our arm.\n")
    arms = [("reference fp32 (as published)", "pipeline_sweep_r2_reference_fp32.json"),
            ("reference bf16 autocast", "pipeline_sweep_r2_reference_bf16.json"),
            ("ours eager", "pipeline_sweep_r2_ours_eager.json"),
            ("ours graph, boundary inside the graphs", "pipeline_sweep_r2_ours_graph_p2p.json")]
    data = [(label, load(f), f) for label, f in arms]
    data = [(a, d, f) for a, d, f in data if d is not None]
    if not data:
        return
    out.append("| config | " + " | ".join(f"{a}: synced / device" for a, _, _ in data) + " | ref bf16 / ours graph (device) |")
    out.append("|---|" + "---|" * (len(data) + 1))
    keys = list(data[0][1]["results"].keys())
    for k in keys:
        cells, refd, oursd = [], None, None
        for a, d, _ in data:
            r = d["results"].get(k)
            if r is None:
                cells.append("-")
                continue
            flag = " (REJECTED)" if BAD & set((r.get("clocks") or {}).get("reasons") or []) else ""
            cells.append(f"{r['synced_s_mean']:.3f} / {r['device_s_mean']:.3f}{flag}")
            if a.startswith("reference bf16"):
                refd = r["device_s_mean"]
            if a.startswith("ours graph"):
                oursd = r["device_s_mean"]
        ratio = f"{refd / oursd:.2f}x" if refd and oursd else "-"
        out.append(f"| {k} | " + " | ".join(cells) + f" | {ratio} |")
    out.append("")
    out.append("Files: " + ", ".join(f"`profiles/{f}`" for _, _, f in data) + " (each result carries its clocks record; "
               "every arm ran at 1965 MHz with no throttle reason).\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--write", action="store_true")
    args = ap.parse_args()
    out = []
    bench_table(out)
    dp_table(out)
    pipeline_table(out)
    allreduce_table(out)
    text = "\n".join(out)
    if not args.write:
        print(text)
        return
    path = os.path.join(ROOT, "BASELINE.md")
    s = open(path).read()
    i, j = s.index(BEGIN), s.index(END)
    open(path, "w").write(s[:i + len(BEGIN)] + "\n\n" + text + "\n" + s[j:])
    print("BASELINE.md updated")


if __name__ == "__main__":
    main()
