"""The two figures of the reference's pipeline notebook, drawn from sweep files of ``benchmarks/pipeline_sweep.py``:

* ``mp_vs_single``: execution time of ``train()`` for model parallel / single GPU / pipeline (split 20) as bars with
  error bars - ``plot()`` in ``R/03 模型并行/03_model_parallel.ipynb#cell8, #cell9, #cell12`` (json L401-431, L564-573);
* ``split_size_tradeoff``: time against pipeline split size, one line per sweep file -
  ``#cell13`` (json L599-624; the reference saves ``split_size_tradeoff.png``).

SVG instead of PNG (there is no plotting library on the box):

    python benchmarks/plot_pipeline.py --out-dir profiles \\
        "ours graph=profiles/pipeline_sweep_r2_ours_graph_p2p.json" "reference bf16=profiles/pipeline_sweep_r2_reference_bf16.json"
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pytorchdistributed_b200.utils.svgplot import bar_chart, line_chart  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("files", nargs="+", help='"label=path.json" (or just a path)')
    ap.add_argument("--out-dir", default=".")
    ap.add_argument("--metric", default="device", choices=["device", "synced", "wall"],
                    help="device = CUDA-event time on cached inputs, synced = the notebook's procedure")
    args = ap.parse_args()
    mk, sk = f"{args.metric}_s_mean", f"{args.metric}_s_std"
    series = []
    for i, spec in enumerate(args.files):
        label, path = spec.split("=", 1) if "=" in spec else (os.path.basename(spec), spec)
        res = json.load(open(path))["results"]
        splits = sorted(int(k.split("_")[1]) for k in res if k.startswith("pipeline_"))
        series.append((label, splits, [res[f"pipeline_{s}"][mk] for s in splits], [res[f"pipeline_{s}"][sk] for s in splits]))
        if i == 0:
            names = [n for n in ("model_parallel", "single", "pipeline_20") if n in res]
            pretty = {"model_parallel": "Model Parallel", "single": "Single GPU", "pipeline_20": "Pipelining Model Parallel"}
            out = os.path.join(args.out_dir, "mp_vs_single.svg")
            bar_chart([res[n][mk] for n in names], [res[n][sk] for n in names], [pretty[n] for n in names],
                      title=f"ResNet50 Execution Time (Second) - {label}", ylabel="seconds per train()", path=out)
            print("wrote", out)
    out = os.path.join(args.out_dir, "split_size_tradeoff.svg")
    line_chart(series, title="ResNet50 Execution Time (Second)", xlabel="Pipeline Split Size",
               ylabel="seconds per train()", path=out)
    print("wrote", out)


if __name__ == "__main__":
    main()
