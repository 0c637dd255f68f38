"""Gradient all-reduce sweep (BASELINE config 5): our kernels vs NCCL, 1 KB - 1 GB, plus DDP's real bucket sizes.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        benchmarks/allreduce_sweep.py [--max-mb 256] [--out gpurun_out/allreduce_N.json]

Arms per size (fp32 buffer, in place, average):
  ours        what ``P2PCollectives.all_reduce_avg`` picks on this box (``mode`` in the JSON: nvls = through the switch
              with multimem.ld_reduce / multimem.st, vmm / ipc = two-shot peer-to-peer);
  two_shot    the peer-to-peer kernel, forced (same buffer, unicast mappings);
  wire_bf16   fp32 in / fp32 out with bf16 on the wire (cast + reduce + cast back in one kernel);
  nccl        ``dist.all_reduce(AVG)`` - what the reference's DDP runs (R/02 DDP基本概念/ddp_gpus.py:35,41).

Per arm: CUDA-event time of `iters` back-to-back calls after 3 warm-up calls, max over ranks; buffers above 126 MB
exceed L2, smaller ones are re-written between arms.  Bus bandwidth is NCCL's convention, 2(N-1)/N * bytes / time (the
bytes a GPU must move through each direction of its port WITHOUT in-switch reduction); ``frac_roofline`` divides it by
770 GB/s, the pool's measured peer-copy rate per direction (B200_PROFILING.md) - a switch reduction can exceed 1.0
there because it moves fewer bytes ((1 + 1/N) * bytes per direction; ``frac_roofline_nvls`` uses that count).
Every row carries the SM clocks / throttle reasons sampled while it ran.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

NVLINK_GBS = 770.0


def timed(fn, iters, dev):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=float, default=256)
    ap.add_argument("--out", default=None)
    ap.add_argument("--mc-blocks", default="", help="comma-separated CTA counts: also time the switch kernel at each")
    ap.add_argument("--sizes", default="", help="comma-separated sizes in bytes (default: 1 KB x4 ... plus 25 MiB)")
    args = ap.parse_args()
    from pytorchdistributed_b200 import _lib
    from pytorchdistributed_b200.comm import ddp_setup, destroy_process_group
    from pytorchdistributed_b200.comm.p2p import P2PCollectives
    from pytorchdistributed_b200.utils import ClockSampler
    lr = ddp_setup()
    dev = torch.device("cuda", lr)
    rank, world = dist.get_rank(), dist.get_world_size()
    be = P2PCollectives(None, dev)
    if args.sizes:
        sizes = [int(x) for x in args.sizes.split(",")]
    else:
        sizes, s = [], 1024
        while s <= args.max_mb * (1 << 20):
            sizes.append(s)
            s *= 4
        sizes = sorted(set(sizes + [b for b in (25 << 20, 64 << 20) if b <= args.max_mb * (1 << 20)]))
    rows = []
    sampler = ClockSampler(lr, period_ms=50).start() if rank == 0 else None       # one sampler for the whole sweep
    for size in sizes:
        n = size // 4
        ours = be.alloc(n, torch.float32, dev)
        ref = torch.empty(n, dtype=torch.float32, device=dev)
        iters = 50 if size <= (16 << 20) else 10
        mark = sampler.mark() if sampler else 0
        res, ok = {}, {}
        reg, off = be._loc(ours)
        st = lambda: torch.cuda.current_stream(dev).cuda_stream  # noqa: E731
        pads = be._pad(0)[0]

        def two_shot():
            _lib.call("ptd_allreduce_inplace", be._ptrs(reg, off), pads, rank, world, (n + 3) // 4 * 4, 0, 1.0 / world,
                      be.blocks, st())

        def nvls():
            _lib.call("ptd_allreduce_mc", reg.mc + off, pads, be._pad(0)[1], rank, world, (n + 3) // 4 * 4, 0, 1.0 / world,
                      be.blocks, st())

        arms = [("ours", lambda: be.all_reduce_avg(ours, wire=False)), ("two_shot", two_shot)]
        if reg.mc:
            arms.append(("nvls", nvls))
        arms += [
                ("wire_bf16", lambda: be.all_reduce_avg(ours, wire=True)),
                ("nccl", lambda: dist.all_reduce(ref, op=dist.ReduceOp.AVG))]
        for name, fn in arms:
            ours.fill_(1.0)
            ref.fill_(1.0)
            res[name] = timed(fn, iters, dev)
            tgt = ref if name == "nccl" else ours
            ok[name] = bool(torch.allclose(tgt, torch.ones_like(tgt)))     # the average of ones stays one
        by_blocks = {}
        if args.mc_blocks and reg.mc and size >= (1 << 20):
            keep = be.blocks
            for b in [int(v) for v in args.mc_blocks.split(",")]:
                be.blocks = b
                by_blocks[b] = round(timed(nvls, iters, dev), 2)
            be.blocks = keep
        clocks = sampler.summary(since=mark) if sampler else {}     # the samples taken while this row ran
        bus = 2.0 * (world - 1) / world * size
        nvls_bytes = (1.0 + 1.0 / world) * size
        row = {"bytes": size, "mode": be.mode, "clocks": clocks}
        for name, t in res.items():
            row[f"{name}_us"] = round(t, 2)
            row[f"{name}_busbw_gbs"] = round(bus / t / 1e3, 1)
            row[f"{name}_ok"] = ok[name]
        row["frac_roofline_of_measured"] = round(bus / res["ours"] / 1e3 / NVLINK_GBS, 3)
        if be.mode == "nvls":
            row["frac_roofline_nvls_of_measured"] = round(nvls_bytes / res["ours"] / 1e3 / NVLINK_GBS, 3)
        row["speedup_vs_nccl"] = round(res["nccl"] / res["ours"], 3)
        if by_blocks:
            row["nvls_us_by_blocks"] = by_blocks
        rows.append(row)
        if rank == 0:
            print(f"{size:>11d} B  " + "  ".join(f"{k} {v:8.1f} us" for k, v in res.items())
                  + f"   ours {row['ours_busbw_gbs']:7.1f} GB/s bus = {row['frac_roofline_of_measured']:.2f} of roofline,"
                  f" {row['speedup_vs_nccl']:.2f}x nccl  ok={all(ok.values())}  clocks={clocks.get('sm_mhz')} {clocks.get('reasons')}"
                  + (f"  nvls by blocks {by_blocks}" if by_blocks else ""), flush=True)
        del ours, ref
    clocks_all = sampler.stop() if sampler else {}
    for row in rows:                  # rows shorter than the sampling period: the whole sweep's record stands in
        if not row["clocks"].get("samples"):
            row["clocks"] = dict(clocks_all, note="whole-sweep record (row shorter than the sampling period)")
    if rank == 0 and args.out:
        json.dump({"world": world, "dtype": "float32", "mode": be.mode, "mode_note": be.mode_note, "clocks": clocks_all,
                   "nvlink_gbs_per_direction_measured": NVLINK_GBS, "rows": rows}, open(args.out, "w"), indent=1)
    be.close()
    destroy_process_group()


if __name__ == "__main__":
    main()
