"""Gradient all-reduce sweep (BASELINE config 5): our peer-to-peer kernel vs NCCL, 1 KB - 1 GB.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        benchmarks/allreduce_sweep.py [--max-mb 1024] [--out gpurun_out/allreduce_N.json]

Per size: CUDA-event time of `iters` back-to-back calls after warm-up, max over ranks.  Bus bandwidth is
2(N-1)/N * bytes / time = the bytes every GPU must push through EACH direction of its NVLink port (it serves its
peers' reads of (N-1)/N of its buffer and writes its reduced slice to N-1 peers; symmetrically on the way in),
so the roofline is bus bandwidth / 770 GB/s, the pool's measured peer-copy rate per direction (B200_PROFILING.md).
`benchmarks/p2p_bw.py` measures what the link gives plain push / pull kernels on the same box (680-715 GB/s).
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
    ap.add_argument("--max-mb", type=float, default=1024)
    ap.add_argument("--out", default=None)
    ap.add_argument("--dtype", default="float32")
    args = ap.parse_args()
    from pytorchdistributed_b200.comm import ddp_setup, destroy_process_group
    from pytorchdistributed_b200.comm.p2p import P2PCollectives
    lr = ddp_setup()
    dev = torch.device("cuda", lr)
    rank, world = dist.get_rank(), dist.get_world_size()
    dtype = getattr(torch, args.dtype)
    esz = torch.empty((), dtype=dtype).element_size()
    be = P2PCollectives(None, dev)
    rows = []
    size = 1024
    while size <= args.max_mb * (1 << 20):
        n = size // esz
        ours = be.alloc(n, dtype, dev)
        ref = torch.empty(n, dtype=dtype, device=dev)
        ours.fill_(1.0)
        ref.fill_(1.0)
        iters = 50 if size <= (16 << 20) else 10
        t_ours = timed(lambda: be.all_reduce_avg(ours), iters, dev)
        t_nccl = timed(lambda: dist.all_reduce(ref, op=dist.ReduceOp.AVG), iters, dev)
        ok = bool(torch.allclose(ours.float(), torch.ones_like(ours, dtype=torch.float32)))
        t_one = None
        if dtype == torch.float32 and size <= (4 << 20):          # one-shot variant (one NVLink round trip)
            t_one = timed(lambda: be.all_reduce_oneshot(ours), iters, dev)
            ok = ok and bool(torch.allclose(ours, torch.ones_like(ours)))
        bus = 2.0 * (world - 1) / world * size
        row = {"bytes": size, "ours_us": round(t_ours, 2), "nccl_us": round(t_nccl, 2),
               "ours_busbw_gbs": round(bus / t_ours / 1e3, 1), "nccl_busbw_gbs": round(bus / t_nccl / 1e3, 1),
               "frac_nvlink_roofline_of_measured": round(bus / t_ours / 1e3 / NVLINK_GBS, 3), "correct": ok}
        if t_one is not None:
            row["ours_oneshot_us"] = round(t_one, 2)
        rows.append(row)
        if rank == 0:
            print(f"{size:>11d} B  ours {t_ours:9.1f} us ({row['ours_busbw_gbs']:7.1f} GB/s bus, "
                  f"{row['frac_nvlink_roofline_of_measured']:.2f} of roofline)   nccl {t_nccl:9.1f} us "
                  f"({row['nccl_busbw_gbs']:7.1f} GB/s)  speedup {t_nccl / t_ours:5.2f}x  ok={ok}"
                  + (f"  one-shot {t_one:7.1f} us" if t_one is not None else ""), flush=True)
        del ours, ref
        size *= 4
    if rank == 0 and args.out:
        json.dump({"world": world, "dtype": args.dtype, "rows": rows}, open(args.out, "w"), indent=1)
    be.close()
    destroy_process_group()


if __name__ == "__main__":
    main()
