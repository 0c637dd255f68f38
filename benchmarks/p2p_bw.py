"""NVLink micro-benchmarks of the peer-to-peer kernels: pure remote read (broadcast pull) vs in-place all-reduce.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29571 benchmarks/p2p_bw.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def timed(fn, iters, dev):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    from pytorchdistributed_b200.comm import ddp_setup, destroy_process_group
    from pytorchdistributed_b200.comm.p2p import P2PCollectives
    lr = ddp_setup()
    dev = torch.device("cuda", lr)
    rank, world = dist.get_rank(), dist.get_world_size()
    be = P2PCollectives(None, dev)
    for mb in (64, 256):
        n = mb * (1 << 20) // 4
        buf = be.alloc(n, torch.float32, dev)
        buf.fill_(1.0)
        for blocks in (32, 64, 128):
            be.blocks = blocks
            t_b = timed(lambda: be.broadcast(buf, 0), 10, dev)
            t_a = timed(lambda: be.all_reduce_avg(buf), 10, dev)
            if rank == 0:
                print(f"{mb:4d} MB blocks={blocks:3d}  broadcast-pull {t_b:8.1f} us = {mb * 1.048576 / t_b * 1e3:7.1f} GB/s remote read per GPU   "
                      f"all-reduce {t_a:8.1f} us = {mb * 1.048576 * (world - 1) / world / t_a * 1e3:7.1f} GB/s per direction per GPU", flush=True)
        # one-directional pushes / pulls between rank 0 and rank 1 (every rank does the same towards its neighbour)
        from pytorchdistributed_b200 import _lib
        reg, off = be._loc(buf)
        peer = reg.ptrs[(rank + 1) % world] + off
        local2 = torch.empty(n, dtype=torch.float32, device=dev)
        st = torch.cuda.current_stream(dev).cuda_stream
        nbytes = n * 4
        res = {}
        res["push st.v4"] = timed(lambda: _lib.call("ptd_peer_copy", peer, local2, nbytes, st), 10, dev)
        res["pull ld.v4"] = timed(lambda: _lib.call("ptd_peer_copy", local2, peer, nbytes, st), 10, dev)
        for blocks in (74, 148, 296):
            res[f"push bulk b{blocks}"] = timed(lambda: _lib.call("ptd_peer_copy_bulk", peer, local2, nbytes, blocks, st), 10, dev)
        res["pull bulk b296"] = timed(lambda: _lib.call("ptd_peer_copy_bulk", local2, peer, nbytes, 296, st), 10, dev)
        if rank == 0:
            for k, v in res.items():
                print(f"{mb:4d} MB {k:16s} {v:8.1f} us = {mb * 1.048576 / v * 1e3:7.1f} GB/s per direction (all ranks concurrently)", flush=True)
        del buf
    be.close()
    destroy_process_group()


if __name__ == "__main__":
    main()
