"""Isolated gradient all-reduce timing: the two flat FP32 gradient vectors of the DefectGAN step (generator
180 MB, discriminator 50 MB) reduced with `dist.all_reduce(AVG)` exactly as `parallel.BucketReducer.finish` does, alone on the GPU.

    NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL torchrun --nproc-per-node N tools/allreduce_probe.py [--bucket-mb 32]

Prints one JSON line per size from rank 0: time per all-reduce (CUDA events on the launching stream, max over ranks), algorithm
bandwidth (bytes / time) and bus bandwidth (x 2(N-1)/N).  The NCCL_DEBUG lines of the same run say which algorithm / protocol /
channel count served it (NVLS or ring)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from de_i2i_gan_b200 import parallel  # noqa: E402


def timed(fn, iters, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--bucket-mb", type=int, default=32)
    args = ap.parse_args()
    rank, local, world = parallel.init_distributed()
    assert world > 1, "run under torchrun with N > 1"
    for name, mb in (("generator", 180), ("discriminator", 50), ("bucket", args.bucket_mb)):
        flat = torch.randn(mb * (1 << 20) // 4, device="cuda")
        ms = timed(lambda: dist.all_reduce(flat, op=dist.ReduceOp.AVG), args.iters)
        nbytes = flat.numel() * 4
        if rank == 0:
            alg = nbytes / (ms * 1e-3) / 1e9
            print(json.dumps(dict(probe="allreduce_fp32_avg", what=name, mbytes=mb, n_gpus=world, ms=round(ms, 4), algbw_gbs=round(alg, 1),
                                  busbw_gbs=round(alg * 2 * (world - 1) / world, 1), nccl=".".join(map(str, torch.cuda.nccl.version())),
                                  max_ctas=os.environ.get("NCCL_MAX_CTAS"))), flush=True)
        del flat
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
