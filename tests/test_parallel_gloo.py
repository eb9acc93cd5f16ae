"""Data-parallel plumbing on CPU: world_size-2 gloo processes run the BucketReducer (de_i2i_gan_b200/parallel.py) over a
flat gradient vector laid out by FlatAdam._flatten and must end with the average of the per-rank gradients - which, for a
mean loss over equal shards, is the gradient of the global batch.  No CUDA, no C-ABI compute calls."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from de_i2i_gan_b200 import parallel
    from de_i2i_gan_b200.optim import FlatAdam
    r, _, w = parallel.init_distributed("gloo")
    assert (r, w) == (rank, world)
    torch.manual_seed(100 + rank)  # deliberately different initial weights: broadcast_module must fix that
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.BatchNorm2d(8), torch.nn.Conv2d(8, 4, 3), torch.nn.Flatten(),
                              torch.nn.Linear(4 * 6 * 6, 2))
    parallel.broadcast_module(net)
    flat = FlatAdam._flatten(list(net.parameters()))
    slices = [(p, o, p.numel()) for p, o in zip(flat["params"], flat["offsets"])]
    red = parallel.BucketReducer(flat["g"], slices, bucket_bytes=1024)  # small buckets: several per network
    assert len(red.buckets) > 2
    g = torch.Generator().manual_seed(7)
    x_all, y_all = torch.randn(8, 3, 8, 8, generator=g), torch.randn(8, 2, generator=g)
    shard = slice(rank * 4, rank * 4 + 4)
    net.eval()  # BN in eval mode: the sharded mean loss is then exactly the global mean loss
    for _ in range(2):  # second round checks that finish() re-arms the bucket counters
        flat["g"].zero_()
        loss = ((net(x_all[shard]) - y_all[shard]) ** 2).mean()
        loss.backward()
        red.finish()
    # single-process reference on the global batch with the broadcast weights
    ref = [torch.zeros_like(p) for p in net.parameters()]
    full = ((net(x_all) - y_all) ** 2).mean()
    for t, gr in zip(ref, torch.autograd.grad(full, list(net.parameters()))):
        t.copy_(gr)
    err = max((p.grad - t).abs().max().item() for p, t in zip(net.parameters(), ref))
    w0 = torch.cat([p.detach().reshape(-1) for p in net.parameters()])
    gathered = [torch.zeros_like(w0) for _ in range(world)]
    dist.all_gather(gathered, w0)
    same = all(torch.equal(gathered[0], t) for t in gathered)
    if rank == 0:
        out.put((err, same))
    dist.destroy_process_group()


def test_bucket_reducer_world2_gloo():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0, f"worker exit code {p.exitcode}"
    err, same = out.get(timeout=5)
    assert same, "broadcast_module left different weights on the ranks"
    assert err < 1e-6, err
