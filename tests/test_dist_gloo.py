"""world_size-2 gloo test (CPU) of the N>1 host logic: batch sharding + gather reproduces the single-process result.
The per-rank compute here is the CPU oracle (the CUDA kernels need a GPU); what is under test is the sharding."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import restormer_oracle as O


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import b200ir  # noqa: F401
    from b200ir import parallel
    torch.set_num_threads(2)
    cfg = dict(dim=16, num_blocks=[1, 1, 1, 1], num_refinement_blocks=1)
    sd = O.synth_state_dict(cfg, seed=3)
    x = O.synth_image(5, 16, 16, seed=7)          # 5 samples over 2 ranks: shards of 3 and 2
    xl = parallel.shard_batch(x, rank, world)
    with torch.no_grad():
        yl = O.restormer_forward(sd, xl, cfg)
    y = parallel.gather_batch(yl, x.shape[0])
    if rank == 0:
        with torch.no_grad():
            ref = O.restormer_forward(sd, x, cfg)
        q.put(float((y - ref).abs().max()))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_gather_two_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    err = q.get(timeout=240)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert err <= 1e-5


def test_shard_range_covers_everything():
    import b200ir  # noqa: F401
    from b200ir import parallel
    for n in (1, 5, 16, 17):
        for w in (1, 2, 3, 8):
            r = [parallel.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(w - 1))
