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


def _ddp_worker(rank, world, port, q):
    """DDP gradient averaging (MoCE-IR-main/src/train.py:131) as b200ir does it: each rank computes the gradients of its
    own batch shard (here with the CPU oracle), ONE sum all-reduce of the flat buffer, 1/world folded into the AdamW
    update.  Result must equal single-process training on the whole batch (mean-reduced L1 over equal shards)."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import b200ir
    from b200ir import parallel
    torch.set_num_threads(2)
    cfg = dict(dim=16, num_blocks=[1, 1, 1, 1], num_refinement_blocks=1)
    sd = O.synth_state_dict(cfg, seed=3)
    clean = O.synth_image(4, 16, 16, seed=7)
    deg = (clean + 0.1 * torch.randn(clean.shape, generator=torch.Generator().manual_seed(1))).clamp(0, 1)
    cl, dl = parallel.shard_batch(clean, rank, world), parallel.shard_batch(deg, rank, world)
    _, grads, _ = O.train_step_reference(sd, cfg, dl, cl, steps=1)
    names = sorted(grads)
    flat = torch.cat([grads[n].reshape(-1) for n in names])
    scale = b200ir.allreduce_flat_grads(flat)
    p0 = torch.cat([sd[n].reshape(-1) for n in names])
    p1, _, _ = O.adamw_reference(p0, flat * scale, torch.zeros_like(p0), torch.zeros_like(p0), 1)
    if rank == 0:
        _, gfull, new_sd = O.train_step_reference(sd, cfg, deg, clean, steps=1)
        gref = torch.cat([gfull[n].reshape(-1) for n in names])
        pref = torch.cat([new_sd[n].reshape(-1) for n in names])
        q.put((float((flat * scale - gref).abs().max() / gref.abs().max()), float((p1 - pref).abs().max()), scale))
    dist.barrier()
    dist.destroy_process_group()


def test_ddp_flat_allreduce_two_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_ddp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    gerr, perr, scale = q.get(timeout=240)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert scale == 0.5
    assert gerr <= 1e-5
    # Adam's first step is sign-like (lr * g/|g|): parameters whose tiny gradient flips sign between the two
    # summation orders may differ by up to 2*lr
    assert perr <= 2 * 2e-4 + 1e-6
