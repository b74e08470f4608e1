"""CPU-only, world_size 2 over gloo: the sharded mapping step (views round-robin over ranks, one sum-allreduce of
the gradient arena, identical Adam update on every rank) reproduces the single-process step."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from monogs_b200.mapping import MappingStep


def _make_params():
    g = torch.Generator().manual_seed(0)
    return {"xyz": torch.randn(50, 3, generator=g).requires_grad_(True),
            "scaling": torch.randn(50, 3, generator=g).requires_grad_(True),
            "opacity": torch.randn(50, 1, generator=g).requires_grad_(True)}


def _views():
    g = torch.Generator().manual_seed(1)
    return [torch.randn(50, 3, generator=g) for _ in range(5)]   # odd count: ranks get 3 and 2 views


def _view_loss(params, view):      # stand-in for render + mapping loss: any differentiable per-view function
    return ((params["xyz"] * view).sum(1) * torch.sigmoid(params["opacity"][:, 0])).abs().mean() + \
           (torch.exp(params["scaling"]) * view.abs()).mean()


def _reg(params):                  # isotropic regulariser (slam_backend.py:244-246)
    s = torch.exp(params["scaling"])
    return 10 * (s - s.mean(dim=1, keepdim=True)).abs().mean()


def _run(steps, world=1, rank=0):
    params = _make_params()
    opt = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(params.values(), (1e-2, 5e-3, 5e-2))], eps=1e-15)
    ms = MappingStep(params, opt, _view_loss, _reg)
    losses = [float(ms.step(_views())) for _ in range(steps)]
    return {k: v.detach().clone() for k, v in params.items()}, losses


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        params, losses = _run(3, world, rank)
        q.put((rank, {k: v.numpy() for k, v in params.items()}, losses))
    finally:
        dist.destroy_process_group()


def test_sharded_mapping_equals_single_process():
    ref_params, ref_losses = _run(3)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    got = [q.get(timeout=120) for _ in range(2)]
    for p in procs: p.join(timeout=60)
    assert all(p.exitcode == 0 for p in procs)
    for rank, params, losses in got:
        for k in ref_params:
            assert torch.allclose(torch.from_numpy(params[k]), ref_params[k], rtol=1e-5, atol=1e-6), (rank, k)
        assert all(abs(a - b) <= 1e-5 * abs(b) for a, b in zip(losses, ref_losses))
    # both ranks hold identical parameters (replicated update, no broadcast needed)
    for k in ref_params:
        assert (got[0][1][k] == got[1][1][k]).all()
