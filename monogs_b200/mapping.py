"""Keyframe-window mapping step, data-parallel over GPUs (BASELINE.json config 4).

One mapping iteration of the reference (utils/slam_backend.py:157-333) renders every keyframe of the window,
sums the losses, back-propagates once and takes one Adam step on all Gaussian parameters.  The per-view
render -> loss -> backward calls are independent given the same Gaussians and autograd SUMS their Gaussian
gradients, so the views shard over ranks with no data-path exchange except one sum-allreduce of the gradient
arena (NCCL over NVLink/NVSwitch on the box, gloo in the CPU tests).  Parameters and Adam state are replicated;
every rank applies the identical update, so no parameter broadcast is needed.

All Gaussian gradients live in ONE contiguous fp32 arena (`grad_arena`): each parameter's .grad is a view into
it, autograd accumulates in place, and the collective is a single all_reduce over 4*(11+3M)*P bytes.
"""
from typing import Callable, Dict, List, Sequence

import torch
import torch.distributed as dist


def shard_views(n_views: int, world: int, rank: int) -> List[int]:
    """Round-robin assignment of view indices to ranks (8-keyframe window -> 4/2/1 views per rank at 2/4/8)."""
    return list(range(rank, n_views, world))


class GradArena:
    """Flat gradient buffer with per-parameter views; p.grad aliases the arena."""

    def __init__(self, params: Sequence[torch.Tensor]):
        self.params = list(params)
        n = sum(p.numel() for p in self.params)
        dev = self.params[0].device if self.params else "cpu"
        self.flat = torch.zeros(n, device=dev, dtype=torch.float32)
        self.views = []
        off = 0
        for p in self.params:
            v = self.flat[off:off + p.numel()].view(p.shape)
            self.views.append(v)
            off += p.numel()

    def attach(self):
        """Zero the arena and (re)install the aliases; call before the backward passes of an iteration."""
        self.flat.zero_()
        for p, v in zip(self.params, self.views):
            p.grad = v

    def nbytes(self):
        return self.flat.numel() * 4


class MappingStep:
    """One data-parallel mapping iteration.

    view_loss_fn(params: dict, view) -> scalar loss for one keyframe (render + mapping loss);
    reg_loss_fn(params) -> scalar regulariser that is not per-view (the isotropic term
    10*mean|s - mean(s)|, slam_backend.py:244-246), added on rank 0 only so the allreduced sum counts it once.
    """

    def __init__(self, params: Dict[str, torch.Tensor], optimizer, view_loss_fn: Callable,
                 reg_loss_fn: Callable = None, group=None):
        self.params = params
        self.opt = optimizer
        self.view_loss_fn = view_loss_fn
        self.reg_loss_fn = reg_loss_fn
        self.group = group
        self.arena = GradArena(list(params.values()))
        self.distributed = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if self.distributed else 1
        self.rank = dist.get_rank(group) if self.distributed else 0

    def step(self, views: Sequence) -> torch.Tensor:
        """views: the FULL window (identical on every rank); each rank renders its round-robin share.
        Returns the (allreduced) total loss as a 0-dim tensor on the parameter device, without syncing."""
        self.arena.attach()
        mine = shard_views(len(views), self.world, self.rank)
        total = torch.zeros((), device=self.arena.flat.device)
        for i in mine:
            loss = self.view_loss_fn(self.params, views[i])
            loss.backward()            # accumulates into the arena views in place
            total = total + loss.detach()
        if self.reg_loss_fn is not None and self.rank == 0:
            reg = self.reg_loss_fn(self.params)
            reg.backward()
            total = total + reg.detach()
        if self.distributed and self.world > 1:
            # the single exchange of the path: sum of per-Gaussian gradients over ranks
            dist.all_reduce(self.arena.flat, op=dist.ReduceOp.SUM, group=self.group)
            dist.all_reduce(total, op=dist.ReduceOp.SUM, group=self.group)
        self.opt.step()                # identical update on every rank (replicated params + Adam state)
        return total
