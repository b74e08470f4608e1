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
import ctypes as C
import os
from typing import Callable, Dict, List, NamedTuple, Optional, Sequence

import torch
import torch.distributed as dist

TILE_SEGMENT = 1024     # MGS_TILE_SEGMENT (include/monogs_b200.h)


def assign_views(n_views: int, world: int, n_window: Optional[int] = None) -> List[List[int]]:
    """View indices per rank: the first n_window views — the keyframe window — go round-robin (8-keyframe
    window -> 4/2/1 views per rank at 2/4/8 ranks); the remaining ones — the two random keyframes outside the window that
    BackEnd.map renders beside it (utils/slam_backend.py:215-242) — go one by one to the least-loaded rank (ties: the
    highest rank, away from rank 0, which also carries the regulariser)."""
    n_window = n_views if n_window is None else min(n_window, n_views)
    per = [list(range(r, n_window, world)) for r in range(world)]
    for i in range(n_window, n_views):
        r = min(range(world), key=lambda q: (len(per[q]), -q))
        per[r].append(i)
    return per


def shard_views(n_views: int, world: int, rank: int, n_window: Optional[int] = None) -> List[int]:
    """This rank's share of assign_views()."""
    return assign_views(n_views, world, n_window)[rank]


def expon_lr(step, lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """gaussian_splatting/utils/general_utils.py:80-95 (`helper`): log-linear decay from lr_init to lr_final over max_steps
    with an optional sine warm-up — the xyz learning rate GaussianModel.update_learning_rate installs every iteration
    (gaussian_model.py:298-312)."""
    import math
    if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
        return 0.0
    if lr_delay_steps > 0:
        delay_rate = lr_delay_mult + (1 - lr_delay_mult) * math.sin(0.5 * math.pi * min(max(step / lr_delay_steps, 0.0), 1.0))
    else:
        delay_rate = 1.0
    t = min(max(step / max_steps, 0.0), 1.0)
    return delay_rate * math.exp(math.log(lr_init) * (1 - t) + math.log(lr_final) * t)


class GradArena:
    """Flat gradient buffer with per-parameter views; p.grad aliases the arena."""

    def __init__(self, params: Sequence[torch.Tensor], extra: int = 0, flat: Optional[torch.Tensor] = None):
        """flat: an existing zero-filled f32 buffer of exactly the arena's length to use instead of a new allocation
        (the peer-visible gradient arena of monogs_b200.peer.PeerBlock)."""
        self.params = list(params)
        # every view starts on a 16-byte boundary: the kernels read / write [P,4] rows as float4
        pad4 = lambda n: (n + 3) // 4 * 4
        n = sum(pad4(p.numel()) for p in self.params)
        dev = self.params[0].device if self.params else "cpu"
        # `extra` trailing floats ride along in the same all-reduce (the summed loss of a mapping iteration)
        self.numel = n + pad4(extra)
        if flat is None:
            flat = torch.zeros(self.numel, device=dev, dtype=torch.float32)
        assert flat.numel() == self.numel and flat.dtype == torch.float32 and flat.is_contiguous()
        self.flat = flat
        self.extra = self.flat[n:n + extra]
        self.views = self.views_over(self.flat)
        # exclusive end of every parameter's (padded) segment, then the trailing extra segment
        self.segment_ends, off = [], 0
        for p in self.params:
            off += pad4(p.numel())
            self.segment_ends.append(off)
        if extra:
            self.segment_ends.append(self.numel)

    def views_over(self, flat: torch.Tensor):
        """Per-parameter views with this arena's layout over another flat buffer of the same length."""
        pad4 = lambda n: (n + 3) // 4 * 4
        views, off = [], 0
        for p in self.params:
            views.append(flat[off:off + p.numel()].view(p.shape))
            off += pad4(p.numel())
        return views

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


class MapView(NamedTuple):
    """One keyframe as the mapping engine consumes it (the fields BackEnd.map reads from a Camera,
    utils/camera_utils.py:94-108, plus its target images)."""
    settings: object                      # GaussianRasterizationSettings (matrices on the device)
    gt_image: torch.Tensor                # f32[3,H,W]
    gt_depth: Optional[torch.Tensor] = None     # f32[1,H,W]; None = monocular loss
    exposure_a: Optional[torch.Tensor] = None   # f32[1] device tensors; None = initialization (image_ab = image)
    exposure_b: Optional[torch.Tensor] = None
    update_stats: bool = True


class MapViewResult(NamedTuple):
    radii: torch.Tensor        # i32[P]  (visibility_filter = radii > 0)
    n_touched: torch.Tensor    # i32[P]  (occ_aware_visibility = n_touched > 0)
    tau: torch.Tensor          # f32[6]  pose tangent (rho, theta) of this view's loss
    dab: torch.Tensor          # f32[2]  dL/d(exposure_a, exposure_b)
    loss: torch.Tensor         # f32[1]


class MapEngine:
    """Device-resident mapping iteration over the C-ABI (mgs_map_begin / mgs_map_view_fwd_bwd / mgs_map_end).

    Holds the RAW Gaussian parameters (gaussian_model.py:77-97: _xyz, _features_dc, _features_rest, _opacity, _scaling,
    _rotation), one contiguous gradient arena that aliases their .grad, the densification statistics and a fused Adam.
    `step(views)` = one iteration of BackEnd.map (slam_backend.py:157-333) up to and including optimizer.step():
    activations once, every view of this rank's round-robin share rendered + back-propagated in place, the isotropic
    regulariser on rank 0, one sum-allreduce of the arena (NCCL over NVLink/NVSwitch; gloo in the CPU tests can not run
    this class — it needs the CUDA library), fused Adam on every rank.  Nothing in the iteration synchronises the host.

    exchange="peer": the all-reduce + Adam pair becomes ONE kernel over NVLink peer mappings (mgs_peer_reduce_adam,
    monogs_b200.peer): parameters and gradients live in a peer-visible block, every rank reduces and updates its
    1/world slice of the arena and stores the new parameters into every rank's block; Adam moments are sharded (each
    rank holds the live moments of its own slice; sync_stats() gathers them before a densify).
    exchange="nvls": the same kernel over an NVSwitch multicast mapping of the blocks (torch symmetric memory): the
    gradient sum happens inside the switch (multimem.ld_reduce) and one multimem.st writes the parameters of all ranks.
    exchange="auto": "peer" for 2..7 ranks of one node, "nvls" for 8 (falling back to "peer" without multicast), else "nccl".
    """

    NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")

    def __init__(self, params: Dict[str, torch.Tensor], lrs: Dict[str, float], sh_degree: int = 0, W: int = 640, H: int = 480,
                 capacity_headroom: float = 2.0, group=None, regulariser_weight: float = 10.0,
                 rgb_boundary_threshold: float = 0.01, alpha: float = 0.95, fast_math: bool = True, overlap: bool = True,
                 exchange: str = "nccl", max_views: int = 16, use_graph: bool = False, tile_segments: bool = True):
        from . import _lib as L
        from .ops import FusedAdam
        # tile_segments: fixed-segment binning of every view (mgs_map_view.tile_segments: the preprocess kernel scatters
        # the instances into per-tile segments of 1024 slots, one kernel sorts them; no scan, emit or long-list sort).
        # Used while the longest tile list of the window leaves 1.3x headroom; an overflow or a list within 10 % of a
        # segment switches the engine back to the general binning at the next poll_capacity().
        self.tile_segments = bool(tile_segments) and os.environ.get("MGS_TILE_SEGMENTS", "1") != "0"
        self._tile_max = 0
        self.fast_math = fast_math   # ex2.approx exponent in the blend kernels (relative error ~4e-7), like PoseTracker
        # overlap: the forward half of view k+1 (preprocess, binning, blend + loss) runs on the caller's stream beside the
        # backward half of view k on a second stream; two workspaces and two loss-gradient slots alternate
        self.overlap = overlap
        self._bwd_streams = self._fwd2_stream = None
        self.n_pipe = 4             # workspaces / slots in flight: two forwards + two backwards
        self.L, self.lib = L, L.lib()
        self.params = {k: params[k] for k in self.NAMES if k in params and params[k] is not None and params[k].numel() > 0}
        self.dev = self.params["xyz"].device
        self.P = self.params["xyz"].size(0)
        self.M = 1 + (self.params["f_rest"].size(1) if "f_rest" in self.params else 0)
        self.sh_degree = sh_degree
        self.isotropic = self.params["scaling"].size(-1) == 1
        for p in self.params.values():
            assert p.is_contiguous() and p.dtype == torch.float32 and p.device == self.dev
        # Side segment behind the gradients, riding in the same exchange (sum over ranks): [0] the iteration's loss,
        # [4 + 20 i ..] the pose record (T_w2c 16, exposure a, b, 0, 0) of window view i, written by its owner only —
        # the all-gather of the window poses at no extra collective
        self.max_views = max_views
        self.side = 4 + 20 * max_views
        self.arena = GradArena(list(self.params.values()), extra=self.side)
        self.grads = dict(zip(self.params.keys(), self.arena.views))
        self.lrs = dict(lrs)
        # use_graph: step() of a window installed with set_window() is captured once into a CUDA graph and replayed (the
        # Adam step number and learning rates then live on the device)
        self.use_graph = use_graph
        self.opt = FusedAdam([{"params": [p], "lr": lrs[k]} for k, p in self.params.items()], lr=0.0, eps=1e-15,
                             capturable=use_graph)
        self._win = None            # set_window(): views, per-view pose / optimiser state
        f = dict(device=self.dev, dtype=torch.float32)
        self.xyz_gradient_accum, self.denom = torch.zeros(self.P, **f), torch.zeros(self.P, **f)
        self.max_radii2D = torch.zeros(self.P, **f)
        self.loss_total = self.arena.extra[:1]
        self.W, self.H = W, H
        self.scratch = torch.empty(self.lib.mgs_map_scratch_bytes(self.P, self.M, W, H), device=self.dev, dtype=torch.uint8)
        self.color, self.depth, self.opacity = torch.empty(3, H, W, **f), torch.empty(1, H, W, **f), torch.empty(1, H, W, **f)
        self._out_extra = []
        self.headroom = capacity_headroom
        self.R_cap, self.ws, self.ws_extra = 0, None, []
        self._seen_max, self.overflow_polls, self.poll_every, self._since_poll = 0, 0, 16, 0
        self._status = torch.zeros(8, 8, dtype=torch.int32).pin_memory() if self.dev.type == "cuda" else None
        self._status_ev = torch.cuda.Event() if self.dev.type == "cuda" else None
        self._status_pending, self._status_n, self._graph = False, 0, None
        self.reg_weight, self.thr, self.alpha = regulariser_weight, rgb_boundary_threshold, alpha
        self.group = group
        self.distributed = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if self.distributed else 1
        self.rank = dist.get_rank(group) if self.distributed else 0
        self._keep = []
        self._slots = {}
        self._concurrent = False
        assert exchange in ("nccl", "peer", "nvls", "auto")
        if exchange == "auto":
            # the measured choice: the unicast kernel up to 4 ranks, the multicast kernel at 8, NCCL whenever
            # the group is not one NVLink domain of at most 8 CUDA ranks of this node
            one_node = int(os.environ.get("LOCAL_WORLD_SIZE", self.world)) == self.world
            fused = (self.distributed and 1 < self.world <= 8 and one_node and self.dev.type == "cuda"
                     and dist.get_backend(group) == "nccl")
            exchange = ("nvls" if self.world >= 8 else "peer") if fused else "nccl"
        self.exchange, self.block = exchange, None
        if exchange == "nvls":
            try:
                self._install_peer(steps=0)
            except RuntimeError:      # no multicast mapping on this box (the same on every rank): the unicast kernel
                self.exchange = "peer"
                self._install_peer(steps=0)
        elif exchange != "nccl":
            self._install_peer(steps=0)

    # ------------------------------------------------------------------------------------------------
    def _install_peer(self, steps: int, moments=None):
        """Move parameters and gradients into a peer-visible block (collective when world > 1): self.params become views
        of the block's parameter arena, the gradient arena IS the block's, the Adam moments are two local flat buffers
        with the same layout whose views are installed as the optimizer state."""
        from .peer import PeerBlock
        from .ops import FusedAdam
        old = self.block
        assert self.exchange == "peer" or (self.distributed and self.world > 1), "exchange=\"nvls\" needs a process group"
        block = PeerBlock(self.arena.numel, self.dev, backend="symm" if self.exchange == "nvls" else "ipc", side_floats=self.side)
        if self.distributed and self.world > 1:
            block.connect(self.group)
        arena = GradArena(list(self.params.values()), extra=self.side, flat=block.grads)
        with torch.no_grad():
            pviews = arena.views_over(block.params)
            for v, p in zip(pviews, self.params.values()):
                v.copy_(p)
        self.params = dict(zip(self.params.keys(), pviews))
        arena.params = list(self.params.values())
        self.arena, self.block = arena, block
        self.grads = dict(zip(self.params.keys(), arena.views))
        self.loss_total = arena.extra[:1]        # where mgs_map_end adds this rank's share
        self._m_flat, self._v_flat = torch.zeros_like(block.params), torch.zeros_like(block.params)
        mv, vv = arena.views_over(self._m_flat), arena.views_over(self._v_flat)
        self.opt = FusedAdam([{"params": [p], "lr": self.lrs[k]} for k, p in self.params.items()], lr=0.0, eps=1e-15,
                             capturable=False)     # peer mode: the optimiser object only carries the moments
        for i, (k, p) in enumerate(self.params.items()):
            if moments is not None:
                mv[i].copy_(moments[0][k]); vv[i].copy_(moments[1][k])
            self.opt.state[p] = dict(step=steps, exp_avg=mv[i], exp_avg_sq=vv[i])
        self._peer_step, self._epoch = steps, 0
        self._peer_x = block.exchange(self._m_flat, self._v_flat, arena.segment_ends,
                                      [self.lrs[k] for k in self.params] + [-1.0])
        if old is not None:
            old.close()

    def _args(self):
        L = self.L
        a = L.MgsMapArgs()
        g, p = a.g, self.params
        g.P, g.sh_coeffs, g.sh_degree, g.isotropic = self.P, self.M, self.sh_degree, int(self.isotropic)
        g.xyz, g.f_dc = p["xyz"].data_ptr(), p["f_dc"].data_ptr()
        g.f_rest = p["f_rest"].data_ptr() if "f_rest" in p else None
        g.opacity_raw, g.scaling_raw, g.rotation_raw = p["opacity"].data_ptr(), p["scaling"].data_ptr(), p["rotation"].data_ptr()
        gr = self.grads
        a.grad_xyz, a.grad_f_dc = gr["xyz"].data_ptr(), gr["f_dc"].data_ptr()
        a.grad_f_rest = gr["f_rest"].data_ptr() if "f_rest" in gr else None
        a.grad_opacity, a.grad_scaling, a.grad_rotation = gr["opacity"].data_ptr(), gr["scaling"].data_ptr(), gr["rotation"].data_ptr()
        a.xyz_gradient_accum, a.denom, a.max_radii2D = self.xyz_gradient_accum.data_ptr(), self.denom.data_ptr(), self.max_radii2D.data_ptr()
        a.loss_total, a.scratch, a.scratch_bytes = self.loss_total.data_ptr(), self.scratch.data_ptr(), self.scratch.numel()
        return a

    def _ensure_capacity(self, views: Sequence[MapView]):
        """First use (and after every regrow): one synchronous preprocess of EVERY view this rank renders to learn their
        num_rendered; capacity = headroom x the largest (all workspaces share one size).  Afterwards the sticky header
        words of the workspaces (overflow_sticky, max_rendered: raised by the scan kernel for ANY view rendered into that
        workspace, never cleared by it) are polled without a host wait, see poll_capacity()."""
        if self.ws is not None:
            return
        from .rasterizer import make_view, _stream
        L, lib = self.L, self.lib
        with torch.no_grad():
            op = torch.sigmoid(self.params["opacity"]).contiguous()
            sc = torch.exp(self.params["scaling"]).contiguous()
            if self.isotropic: sc = sc.repeat(1, 3).contiguous()
            rot = torch.nn.functional.normalize(self.params["rotation"]).contiguous()
            shs = (torch.cat((self.params["f_dc"], self.params["f_rest"]), dim=1).contiguous() if "f_rest" in self.params
                   else self.params["f_dc"])
        base = lib.mgs_raster_workspace_bytes(self.P, self.W, self.H, 0)
        tmp = torch.zeros(base, device=self.dev, dtype=torch.uint8)     # zero: the sticky max_tile_count is read below
        radii = torch.zeros(self.P, device=self.dev, dtype=torch.int32)
        n_host = torch.zeros(max(len(views), 1), dtype=torch.int32).pin_memory()
        for k, view in enumerate(views):
            keep = []
            v = make_view(view.settings._replace(sh_degree=self.sh_degree), self.M, keep)
            L.check(lib.mgs_raster_preprocess(C.byref(v), self.P, L.ptr(self.params["xyz"]), L.ptr(op), L.ptr(shs), None, L.ptr(sc),
                                              L.ptr(rot), None, L.ptr(radii), L.ptr(tmp), tmp.numel(),
                                              C.c_void_p(n_host[k:].data_ptr()), _stream(self.dev)), "mgs_raster_preprocess")
        L.check(lib.mgs_raster_status_async(L.ptr(tmp), C.c_void_p(self._status.data_ptr()), _stream(self.dev)), "mgs_raster_status_async")
        torch.cuda.current_stream(self.dev).synchronize()
        seen = max(int(n_host.max()), self._seen_max)
        self.R_cap = max(int(seen * self.headroom) + 4096, 4096)
        self._tile_max = max(self._tile_max, int(self._status[0, 7]))
        if self.tile_segments and self._tile_max * 1.3 > TILE_SEGMENT:
            self.tile_segments = False
        if self.tile_segments:
            self.R_cap = max(self.R_cap, ((self.W + 15) // 16) * ((self.H + 15) // 16) * TILE_SEGMENT)
        # zero-filled: the sticky header words are read by poll_capacity()
        self.ws = torch.zeros(lib.mgs_raster_workspace_bytes(self.P, self.W, self.H, self.R_cap), device=self.dev, dtype=torch.uint8)
        self.ws_extra, self._out_extra = [], []
        self._status_pending = False
        self._graph = None

    def poll_capacity(self):
        """Overflow watch without a host wait (called by step() every `poll_every` iterations): enqueue a 32-byte copy of
        every workspace header to pinned memory, and look at the copies enqueued at the previous poll (event.query(), never
        a synchronize).  Any view of any iteration in between that exceeded R_cap has raised overflow_sticky in its
        workspace: the engine counts it (`overflow_polls`), drops the workspaces and re-estimates the capacity from the
        largest num_rendered seen, before the next iteration.  A count within 10 % of the capacity regrows proactively.
        Returns True when the workspaces were dropped."""
        from .rasterizer import _stream
        dropped = False
        if self._status_pending and self._status_ev.query():
            self._status_pending = False
            n = self._status_n
            over = bool(self._status[:n, 4].any())
            self._seen_max = max(self._seen_max, int(self._status[:n, 5].max()))
            self._tile_max = max(self._tile_max, int(self._status[:n, 7].max()))
            leave = self.tile_segments and (over or self._tile_max * 1.1 > TILE_SEGMENT)
            if over:
                self.overflow_polls += 1
            if leave:
                self.tile_segments = False       # a tile list reached the segment size: general binning from here on
            if over or leave or self._seen_max * 1.1 > self.R_cap:
                self.ws, self.ws_extra = None, []
                dropped = True
        if not self._status_pending and self.ws is not None:
            wss = [self.ws] + list(self.ws_extra)
            with torch.cuda.device(self.dev):
                for k, ws in enumerate(wss):
                    self.L.check(self.lib.mgs_raster_status_async(self.L.ptr(ws), C.c_void_p(self._status[k:].data_ptr()),
                                                                  _stream(self.dev)), "mgs_raster_status_async")
                self._status_ev.record(torch.cuda.current_stream(self.dev))
            self._status_n, self._status_pending = len(wss), True
        return dropped

    def _slot(self, i):
        """Per-view output buffers (kept across iterations: the same window is mapped many times)."""
        s = self._slots.get(i)
        if s is None:
            s = self._slots[i] = MapViewResult(torch.zeros(self.P, device=self.dev, dtype=torch.int32),
                                               torch.zeros(self.P, device=self.dev, dtype=torch.int32),
                                               torch.zeros(6, device=self.dev), torch.zeros(2, device=self.dev),
                                               torch.zeros(1, device=self.dev))
        return s

    def _view_desc(self, view: MapView, slot: MapViewResult, k: int = 0):
        from .rasterizer import make_view
        L = self.L
        mv = L.MgsMapView()
        keep = []
        mv.view = make_view(view.settings._replace(sh_degree=self.sh_degree), self.M, keep)
        self._keep.append(keep)
        mv.gt_image = view.gt_image.data_ptr()
        mv.gt_depth = view.gt_depth.data_ptr() if view.gt_depth is not None else None
        mv.exposure_a = view.exposure_a.data_ptr() if view.exposure_a is not None else None
        mv.exposure_b = view.exposure_b.data_ptr() if view.exposure_b is not None else None
        mv.rgb_boundary_threshold, mv.alpha, mv.update_stats = self.thr, self.alpha, int(view.update_stats)
        r = k % self.n_pipe if self.overlap else 0      # resource set of this view: workspace, slot, output images
        self._ensure_sets(r + 1)
        ws = self.ws if r == 0 else self.ws_extra[r - 1]
        color, depth, opacity = (self.color, self.depth, self.opacity) if r == 0 else self._out_extra[r - 1]
        mv.fast_math, mv.slot, mv.concurrent_backward = int(self.fast_math), r, int(self.overlap and self._concurrent)
        mv.tile_segments = int(self.tile_segments)
        mv.out_color, mv.out_depth, mv.out_opacity = color.data_ptr(), depth.data_ptr(), opacity.data_ptr()
        mv.radii, mv.n_touched = slot.radii.data_ptr(), slot.n_touched.data_ptr()
        mv.tau, mv.dab, mv.loss = slot.tau.data_ptr(), slot.dab.data_ptr(), slot.loss.data_ptr()
        mv.workspace, mv.workspace_bytes, mv.R_cap = ws.data_ptr(), ws.numel(), self.R_cap
        return mv

    def _ensure_sets(self, n):
        """Resource sets 1 .. n-1 (set 0 is self.ws / self.color ...), allocated on first use: a one-view iteration needs
        only set 0.  The zero fill runs on the CALLER's stream, so it must be enqueued before the event the other streams
        wait for (_views_overlapped allocates before it records `begin`)."""
        while n - 1 > len(self.ws_extra):
            self.ws_extra.append(torch.zeros_like(self.ws))
            self._out_extra.append((torch.empty_like(self.color), torch.empty_like(self.depth), torch.empty_like(self.opacity)))

    def render_backward_view(self, args, view: MapView, slot: MapViewResult):
        """One view, forward and backward back to back on the current stream."""
        from .rasterizer import _stream
        mv = self._view_desc(view, slot, 0)
        self.L.check(self.lib.mgs_map_view_fwd_bwd(C.byref(args), C.byref(mv), _stream()), "mgs_map_view_fwd_bwd")

    def _views_overlapped(self, args, todo):
        """Views in flight on four streams: forward(k) alternates between the caller's stream and a second forward
        stream, backward(k) follows it on one of two backward streams (their gradient / statistics accumulation is
        atomic, mgs_map_view.concurrent_backward); forward(k + n_pipe) waits for backward(k) because it reuses that
        view's workspace, scratch slot and output images."""
        L, lib = self.L, self.lib
        main = torch.cuda.current_stream()
        if self._bwd_streams is None:
            self._bwd_streams = (torch.cuda.Stream(device=self.dev), torch.cuda.Stream(device=self.dev))
            self._fwd2_stream = torch.cuda.Stream(device=self.dev)
        fwd = (main, self._fwd2_stream)
        self._ensure_sets(min(len(todo), self.n_pipe))   # their zero fill is on `main`, ahead of the `begin` event
        begin = torch.cuda.Event()
        begin.record(main)                       # mgs_map_begin (activations, zeroed accumulators) is on the caller's stream
        fwd[1].wait_event(begin)
        self._concurrent = True
        bwd_done = []
        for k, (view, slot) in enumerate(todo):
            mv = self._view_desc(view, slot, k)
            f, b = fwd[k & 1], self._bwd_streams[k & 1]
            if k >= self.n_pipe:
                f.wait_event(bwd_done[k - self.n_pipe])
            L.check(lib.mgs_map_view_forward(C.byref(args), C.byref(mv), C.c_void_p(f.cuda_stream)), "mgs_map_view_forward")
            ev = torch.cuda.Event()
            ev.record(f)
            b.wait_event(ev)
            L.check(lib.mgs_map_view_backward(C.byref(args), C.byref(mv), C.c_void_p(b.cuda_stream)), "mgs_map_view_backward")
            ev = torch.cuda.Event()
            ev.record(b)
            bwd_done.append(ev)
        self._concurrent = False
        for ev in bwd_done[-2:]:
            main.wait_event(ev)

    # ---- the keyframe window: placement, window poses, learning-rate schedule ----------------------------------------
    def set_window(self, views: Sequence[MapView], T_w2c: Optional[Sequence[torch.Tensor]] = None,
                   frozen: Optional[Sequence[bool]] = None, n_window: Optional[int] = None,
                   pose_lrs=(5e-4, 1.5e-3, 0.01, 0.01)):
        """Install the views of the coming iterations (identical on every rank): the first n_window are the keyframe window,
        the rest the extra keyframes BackEnd.map renders beside it (utils/slam_backend.py:215-242); see assign_views().

        T_w2c (one f32[4,4] world-to-camera matrix per view) switches on the window-pose side of the iteration
        (slam_backend.py:325-332): every view gets (rho, theta, exposure a, b) + Adam moments on the device, the engine owns
        the view / projection matrices the kernels read, and after the backward of an iteration the rank that rendered a
        view takes its Adam step (lrs = cam_trans_delta, cam_rot_delta, exposure_a, exposure_b: the backend builds the
        window optimiser at half the tracking rates, :437-489), updates T <- Exp(rho, theta) T and rewrites the matrices in
        place.  frozen[i] keeps view i fixed (the first keyframe; views outside the window are never optimised).  The
        owners' poses ride in the side segment of the gradient exchange, so window_poses() is complete on every rank."""
        V = len(views)
        assert 0 < V <= self.max_views, f"at most max_views={self.max_views} views per iteration"
        n_window = V if n_window is None else min(n_window, V)
        win = dict(views=list(views), n_window=n_window, per_rank=assign_views(V, self.world, n_window), poses=None)
        win["mine"] = win["per_rank"][self.rank]
        if T_w2c is not None:
            assert len(T_w2c) == V
            f = dict(device=self.dev, dtype=torch.float32)
            st = dict(T=torch.zeros(V, 4, 4, **f), view=torch.zeros(V, 4, 4, **f), proj=torch.zeros(V, 4, 4, **f),
                      params=torch.zeros(V, 8, **f), m=torch.zeros(V, 8, **f), v=torch.zeros(V, 8, **f),
                      step=torch.zeros(V, device=self.dev, dtype=torch.int32),
                      proj_raw=views[0].settings.projmatrix_raw.to(**f).contiguous().clone(), lrs=tuple(float(x) for x in pose_lrs),
                      frozen=[bool(frozen[i]) if frozen is not None else False for i in range(V)])
            new_views = []
            for i, v in enumerate(views):
                st["T"][i].copy_(T_w2c[i])
                st["view"][i].copy_(st["T"][i].t())
                st["proj"][i].copy_(st["view"][i] @ st["proj_raw"])
                st["params"][i, 6] = 1.0
                if v.exposure_a is not None:
                    st["params"][i, 6:7].copy_(v.exposure_a.detach().reshape(1)); st["params"][i, 7:8].copy_(v.exposure_b.detach().reshape(1))
                rs = v.settings
                alias = rs.campos is not None and rs.campos.data_ptr() == rs.viewmatrix.data_ptr()
                rs = rs._replace(viewmatrix=st["view"][i], projmatrix=st["proj"][i], projmatrix_raw=st["proj_raw"],
                                 campos=st["view"][i] if (alias or rs.campos is None or rs.campos.numel() >= 16) else rs.campos)
                new_views.append(v._replace(settings=rs,
                                            exposure_a=st["params"][i, 6:7] if v.exposure_a is not None else None,
                                            exposure_b=st["params"][i, 7:8] if v.exposure_a is not None else None))
                if i >= n_window:
                    st["frozen"][i] = True
            win["views"], win["poses"] = new_views, st
        self._win, self._graph = win, None
        self.arena.extra[4:].zero_()
        return win["mine"]

    def reset_pose_optimizer(self):
        """A new keyframe_optimizers object (the backend rebuilds it for every map() call, slam_backend.py:437-489)."""
        st = self._win["poses"]
        st["m"].zero_(); st["v"].zero_(); st["step"].zero_()

    def window_poses(self):
        """(T_w2c [V,4,4], exposure [V,2]) of every view of the window.  Views this rank rendered are live; the others come
        out of the side segment of the last exchange (zeros before the first iteration of a window on a multi-rank group)."""
        st = self._win["poses"]
        T, ab = st["T"].clone(), st["params"][:, 6:8].clone()
        if self.world > 1:
            side = self.block.loss_out if self.exchange != "nccl" else self.arena.extra
            rec = side[4:4 + 20 * T.size(0)].view(-1, 20)
            for r, idxs in enumerate(self._win["per_rank"]):
                if r != self.rank:
                    for i in idxs:
                        T[i].copy_(rec[i, :16].view(4, 4)); ab[i].copy_(rec[i, 16:18])
        return T, ab

    def _pose_step(self, out):
        """keyframe_optimizers.step() + update_pose for the views this rank rendered (one launch)."""
        from .rasterizer import _stream
        st, L = self._win["poses"], self.L
        mine = self._win["mine"]
        if not mine:
            return
        arr = (L.MgsPoseSlot * len(mine))()
        for k, i in enumerate(mine):
            s = arr[k]
            s.T_w2c, s.params = st["T"][i].data_ptr(), st["params"][i].data_ptr()
            s.exp_avg, s.exp_avg_sq, s.step = st["m"][i].data_ptr(), st["v"][i].data_ptr(), st["step"][i:i + 1].data_ptr()
            s.tau, s.dab = out[i].tau.data_ptr(), out[i].dab.data_ptr()
            s.viewmatrix, s.projmatrix = st["view"][i].data_ptr(), st["proj"][i].data_ptr()
            s.gather_out = self.arena.extra[4 + 20 * i:].data_ptr()
            for q in range(4):
                s.lr[q] = st["lrs"][q]
            s.frozen = int(st["frozen"][i])
        L.check(self.lib.mgs_window_pose_step(arr, len(mine), L.ptr(st["proj_raw"]), 0.9, 0.999, 1e-8, _stream()),
                "mgs_window_pose_step")

    def set_learning_rates(self, lrs: Dict[str, float]):
        """New learning rates for some parameter groups, effective from the next step — also under CUDA-graph replay (the
        rates live on the device) and on the fused exchanges (the segment table of mgs_peer_reduce_adam is rewritten; the
        block is not re-installed)."""
        names = list(self.params.keys())
        for k, lr in lrs.items():
            if k not in self.params:
                continue
            self.lrs[k] = float(lr)
            self.opt.set_lr(names.index(k), float(lr))
            if self.exchange != "nccl":
                self._peer_x.segment_lr[names.index(k)] = float(lr)

    def update_learning_rate(self, iteration, lr_init, lr_final, lr_delay_mult=1.0, max_steps=1000000):
        """GaussianModel.update_learning_rate (gaussian_model.py:298-312): the xyz rate of this iteration; returns it."""
        lr = expon_lr(iteration, lr_init, lr_final, lr_delay_mult=lr_delay_mult, max_steps=max_steps)
        self.set_learning_rates({"xyz": lr})
        return lr

    def reset_opacity(self, value: float = 0.01, visibility_filters: Optional[Sequence[torch.Tensor]] = None):
        """GaussianModel.reset_opacity (gaussian_model.py:364-367, value 0.01) and reset_opacity_nonvisible (:369-377,
        value 0.4: Gaussians seen by any of `visibility_filters` — bool[P] or radii i32[P] per view — keep their opacity),
        with replace_tensor_to_optimizer (:470-481): the opacity moments are zeroed, the step count stays.  In a sharded
        window every rank passes the filters of ITS views; they are OR-ed over the group."""
        import math
        raw = math.log(value / (1.0 - value))          # inverse_sigmoid
        op = self.params["opacity"]
        with torch.no_grad():
            if visibility_filters is None:
                op.fill_(raw)
            else:
                seen = torch.zeros(self.P, device=self.dev, dtype=torch.int32)
                for fl in visibility_filters:
                    seen |= (fl.reshape(-1) > 0).to(torch.int32) if fl.dtype != torch.bool else fl.reshape(-1).to(torch.int32)
                if self.distributed and self.world > 1:
                    dist.all_reduce(seen, op=dist.ReduceOp.MAX, group=self.group)
                op.copy_(torch.where(seen.view_as(op) > 0, op, torch.full_like(op, raw)))
            st = self.opt.state.get(op)
            if st is not None:
                st["exp_avg"].zero_(); st["exp_avg_sq"].zero_()

    # ---- one iteration ----------------------------------------------------------------------------------------------
    def backward_only(self, views: Optional[Sequence[MapView]] = None) -> Dict[int, MapViewResult]:
        """Everything of an iteration up to the raw-parameter gradients of THIS rank's views (no exchange, no Adam).
        views=None: the window installed with set_window()."""
        from .rasterizer import _stream
        L = self.L
        if views is None:
            views, mine = self._win["views"], list(self._win["mine"])
        else:
            mine = list(shard_views(len(views), self.world, self.rank))
        for p, v in zip(self.arena.params, self.arena.views):
            p.grad = v
        self._keep = []
        args = self._args()
        self._ensure_capacity([views[i] for i in mine])
        if self._win is not None and self._win["poses"] is not None and self.exchange == "nccl" and self.world > 1:
            self.arena.extra[4:].zero_()      # the all-reduce left the sums here; only owners write their slots again
        L.check(self.lib.mgs_map_begin(C.byref(args), _stream()), "mgs_map_begin")
        out = {}
        for i in mine:
            out[i] = self._slot(i)
        if self.overlap and len(mine) > 1:
            self._views_overlapped(args, [(views[i], out[i]) for i in mine])
        else:
            for i in mine:
                self.render_backward_view(args, views[i], out[i])
        L.check(self.lib.mgs_map_end(C.byref(args), int(self.rank == 0), self.reg_weight, _stream()), "mgs_map_end")
        return out

    def _iteration(self, views):
        """backward of this rank's views, window poses, the exchange, Adam — the launch sequence of one iteration."""
        out = self.backward_only(views)
        if views is None and self._win["poses"] is not None:
            self._pose_step(out)
        if self.exchange != "nccl":
            return out          # the fused exchange carries host-side step / epoch numbers: launched by step(), never captured
        if self.distributed and self.world > 1:
            dist.all_reduce(self.arena.flat, op=dist.ReduceOp.SUM, group=self.group)   # the single exchange of the path
                                                                                     # (gradients + the side segment)
        self.opt.step()
        return out

    def step(self, views: Optional[Sequence[MapView]] = None):
        """One data-parallel mapping iteration.  Returns (total loss f32[1] on the device, {view index: MapViewResult}).
        views=None runs the window installed with set_window(); with use_graph the launch sequence of that window is
        captured once (after one eager iteration that sizes every buffer) and replayed."""
        self._since_poll += 1
        if self._since_poll >= self.poll_every:
            self._since_poll = 0
            self.poll_capacity()        # may drop the workspaces: the captured graph is then rebuilt
        if views is None and self.use_graph:
            g = self._graph
            if g is None or g["win"] is not self._win or g["ws"] is not self.ws:
                # first iteration of this window (or after a regrow): eager — it sizes the workspaces, the resource sets,
                # the per-view slots and the optimiser state the captured sequence will point at
                out = self._iteration(None)
                self._graph = dict(win=self._win, ws=self.ws, graph=None, out=out)
            else:
                if g["graph"] is None:
                    torch.cuda.synchronize(self.dev)
                    cg = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(cg):
                        g["out"] = self._iteration(None)
                    g["graph"] = cg
                g["graph"].replay()
                out = g["out"]
        else:
            out = self._iteration(views)
        if self.exchange != "nccl":
            self._launch_peer_exchange()
            return self.block.loss_out[:1], out
        return self.loss_total, out

    def _launch_peer_exchange(self):
        """Gradient reduce-scatter + Adam on this rank's slice + parameter all-gather, one kernel over peer memory."""
        from .rasterizer import _stream
        self._peer_step += 1
        self._epoch += 1
        for st in self.opt.state.values():
            st["step"] = self._peer_step
        self.L.check(self.lib.mgs_peer_reduce_adam(C.byref(self._peer_x), self._peer_step, self._epoch, 0.9, 0.999, 1e-15,
                                                   _stream()), "mgs_peer_reduce_adam")

    def sync_stats(self):
        """Side channels needed only before densify / prune: sum of the locally accumulated
        statistics, max of the radii."""
        if self.distributed and self.world > 1:
            dist.all_reduce(self.xyz_gradient_accum, op=dist.ReduceOp.SUM, group=self.group)
            dist.all_reduce(self.denom, op=dist.ReduceOp.SUM, group=self.group)
            dist.all_reduce(self.max_radii2D, op=dist.ReduceOp.MAX, group=self.group)
            if self.exchange != "nccl":      # sharded Adam moments: every rank fetches the live slices of the others
                for r in range(self.world):
                    b, e = self.block.slice_of(r)
                    src = dist.get_global_rank(self.group, r) if self.group is not None else r
                    dist.broadcast(self._m_flat[b:e], src=src, group=self.group)
                    dist.broadcast(self._v_flat[b:e], src=src, group=self.group)

    def check_capacity(self):
        """Synchronous check of the device-side overflow flag of the last rendered view(s); regrows on overflow."""
        from .rasterizer import _stream
        n_max, over = 0, False
        wss = [ws for ws in [self.ws] + list(self.ws_extra) if ws is not None]
        if wss:
            st = torch.zeros(len(wss), 8, dtype=torch.int32).pin_memory()
            for k, ws in enumerate(wss):
                self.L.check(self.lib.mgs_raster_status_async(self.L.ptr(ws), C.c_void_p(st[k:].data_ptr()), _stream()),
                             "mgs_raster_status_async")
            torch.cuda.current_stream(self.dev).synchronize()
            # sticky words: ANY view rendered into these workspaces since they were allocated, not only the last one
            n_max, over = int(st[:, 5].max()), bool(st[:, 4].any())
            self._seen_max = max(self._seen_max, n_max)
            self._tile_max = max(self._tile_max, int(st[:, 7].max()))
        if over:
            self.tile_segments = False
            self.ws, self.ws_extra = None, []
        if self.block is not None:
            epoch = self.block.error_epoch(_stream())
            if epoch:
                raise RuntimeError(f"peer exchange {epoch}: a rank of the group never arrived (device-side wait timed out); "
                                   "parameters on this rank are stale from that iteration on")
        return n_max, over

    # ------------------------------------------------------------------------------------------------
    def _adam_moments(self):
        m, v = {}, {}
        for k, p in self.params.items():
            st = self.opt.state.get(p)
            m[k] = st["exp_avg"] if st is not None else torch.zeros_like(p)
            v[k] = st["exp_avg_sq"] if st is not None else torch.zeros_like(p)
        return m, v

    def _adopt(self, out):
        """Install the arrays a densify / prune produced: parameters, Adam moments (step counts carried over, like the
        reference's replace-in-place of optimizer.state), a fresh gradient arena, zeroed statistics
        (densification_postfix, gaussian_model.py:592-594), re-sized scratch; the instance capacity is re-estimated."""
        from .ops import FusedAdam
        steps = {k: (self.opt.state[p]["step"] if p in self.opt.state else 0) for k, p in self.params.items()}
        if self.opt.capturable and self.opt.device_step() is not None:      # the live count is on the device
            steps = {k: int(self.opt.device_step()) for k in self.params}
        self.params = {k: out["params"][k] for k in self.params}
        self.P = self.params["xyz"].size(0)
        self.arena = GradArena(list(self.params.values()), extra=self.side)
        self.loss_total = self.arena.extra[:1]
        self.grads = dict(zip(self.params.keys(), self.arena.views))
        self.opt = FusedAdam([{"params": [p], "lr": self.lrs[k]} for k, p in self.params.items()], lr=0.0, eps=1e-15,
                             capturable=self.use_graph and self.exchange == "nccl")
        for k, p in self.params.items():
            if steps[k] > 0:
                self.opt.state[p] = dict(step=steps[k], exp_avg=out["exp_avg"][k], exp_avg_sq=out["exp_avg_sq"][k])
        if self.opt.capturable:
            self.opt.set_step(max(steps.values()))
        self._graph = None
        f = dict(device=self.dev, dtype=torch.float32)
        self.xyz_gradient_accum, self.denom = torch.zeros(self.P, **f), torch.zeros(self.P, **f)
        self.max_radii2D = torch.zeros(self.P, **f)
        self.scratch = torch.empty(self.lib.mgs_map_scratch_bytes(self.P, self.M, self.W, self.H), device=self.dev, dtype=torch.uint8)
        self.ws, self._slots, self._seen_max = None, {}, 0
        if self.exchange != "nccl":
            assert len(set(steps.values())) == 1, "the engine's parameters step together"
            self._install_peer(steps=next(iter(steps.values())), moments=(out["exp_avg"], out["exp_avg_sq"]))

    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, percent_dense=0.01, unique_kfIDs=None,
                          n_obs=None, noise=None, generator=None):
        """GaussianModel.densify_and_prune on the engine's arena (gaussian_model.py:676-691).  Call sync_stats() first
        in a data-parallel group, and draw from identically seeded generators on every rank."""
        from . import densify as DZ
        m, v = self._adam_moments()
        out = DZ.densify_and_prune(self.params, m, v, self.xyz_gradient_accum, self.denom, max_grad, min_opacity, extent,
                                   max_screen_size, percent_dense=percent_dense, unique_kfIDs=unique_kfIDs, n_obs=n_obs,
                                   noise=noise, generator=generator)
        self._adopt(out)
        return out

    def prune_points(self, mask, unique_kfIDs=None, n_obs=None):
        """GaussianModel.prune_points(mask) on the engine's arena (gaussian_model.py:512-527)."""
        from . import densify as DZ
        m, v = self._adam_moments()
        out = DZ.prune_points(self.params, m, v, mask, unique_kfIDs=unique_kfIDs, n_obs=n_obs)
        self._adopt(out)
        return out
