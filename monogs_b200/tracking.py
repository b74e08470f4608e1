"""Device-resident pose tracking loop on top of the C-ABI (the reference's FrontEnd.tracking inner loop,
utils/slam_frontend.py:455-630, first-order phase).

One iteration = render -> fused tracking loss (fwd+bwd) -> pose-only backward -> Adam on
(cam_trans_delta lr 1e-3, cam_rot_delta lr 3e-3, exposure_a 2e-2, exposure_b 2e-2; configs/mono/tum/base_config.yaml
`lr:` block) -> T <- Exp(tau) T.  Every stage is a call into libmonogs_b200.so on the caller's stream with
device-side counts, so the iteration has no host synchronisation and is captured once into a CUDA graph and
replayed (the reference blocks on cudaMemcpy of num_rendered every forward, rasterizer_impl.cu:284, and pays a
Python + autograd round trip per stage).  torch provides memory, streams and the graph object only.
"""
import ctypes as C
import os

import torch

from . import _lib as L
from .rasterizer import make_view, GaussianRasterizationSettings, _stream

KERNELS_PER_ITER = 7    # own kernels per fused iteration (SH degree 0): preprocess (+ pose basis, tile counts), tile_scan, emit,
                        # tile_sort_records x2 (short / long lists), blend_track (forward + loss + pose tangent),
                        # track_update (loss norm + Adam + pose update)
KERNELS_PER_ITER_SEGMENTS = 4   # fixed tile segments: preprocess (+ scatter), tile_sort_segments, blend_track, track_update
KERNELS_PER_ITER_COMPOSED = 14
TILE_SEGMENT = 1024     # MGS_TILE_SEGMENT (include/monogs_b200.h)


class PoseTracker:
    def __init__(self, means3D, opacities, shs, scales, rotations, sh_degree, cam, bg, T_w2c, mode=L.LOSS_TRACK_HUBER,
                 use_depth=False, lr=(1e-3, 3e-3, 2e-2, 2e-2), capacity_headroom=1.3, use_graph=True, fused=True, fast_math=True,
                 adam_eps=1e-8, tile_segments=True):
        """tile_segments: fixed-segment binning of the fused SH-degree-0 iteration (mgs_track_args.tile_segments: four kernels
        per iteration instead of seven; the workspace holds tiles x 1024 instance slots).  A tile list beyond 1024 instances
        raises the overflow status; poll_capacity() / check_capacity() then switch the tracker back to the general binning.
        adam_eps: the reference builds the pose / exposure optimiser as torch.optim.Adam(opt_params) with torch's default
        eps = 1e-8 (utils/slam_frontend.py:453); only the Gaussian optimiser uses 1e-15 (gaussian_model.py:285)."""
        lib = self.lib = L.lib()
        dev = self.dev = means3D.device
        f = dict(device=dev, dtype=torch.float32)
        self.P, self.H, self.W = means3D.size(0), cam.H, cam.W
        self.means3D, self.opacities, self.shs = means3D.contiguous(), opacities.contiguous(), shs.contiguous()
        self.scales, self.rotations = scales.contiguous(), rotations.contiguous()
        self.M = shs.size(1)
        # pose + matrices live in fixed device buffers that the view struct points at
        self.T = T_w2c.to(**f).contiguous().clone()
        self.proj_raw = cam.projmatrix_raw.to(**f).contiguous()
        self.viewmatrix = self.T.t().contiguous()
        self.projmatrix = (self.viewmatrix @ self.proj_raw).contiguous()
        self.bg = bg.to(**f).contiguous()
        rs = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, self.bg, 1.0, self.viewmatrix,
                                           self.projmatrix, self.proj_raw, sh_degree, self.viewmatrix, False, False)
        self._keep = []
        self.view = make_view(rs, self.M, self._keep)
        assert self.view.viewmatrix == self.viewmatrix.data_ptr() and self.view.campos == self.viewmatrix.data_ptr()
        # optimiser state: params = [rho(3), theta(3), a(1), b(1)], grads = [tau(6), dab(2)]
        self.params = torch.zeros(8, **f)
        self.params[6] = 1.0            # exposure_a starts at 1 (utils/camera_utils.py:56-58)
        self.tau = torch.zeros(6, **f)
        self.dab = torch.zeros(2, **f)
        self.exp_avg, self.exp_avg_sq = torch.zeros(8, **f), torch.zeros(8, **f)
        self.lr, self.adam_eps = lr, float(adam_eps)
        self.step_dev = torch.zeros(1, device=dev, dtype=torch.int32)   # Adam step number, kept on the device
        self.mode, self.use_depth = mode, use_depth
        # outputs / intermediates
        self.color, self.depth = torch.empty(3, self.H, self.W, **f), torch.empty(1, self.H, self.W, **f)
        self.opacity = torch.empty(1, self.H, self.W, **f)
        self.radii = torch.zeros(self.P, device=dev, dtype=torch.int32)
        self.n_touched = torch.zeros(self.P, device=dev, dtype=torch.int32)
        self.g_color, self.g_depth = torch.zeros(3, self.H, self.W, **f), torch.zeros(1, self.H, self.W, **f)
        self.loss = torch.zeros(1, **f)
        self.converged = torch.zeros(1, device=dev, dtype=torch.int32)
        self.gt_image = torch.zeros(3, self.H, self.W, **f)
        self.gt_depth = torch.zeros(1, self.H, self.W, **f) if use_depth else None
        self.grad_mask = None
        self.loss_scratch = torch.empty(lib.mgs_loss_scratch_bytes(self.H, self.W), device=dev, dtype=torch.uint8)
        self.bwd_scratch = torch.empty(max(lib.mgs_raster_backward_pose_scratch_bytes(self.P, self.W, self.H), 256),
                                       device=dev, dtype=torch.uint8)
        self.fused, self.fast_math = fused, fast_math
        self.tile_segments = bool(tile_segments) and fused and sh_degree == 0 and os.environ.get("MGS_TILE_SEGMENTS", "1") != "0"
        self.tiles = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        # zero-filled once: holds the tangent accumulator and the update ticket, which every iteration leaves zero again
        self.track_scratch = torch.zeros(lib.mgs_track_scratch_bytes(self.P, self.W, self.H), device=dev, dtype=torch.uint8)
        self.headroom = capacity_headroom
        self.R_cap, self.ws = 0, None
        self.use_graph, self.graph = use_graph, None
        self._seen_max, self._tile_max = 0, 0
        self._status = torch.zeros(8, dtype=torch.int32).pin_memory()
        self._status_ev = torch.cuda.Event()
        self._status_pending = False
        # cross-iteration record of the tracker on the device (include/monogs_b200.h, lm_state): LM damping, pending
        # second-order step, best pose by scalar loss
        self.lm_state = torch.zeros(64, **f)
        self._sk = None             # second-order sketch buffers (set_sketch)
        self._graph2 = None
        self._host_bind = None
        self.overflow_frames = 0     # frames during which at least one iteration dropped instances (see poll_capacity)
        self._size_workspace()

    # ------------------------------------------------------------------------------------------------
    def _size_workspace(self):
        """One synchronous preprocess to learn num_rendered, then capacity = headroom x that."""
        lib = self.lib
        base = lib.mgs_raster_workspace_bytes(self.P, self.W, self.H, 0)
        tmp = torch.zeros(base, device=self.dev, dtype=torch.uint8)     # zero: the sticky header words are read below
        n_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        L.check(lib.mgs_raster_preprocess(C.byref(self.view), self.P, L.ptr(self.means3D), L.ptr(self.opacities),
                                          L.ptr(self.shs), None, L.ptr(self.scales), L.ptr(self.rotations), None,
                                          L.ptr(self.radii), L.ptr(tmp), tmp.numel(), C.c_void_p(n_host.data_ptr()),
                                          _stream(self.dev)), "mgs_raster_preprocess")
        L.check(lib.mgs_raster_status_async(L.ptr(tmp), C.c_void_p(self._status.data_ptr()), _stream(self.dev)), "mgs_raster_status_async")
        torch.cuda.current_stream(self.dev).synchronize()
        R = int(n_host.item())
        self.R_cap = max(int(max(R, self._seen_max) * self.headroom) + 4096, 4096)
        self._tile_max = max(self._tile_max, int(self._status[7]))
        if self.tile_segments and self._tile_max * self.headroom > TILE_SEGMENT:
            self.tile_segments = False      # a tile list is (close to) longer than a segment: general binning
        if self.tile_segments:
            self.R_cap = max(self.R_cap, self.tiles * TILE_SEGMENT)     # every tile owns a fixed segment
        # zero-filled: the sticky header words (overflow_sticky, max_rendered) are read by poll_capacity()
        self.ws = torch.zeros(lib.mgs_raster_workspace_bytes(self.P, self.W, self.H, self.R_cap), device=self.dev,
                              dtype=torch.uint8)
        self.graph = None
        self._status_pending = False
        if getattr(self, "_host_bind", None) is not None:
            self.bind_host(*self._host_bind)     # the captured host graphs point into the old workspace

    def _loss_args(self):
        a = L.MgsLossArgs()
        a.mode, a.H, a.W, a.use_depth = self.mode, self.H, self.W, int(self.use_depth)
        a.rgb_boundary_threshold, a.alpha, a.huber_delta, a.loss_scale = 0.01, 0.95, 0.01, 1.0
        a.image, a.depth, a.opacity = self.color.data_ptr(), self.depth.data_ptr(), self.opacity.data_ptr()
        a.gt_image = self.gt_image.data_ptr()
        a.gt_depth = self.gt_depth.data_ptr() if self.gt_depth is not None else None
        a.grad_mask = self.grad_mask.data_ptr() if self.grad_mask is not None else None
        a.exposure_a, a.exposure_b = self.params[6:7].data_ptr(), self.params[7:8].data_ptr()
        return a

    def _track_args(self):
        t = L.MgsTrackArgs()
        t.view = self.view
        t.P = self.P
        t.means3D, t.opacities, t.shs = self.means3D.data_ptr(), self.opacities.data_ptr(), self.shs.data_ptr()
        t.scales, t.rotations = self.scales.data_ptr(), self.rotations.data_ptr()
        t.loss_mode, t.use_depth, t.fast_math = self.mode, int(self.use_depth), int(self.fast_math)
        t.rgb_boundary_threshold, t.alpha, t.huber_delta = 0.01, 0.95, 0.01
        t.gt_image = self.gt_image.data_ptr()
        t.gt_depth = self.gt_depth.data_ptr() if self.gt_depth is not None else None
        t.grad_mask = self.grad_mask.data_ptr() if self.grad_mask is not None else None
        t.T_w2c, t.params = self.T.data_ptr(), self.params.data_ptr()
        t.exp_avg, t.exp_avg_sq, t.step_counter = self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.step_dev.data_ptr()
        for i in range(4):
            t.lr[i] = self.lr[i]
        t.beta1, t.beta2, t.adam_eps, t.converged_threshold = 0.9, 0.999, self.adam_eps, 1e-4
        t.out_color, t.out_depth, t.out_opacity = self.color.data_ptr(), self.depth.data_ptr(), self.opacity.data_ptr()
        t.radii, t.n_touched = self.radii.data_ptr(), self.n_touched.data_ptr()
        t.loss, t.tau, t.dab, t.converged = self.loss.data_ptr(), self.tau.data_ptr(), self.dab.data_ptr(), self.converged.data_ptr()
        t.workspace, t.workspace_bytes, t.R_cap = self.ws.data_ptr(), self.ws.numel(), self.R_cap
        t.scratch, t.scratch_bytes = self.track_scratch.data_ptr(), self.track_scratch.numel()
        t.lm_state, t.status_out = self.lm_state.data_ptr(), None
        t.tile_segments = int(self.tile_segments)
        return t

    @property
    def kernels_per_iter(self):
        if not self.fused:
            return KERNELS_PER_ITER_COMPOSED
        if self.tile_segments:
            return KERNELS_PER_ITER_SEGMENTS
        # SH degree > 0: the colours depend on the pose, so the blend is two kernels (forward + loss, pose-only backward)
        return KERNELS_PER_ITER + (1 if self.view.sh_degree > 0 else 0)

    def _enqueue(self):
        """The whole iteration on the current stream, no host sync anywhere: one fused C-ABI call
        (mgs_track_iteration) or, with fused=False, the same iteration composed from the individual entry points."""
        if self.fused:
            t = self._track_args()
            L.check(self.lib.mgs_track_iteration(C.byref(t), _stream(self.dev)), "mgs_track_iteration")
            return
        lib, s, v = self.lib, _stream(self.dev), C.byref(self.view)
        L.check(lib.mgs_raster_preprocess(v, self.P, L.ptr(self.means3D), L.ptr(self.opacities), L.ptr(self.shs), None,
                                          L.ptr(self.scales), L.ptr(self.rotations), None, L.ptr(self.radii),
                                          L.ptr(self.ws), self.ws.numel(), None, s), "mgs_raster_preprocess")
        L.check(lib.mgs_raster_render(v, self.P, None, L.ptr(self.radii), L.ptr(self.ws), self.ws.numel(), self.R_cap,
                                      L.ptr(self.color), L.ptr(self.depth), L.ptr(self.opacity), L.ptr(self.n_touched),
                                      s), "mgs_raster_render")
        la = self._loss_args()
        L.check(lib.mgs_loss_fwd_bwd(C.byref(la), L.ptr(self.loss), L.ptr(self.dab), L.ptr(self.g_color),
                                     L.ptr(self.g_depth), L.ptr(self.loss_scratch), self.loss_scratch.numel(), s),
                "mgs_loss_fwd_bwd")
        L.check(lib.mgs_raster_backward_pose(v, self.P, L.ptr(self.means3D), L.ptr(self.shs), None, L.ptr(self.scales),
                                             L.ptr(self.rotations), None, L.ptr(self.radii), L.ptr(self.ws),
                                             self.ws.numel(), L.ptr(self.bwd_scratch), self.bwd_scratch.numel(),
                                             L.ptr(self.g_color), L.ptr(self.g_depth), None, L.ptr(self.tau), s),
                "mgs_raster_backward_pose")
        groups = (L.MgsAdamGroup * 4)()
        spans = ((0, 3, self.tau, 0), (3, 6, self.tau, 3), (6, 7, self.dab, 0), (7, 8, self.dab, 1))
        for i, (lo, hi, gsrc, goff) in enumerate(spans):
            groups[i].param = self.params[lo:hi].data_ptr()
            groups[i].grad = gsrc[goff:goff + (hi - lo)].data_ptr()
            groups[i].exp_avg, groups[i].exp_avg_sq = self.exp_avg[lo:hi].data_ptr(), self.exp_avg_sq[lo:hi].data_ptr()
            groups[i].numel, groups[i].lr = hi - lo, self.lr[i]
        L.check(lib.mgs_adam_step_device(groups, 4, L.ptr(self.step_dev), 0.9, 0.999, self.adam_eps, s), "mgs_adam_step_device")
        L.check(lib.mgs_pose_update(L.ptr(self.T), L.ptr(self.params[0:3]), L.ptr(self.params[3:6]), L.ptr(self.proj_raw),
                                    L.ptr(self.viewmatrix), L.ptr(self.projmatrix), L.ptr(self.converged), 1e-4, s),
                "mgs_pose_update")

    # ------------------------------------------------------------------------------------------------
    def poll_capacity(self):
        """Overflow watch without a host wait.  The scan kernel raises two sticky words in the workspace header
        (overflow_sticky, max_rendered) whenever an iteration's tile instances exceed R_cap; once per frame the tracker
        enqueues a 32-byte copy of the header to pinned memory and looks at the copy it enqueued one frame earlier (long
        complete: event.query(), never synchronize).  An overflow regrows the workspace (and re-captures the graph) before
        the next frame and is counted in `overflow_frames`; a count that merely approaches the capacity regrows it
        proactively.  Returns True when the workspace was regrown."""
        regrown = False
        if self._status_pending and self._status_ev.query():
            self._status_pending = False
            over, seen = bool(self._status[4]), int(self._status[5])
            self._seen_max = max(self._seen_max, seen)
            self._tile_max = max(self._tile_max, int(self._status[7]))
            leave = self.tile_segments and (over or self._tile_max * 1.1 > TILE_SEGMENT)
            if over:
                self.overflow_frames += 1
            if leave:
                self.tile_segments = False      # in segment mode overflow = a tile list beyond 1024 instances
            if over or leave or seen * 1.1 > self.R_cap:
                self._size_workspace()
                regrown = True
        if not self._status_pending:
            with torch.cuda.device(self.dev):
                L.check(self.lib.mgs_raster_status_async(L.ptr(self.ws), C.c_void_p(self._status.data_ptr()), _stream(self.dev)),
                        "mgs_raster_status_async")
                self._status_ev.record(torch.cuda.current_stream(self.dev))
            self._status_pending = True
        return regrown

    def set_frame(self, gt_image, gt_depth=None, grad_mask=None, non_blocking=True, exposure=(1.0, 0.0)):
        """New target frame (host pinned or device tensors); resets the per-frame optimiser state like the
        reference, which builds a fresh Adam per frame (slam_frontend.py:364-392,453) and a fresh Camera whose exposure
        starts at a = 1, b = 0 (utils/camera_utils.py:55-61: image_ab = (|a| + 1e-8) * image + b); only the pose is
        carried over.  exposure=None keeps the previous frame's (a, b)."""
        self.poll_capacity()
        self.gt_image.copy_(gt_image, non_blocking=non_blocking)
        if self.gt_depth is not None and gt_depth is not None:
            self.gt_depth.copy_(gt_depth, non_blocking=non_blocking)
        if grad_mask is not None:
            if self.grad_mask is None:
                self.grad_mask = torch.zeros(1, self.H, self.W, device=self.dev)
                self.graph = None
            self.grad_mask.copy_(grad_mask, non_blocking=non_blocking)
        self.exp_avg.zero_(); self.exp_avg_sq.zero_(); self.params[:6].zero_(); self.step_dev.zero_()
        self.lm_state.zero_()
        if exposure is not None:
            self.params[6].fill_(float(exposure[0])); self.params[7].fill_(float(exposure[1]))

    def step(self):
        """One tracking iteration: the first call captures the C-ABI call sequence into a CUDA graph (the Adam step
        number lives on the device, so one graph serves every iteration); later calls replay it."""
        if not self.use_graph:
            self._enqueue()
            return
        if self.graph is None:
            torch.cuda.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._enqueue()
        self.graph.replay()

    # ---- second-order (sketched Gauss-Newton / LM) phase: utils/slam_frontend.py:455-720, in_second_order ----------------
    LM_FIELDS = dict(lam=0, prev_scalar=1, have_prev=2, pending=3, x=slice(4, 12), best_scalar=12, best_T=slice(13, 29),
                     best_a=29, best_b=30, converged=31, iters=32, scalar=33, step_norm=34, have_best=35)

    @staticmethod
    def gen_sketch_args(H, W, stack_dim=16, sketch_dim=64, device="cuda", generator=None):
        """One repeat of FrontEnd.gen_forward_sketch_args (utils/slam_frontend.py:269-338): a random permutation of the
        pixels cut into stack_dim x sketch_dim buckets of m // (stack_dim*sketch_dim) pixels -> sketch_indices
        i32[stack,H,W] (-1 = pixel not in that stack), and Rademacher rand_weights f32[H,W]."""
        m, d = H * W, stack_dim * sketch_dim
        chunk = m // d
        perm = torch.randperm(m, device=device, generator=generator)[:chunk * d]
        idx = torch.full((stack_dim * m,), -1, dtype=torch.int32, device=device)
        bucket = torch.arange(d, device=device).repeat_interleave(chunk)          # bucket of the k-th drawn pixel
        idx[(bucket // sketch_dim) * m + perm] = (bucket % sketch_dim).to(torch.int32)
        w = torch.randint(0, 2, (H, W), device=device, generator=generator, dtype=torch.float32) * 2 - 1
        return idx.view(stack_dim, H, W), w

    def begin_second_order(self, initial_lambda=1e-3, use_first_order_best=True):
        """Switch of FrontEnd.tracking to the second-order phase (:468-473): optionally go back to the best first-order
        pose / exposure (best_viewpoint_params.assign), reset the LM record, lambda = initial_lambda."""
        st = self.lm_state
        if use_first_order_best:
            have = st[35:36]                     # device-side select, no host read
            Tb = st[13:29].view(4, 4)
            self.T.copy_(torch.where(have.bool(), Tb, self.T))
            self.params[6:7].copy_(torch.where(have.bool(), st[29:30], self.params[6:7]))
            self.params[7:8].copy_(torch.where(have.bool(), st[30:31], self.params[7:8]))
            self.params[:6].zero_()
            self.viewmatrix.copy_(self.T.t())
            self.projmatrix.copy_(self.viewmatrix @ self.proj_raw)
        st[0] = float(initial_lambda)
        st[1:12].zero_(); st[31:35].zero_()

    def set_sketch(self, sketch_indices, rand_weights, sketch_dim):
        """This iteration's sketch (device tensors; copied into fixed buffers so that the captured graph stays valid)."""
        idx = sketch_indices.to(torch.int32)
        if self._sk is None or self._sk[0].shape != idx.shape or self._sk[3] != sketch_dim:
            n_stacks = idx.shape[0]
            self._sk = (torch.empty_like(idx).contiguous(), torch.empty(self.H, self.W, device=self.dev),
                        torch.zeros(n_stacks * sketch_dim, 9, device=self.dev), sketch_dim)
            self._graph2 = None
        self._sk[0].copy_(idx, non_blocking=True)
        self._sk[1].copy_(rand_weights, non_blocking=True)

    def _enqueue_second_order(self, min_lambda, max_lambda, increase_factor, decrease_factor, converged_threshold):
        a = L.MgsTrack2Args()
        a.t = self._track_args()
        idx, w, rows, d = self._sk
        a.n_stacks, a.sketch_dim = idx.shape[0], d
        a.sketch_indices, a.rand_weights, a.rows = idx.data_ptr(), w.data_ptr(), rows.data_ptr()
        a.lm_state = self.lm_state.data_ptr()
        a.min_lambda, a.max_lambda, a.increase_factor, a.decrease_factor = min_lambda, max_lambda, increase_factor, decrease_factor
        a.converged_threshold = converged_threshold
        L.check(self.lib.mgs_track_second_order(C.byref(a), _stream(self.dev)), "mgs_track_second_order")

    def second_order_step(self, min_lambda=1e-6, max_lambda=1e7, increase_factor=5.0, decrease_factor=5.0,
                          converged_threshold=1e-5):
        """One second-order iteration with the sketch given by set_sketch(): apply the pending step, render, scalar loss,
        lambda, SJ / Sf, damped solve, new pending step (defaults: configs/mono/tum/base_config.yaml RGN.second_order).
        Captured once into a CUDA graph and replayed.  Results on the device: self.sketch_rows [stack*sketch, 9] (SJ | Sf),
        self.lm_state (see LM_FIELDS), self.loss (the scalar ||r||_1)."""
        assert self._sk is not None, "set_sketch() first"
        args = (min_lambda, max_lambda, increase_factor, decrease_factor, converged_threshold)
        if not self.use_graph:
            self._enqueue_second_order(*args)
            return
        if self._graph2 is None or self._graph2[1] != args or self._graph2[2] is not self.ws:
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._enqueue_second_order(*args)
            self._graph2 = (g, args, self.ws)
        self._graph2[0].replay()

    @property
    def sketch_rows(self):
        return self._sk[2]

    def finish_second_order(self, use_best_loss=True):
        """End of the frame (slam_frontend.py:741-790): the last pending step is NOT applied by the reference's loop; with
        use_best_loss the pose / exposure of the best scalar loss seen in either phase are restored."""
        if use_best_loss:
            st = self.lm_state
            have = st[35:36].bool()
            self.T.copy_(torch.where(have, st[13:29].view(4, 4), self.T))
            self.params[6:7].copy_(torch.where(have, st[29:30], self.params[6:7]))
            self.params[7:8].copy_(torch.where(have, st[30:31], self.params[7:8]))
            self.viewmatrix.copy_(self.T.t())
            self.projmatrix.copy_(self.viewmatrix @ self.proj_raw)
        self.lm_state[3] = 0.0

    def bind_host(self, gt_host, loss_host, tau_host, gt_depth_host=None):
        """Host-resident frame in, loss + pose tangent out, one graph launch per iteration (step_host()).
        gt_host [3,H,W], loss_host [S] and tau_host [S,6] (or [6] when S = 1) are pinned host tensors.  The captured graph is
            gt_image <- staging;  { the iteration's kernels  ||  staging <- gt_host on a second stream };
            loss_host[s] <- loss, tau_host[s] <- tau
        so every iteration moves one frame host -> device and its results device -> host, with the upload of the
        NEXT iteration's frame hidden behind the current iteration's kernels: the contents of gt_host at replay k
        are what iteration k+1 tracks against (the frame stays the same for the ~100 iterations of a tracking loop,
        slam_frontend.py:455); bind_host() itself uploads the first frame.
        S = 2 result slots enable step_host_lagged(): iteration k is launched before the host waits for iteration
        k-1, so the GPU never idles while the host turns around (one graph per slot, used alternately)."""
        assert gt_host.is_pinned() and loss_host.is_pinned() and tau_host.is_pinned()
        assert (gt_depth_host is not None) == (self.gt_depth is not None), "a depth tracker takes a pinned depth frame too"
        self._host_bind = (gt_host, loss_host, tau_host, gt_depth_host)
        S = loss_host.numel()
        tau_slots = tau_host.view(S, 6)
        self._stage = torch.empty_like(self.gt_image)
        self._stage.copy_(gt_host, non_blocking=True)
        self._stage_depth = None
        if gt_depth_host is not None:
            assert gt_depth_host.is_pinned()
            self._stage_depth = torch.empty_like(self.gt_depth)
            self._stage_depth.copy_(gt_depth_host, non_blocking=True)
        torch.cuda.synchronize()
        side = torch.cuda.Stream(device=self.dev)
        self.host_graphs, self._host_ev, self._host_k = [], [torch.cuda.Event() for _ in range(S)], 0
        for slot in range(S):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                main = torch.cuda.current_stream(self.dev)
                self.gt_image.copy_(self._stage, non_blocking=True)
                if self._stage_depth is not None:
                    self.gt_depth.copy_(self._stage_depth, non_blocking=True)
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    self._stage.copy_(gt_host, non_blocking=True)
                    if self._stage_depth is not None:
                        self._stage_depth.copy_(gt_depth_host, non_blocking=True)
                self._enqueue()
                loss_host[slot:slot + 1].copy_(self.loss, non_blocking=True)
                tau_slots[slot].copy_(self.tau, non_blocking=True)
                main.wait_stream(side)
            self.host_graphs.append(g)
        self.host_graph = self.host_graphs[0]

    def step_host(self, sync=True):
        """One iteration of the host-bound pipeline set up by bind_host(); results in slot 0."""
        self.host_graph.replay()
        if sync:
            torch.cuda.current_stream(self.dev).synchronize()

    def step_host_lagged(self):
        """Launch iteration k, then wait for iteration k-1 only.  Returns the result slot (0 / 1) that now holds
        iteration k-1's loss and tangent, or None for the first call.  Needs two result slots (bind_host)."""
        S = len(self.host_graphs)
        assert S >= 2, "bind_host() with loss_host[2], tau_host[2, 6]"
        k = self._host_k
        self.host_graphs[k % S].replay()
        self._host_ev[k % S].record()
        self._host_k = k + 1
        if k == 0:
            return None
        self._host_ev[(k - 1) % S].synchronize()
        return (k - 1) % S

    def check_capacity(self):
        """Synchronous check of the device-side overflow flag; regrows the workspace if the splat count outgrew it."""
        n, ov = C.c_int32(0), C.c_int32(0)
        L.check(self.lib.mgs_raster_query(L.ptr(self.ws), C.byref(n), C.byref(ov), _stream(self.dev)), "mgs_raster_query")
        self._seen_max = max(self._seen_max, n.value)
        if ov.value:
            self.tile_segments = False
            self._size_workspace()
        return n.value, bool(ov.value)
