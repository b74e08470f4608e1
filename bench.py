#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json): tracking forward+backward iterations/s at
640x480 with 200k Gaussians (config C2: mono tracking loop, synthetic frames), one B200.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Own arm.  A step = one tracking iteration of monogs_b200.tracking.PoseTracker (render -> fused Huber/L2 tracking
loss -> pose-only backward -> Adam(pose, exposure) -> pose update), replayed from a CUDA graph.
  value : iterations/s with every input resident in HBM; per-step CUDA events, L2 flushed between steps.
  e2e   : the same iteration driven per step from HOST buffers through the C-ABI tracker: H2D of the target frame
          from pinned memory and D2H of loss + pose tangent inside the timed region.
  N > 1 : tracking does not shard ("replicas only") — every rank tracks its own frame and value is
          the sum; the sharded keyframe-window mapping step (the one place with a real exchange: one NCCL
          allreduce of the Gaussian-gradient arena) is measured at N ranks and reported under "mapping".
Reference arm (--impl reference): the reference has no CPU rasterizer, so its CPU implementation of the path is
the PyTorch-CPU golden renderer (oracle/golden.py) on the host cores, on a bounded tile sample of the same config.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "tracking fwd+bwd iters/sec @640x480 200k Gaussians"
WORKLOADS = {
    "C1": "C1: mono tracking loop on the 50k-Gaussian scene of the rasterizer fwd+bwd config, 640x480, SH degree 0",
    "C2": "C2: mono TUM fr3_office-shaped tracking loop, synthetic frames, 200k Gaussians, 640x480, SH degree 0",
    "C3": "C3: RGB-D Replica office0-shaped tracking loop with depth loss (alpha 0.95), 500k Gaussians, 1200x680, SH degree 0",
    "C5": "C5: stress, 2M Gaussians at 1920x1080, SH degree 3: tracking loop (forward + fused loss + pose-only backward)",
}
WORKLOAD = WORKLOADS["C2"]
ISSUE_PEAK_WARP_INST_PER_S = 148 * 4 * 1.965e9      # one warp instruction per SM sub-partition per clock at sm_max_mhz


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2")
    ap.add_argument("--no-extras", action="store_true", help="skip cpu_baseline / ref_cuda / mapping / profile legs")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:                       # make sure the poller is gone before anything else is timed (it queries the driver every 200 ms)
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def build_tracker(cfg, dev, use_graph=True, rank=0):
    from monogs_b200 import scene as S
    from monogs_b200.tracking import PoseTracker
    from monogs_b200 import rasterizer as RZ
    sc, cam = S.make_config(cfg, seed=rank)
    scd = S.Scene(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in sc])
    bg = torch.zeros(3, device=dev)
    # target frame: the same scene rendered at the pose perturbed by GT_TAU
    T_gt = S.se3_exp(torch.tensor(S.GT_TAU)).float()
    cam_gt = S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, T_gt)
    cam_gt = S.Camera(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in cam_gt])
    rs = RZ.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg, 1.0, cam_gt.viewmatrix,
                                          cam_gt.projmatrix, cam_gt.projmatrix_raw, sc.sh_degree, cam_gt.campos, False, False)
    gt, _, gt_depth, _, _, _ = RZ.raster_forward(rs, scd.means3D, scd.opacities, scd.shs, None, scd.scales, scd.rotations, None)
    rgbd = cfg == "C3"       # config 3 tracks with the depth loss (get_loss_tracking_rgbd), the others with the mono Huber objective
    from monogs_b200 import _lib as L
    tr = PoseTracker(scd.means3D, scd.opacities, scd.shs, scd.scales, scd.rotations, sc.sh_degree, cam, bg,
                     torch.eye(4), use_graph=use_graph, mode=L.LOSS_TRACK_L1 if rgbd else L.LOSS_TRACK_HUBER, use_depth=rgbd)
    tr.bench_gt_depth = gt_depth if rgbd else None
    tr.set_frame(gt, gt_depth=tr.bench_gt_depth)
    return tr, scd, cam, gt


def timed_steps(step_fn, steps, warmup, flush, barrier):
    """W untimed + K timed steps, one CUDA-event pair per step with the L2 flush outside the pairs."""
    for _ in range(warmup):
        flush.zero_()
        step_fn()
    torch.cuda.synchronize()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        flush.zero_()
        a.record()
        step_fn()
        b.record()
    torch.cuda.synchronize()
    barrier()
    return sum(a.elapsed_time(b) for a, b in ev)   # ms for K steps


def kernel_profile(cfg, dev, steps):
    """Per-kernel CUDA-event times over `steps` un-graphed iterations (library-side events on the launch stream)."""
    import ctypes as C
    from monogs_b200 import _lib as L
    tr, _, _, _ = build_tracker(cfg, dev, use_graph=False)
    for _ in range(5):
        tr.step()
    torch.cuda.synchronize()
    lib = L.lib()
    lib.mgs_profile_enable(1)
    for _ in range(steps):
        tr.step()
    buf = C.create_string_buffer(8192)
    lib.mgs_profile_report(buf, 8192)
    lib.mgs_profile_enable(0)
    prof = {}
    for line in buf.value.decode().strip().splitlines():
        name, n, ms = line.split()
        prof[name] = {"launches": int(n), "avg_us": 1e3 * float(ms) / int(n)}
    n_rendered, _ = tr.check_capacity()
    p_vis = int((tr.radii > 0).sum())
    return prof, n_rendered, p_vis, tr.R_cap


PROFILE_FILE = "profiles/r02_kernel_metrics.txt"     # ncu per-kernel metric table of the committed code (scripts/summarise_profiles.py)


def profile_row(kernel_prefix):
    """Row of the committed ncu metric table for the kernel whose name starts with `kernel_prefix` (tracking section), as a
    dict {column: float}, plus whether the capture still describes the code that is running: the table's header records
    the sha1 of every csrc file at capture time; a kernel source that changed since makes the row STALE and the caller
    reports no traffic rather than an old number."""
    import hashlib
    import re
    path = os.path.join(ROOT, PROFILE_FILE)
    if not os.path.exists(path):
        return None, "missing " + PROFILE_FILE
    lines = open(path).read().splitlines()
    shas = {}
    for ln in lines:
        if ln.startswith("# source-sha1:"):
            for tok in ln.split(":", 1)[1].split():
                f, _, h = tok.partition("=")
                shas[f] = h
    # only the translation unit that defines the kernel and the headers it includes (transitively) decide staleness
    csrc = os.path.join(ROOT, "monogs_b200", "csrc")
    relevant, todo = set(), [f for f in os.listdir(csrc) if f.endswith(".cu") and (kernel_prefix + "(") in open(os.path.join(csrc, f)).read()
                             or f.endswith(".cu") and ("\n" + kernel_prefix) in open(os.path.join(csrc, f)).read()]
    while todo:
        f = todo.pop()
        if f in relevant or not os.path.exists(os.path.join(csrc, f)):
            continue
        relevant.add(f)
        todo += re.findall(r'#include "([\w.]+)"', open(os.path.join(csrc, f)).read())
    stale = []
    for f, h in shas.items():
        if relevant and f not in relevant:
            continue
        fp = os.path.join(csrc, f)
        if not os.path.exists(fp) or hashlib.sha1(open(fp, "rb").read()).hexdigest()[:12] != h:
            stale.append(f)
    header, row = None, None
    for ln in lines:
        if ln.startswith("kernel ") and header is None:
            header = ln.split()
        elif header is not None and ln.startswith(kernel_prefix):
            m = re.match(r"^(.*?)\s+(\d+)((?:\s+[-0-9.eE+]+){%d})\s*$" % (len(header) - 2), ln)
            if m:
                row = dict(zip(header[1:], [float(m.group(2))] + [float(x) for x in m.group(3).split()]))
                row["kernel"] = m.group(1).strip()
                break
        elif ln.startswith("## mapping"):
            break
    if row is None:
        return None, f"{kernel_prefix} not in {PROFILE_FILE}"
    if not shas:
        return row, "capture does not record source hashes"
    return row, ("stale: changed since the capture: " + ", ".join(stale)) if stale else None


def algorithmic_bytes(P, P_vis, R, R_cap, W, H, M=1, segments=False):
    """Per-launch algorithmic bytes of every kernel of the fused tracking iteration."""
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    if segments:
        # fixed tile segments (DESIGN.md redesign 23): the preprocess kernel scatters one 8-byte word per instance (no depth /
        # mean2D arrays), one kernel sorts the segments (words in, ids out, counts in, ranges + order out)
        return {
            "preprocess": 12 * P + 8 * P + P_vis * (32 + 12 * M) + P_vis * 176 + 8 * R,
            "tile_sort_segments": R * (8 + 4) + (4 + 8 + 4) * tiles,
            "blend_track": R * (4 + 176) + (20 + 12) * W * H + 44 * tiles,
            "track_update": 32 * tiles,
        }
    return {
        # fused tracker, packed-row mode (DESIGN.md redesign 22): per visible Gaussian depth + mean2D (12 B) and one 176-B row
        "preprocess": 12 * P + 8 * P + P_vis * (32 + 12 * M) + P_vis * (4 + 8) + P_vis * 176,
        "preprocess+scan": 12 * P + 8 * P + P_vis * (32 + 12 * M) + P_vis * (4 + 8 + 16 + 16) + P_vis * 128 + 16 * tiles,
        "tile_scan": 16 * tiles,
        "emit": 4 * P + P_vis * 12 + 8 * R,
        "tile_sort_records": R * (8 + 4) + 8 * tiles,          # 64-bit words in, sorted ids out
        "tile_sort_records_large": 8 * tiles,
        "blend_track": R * (4 + 176) + (20 + 12) * W * H + 40 * tiles,   # id + gathered row per instance
        "blend_forward": R * 48 + (28 + 12 + 12) * W * H + 40 * tiles,
        "blend_backward_pose": R * (48 + 144) + 20 * W * H + 8 * tiles,
        "track_update": 32 * tiles,
    }


_CPU_STATE = {}


def cpu_oracle_rate(cfg, tile_stride=1, gauss_stride=1, stage_scale=None):
    """CPU oracle (oracle/golden.py) on the host cores, rasterizer forward+backward of the config.  A bounded sample:
    the alpha blend (fwd+bwd) runs on every `tile_stride`-th tile with its real tile lists, the per-Gaussian stages
    (preprocess, per-Gaussian backward) on every `gauss_stride`-th Gaussian.  Each stage time is scaled to the whole
    workload by `stage_scale` (four factors: full-pass time / sample time of that stage, measured by the caller) —
    the oracle's vectorised stages have fixed costs, so the inverse work fraction (the default when stage_scale is
    None) over-estimates the full pass for small samples.  Scene and tile lists are built once, untimed."""
    from monogs_b200 import scene as S
    from oracle import golden as G
    torch.set_num_threads(os.cpu_count())
    st = _CPU_STATE.get(cfg)
    if st is None:
        sc, cam = S.make_config(cfg)
        gc, gd = S.output_grads(cam.W, cam.H)
        pre = G.preprocess(sc.means3D, sc.opacities, cam, shs=sc.shs, scales=sc.scales, rotations=sc.rotations, sh_degree=sc.sh_degree)
        keys, gid, ranges = G.bin_tiles(pre)
        st = _CPU_STATE[cfg] = dict(sc=sc, cam=cam, gc=gc, gd=gd, pre=pre, gid=gid, ranges=ranges)
    sc, cam, gc, gd, pre, gid, ranges = (st[k] for k in ("sc", "cam", "gc", "gd", "pre", "gid", "ranges"))
    bg = torch.zeros(3)
    P = sc.means3D.shape[0]
    sel = torch.arange(0, P, gauss_stride)
    gfrac = sel.numel() / P
    t0 = time.perf_counter()
    pre_s = G.preprocess(sc.means3D[sel], sc.opacities[sel], cam, shs=sc.shs[sel], scales=sc.scales[sel],
                         rotations=sc.rotations[sel], sh_degree=sc.sh_degree)
    G.bin_tiles(pre_s)
    t1 = time.perf_counter()
    sub = ranges.clone()
    keep = torch.zeros(ranges.shape[0], dtype=torch.bool)
    keep[::tile_stride] = True
    sub[~keep] = 0
    tfrac = float((sub[:, 1] - sub[:, 0]).sum()) / max(1.0, float((ranges[:, 1] - ranges[:, 0]).sum()))
    fw = G.blend_forward(pre, gid, sub, cam, bg)
    t2 = time.perf_counter()
    g2d = G.blend_backward(pre, gid, sub, cam, bg, fw, gc, gd)
    t3 = time.perf_counter()
    g2d_s = {k: v[sel] for k, v in g2d.items()}
    G.gaussian_backward(pre_s, g2d_s, sc.means3D[sel], cam, shs=sc.shs[sel], scales=sc.scales[sel], rotations=sc.rotations[sel],
                        sh_degree=sc.sh_degree)
    t4 = time.perf_counter()
    stages = [t1 - t0, t2 - t1, t3 - t2, t4 - t3]
    if stage_scale is None:
        stage_scale = [1.0 / gfrac, 1.0 / tfrac, 1.0 / tfrac, 1.0 / gfrac]
    full = sum(t * k for t, k in zip(stages, stage_scale))
    return 1.0 / full, {"measured_s": t4 - t0, "tile_fraction": tfrac, "gaussian_fraction": gfrac, "tile_stride": tile_stride,
                        "gauss_stride": gauss_stride, "stages": stages}


def _sample_text(info):
    return (f"rasterizer fwd+bwd of the C2 scene: alpha blend fwd+bwd on every {info['tile_stride']}th tile "
            f"({info['tile_fraction']:.3f} of the tile instances), per-Gaussian preprocess/backward on every "
            f"{info['gauss_stride']}th Gaussian, each stage scaled to the whole; {info['measured_s']:.2f} s of CPU work per step")


def run_reference(args, rank):
    """--impl reference: the reference's CPU implementation of the path = the CPU oracle, rank 0 only."""
    if rank != 0:
        return
    # Size the per-step sample so that K+W steps end within ~3 minutes.  The oracle's stages have fixed costs, so a
    # sample is NOT scaled by its work fraction: two untimed full passes (the second one warm) give each stage's
    # full-pass time, the chosen sample is timed the same way, and every step's stage times are scaled by
    # full / sample of that stage.  With stride 1 the factors are 1 and a step is the whole pass.
    n = args.warmup + args.steps
    cpu_oracle_rate(args.config)
    _, full = cpu_oracle_rate(args.config)
    t_full = sum(full["stages"])
    stride, scale, cal = 1, None, full
    while stride < 64 and n * sum(cal["stages"]) > 180.0:
        stride *= 2
        cpu_oracle_rate(args.config, tile_stride=stride, gauss_stride=stride)
        _, cal = cpu_oracle_rate(args.config, tile_stride=stride, gauss_stride=stride)
    if stride > 1:
        scale = [f / max(c, 1e-9) for f, c in zip(full["stages"], cal["stages"])]
    vals = []
    for i in range(n):
        v, info = cpu_oracle_rate(args.config, tile_stride=stride, gauss_stride=stride, stage_scale=scale)
        if i >= args.warmup:
            vals.append(v)
    v = len(vals) / sum(1.0 / x for x in vals)
    sample = _sample_text(info) + (f"; stage times scaled by full-pass / sample time measured once before the steps "
                                   f"(full pass {t_full:.2f} s)" if stride > 1 else "")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "iters/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "note": "reference has no CPU rasterizer; PyTorch-CPU golden renderer "
                       "(oracle/golden.py) timed on the host cores, rasterizer fwd+bwd only"},
            "cpu_baseline": {"value": v, "unit": "iters/s", "cores": os.cpu_count(), "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def ref_cuda_leg(cfg, dev, steps, warmup, flush):
    """The unmodified reference CUDA rasterizer (oracle/_ref) on this GPU: rasterizer fwd+bwd and the reference-style
    tracking iteration (eager loss, torch Adam, torch pose update)."""
    from oracle import ref_gpu, ref_tracking
    from monogs_b200 import scene as S
    if not ref_gpu.available():
        return {"unavailable": "oracle/_ref/libref_rasterizer.so not built"}
    sc, cam = S.make_config(cfg)
    scd = S.Scene(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in sc])
    camd = S.Camera(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in cam])
    bg = torch.zeros(3, device=dev)
    gc, gd = S.output_grads(cam.W, cam.H)
    gc, gd = gc.to(dev), gd.to(dev)
    state = {}

    def raster():
        state["fw"] = ref_gpu.forward(camd, bg, scd.means3D, scd.opacities, shs=scd.shs, scales=scd.scales,
                                      rotations=scd.rotations, sh_degree=sc.sh_degree)
        ref_gpu.backward(camd, bg, state["fw"], scd.means3D, gc, gd, shs=scd.shs, scales=scd.scales,
                         rotations=scd.rotations, sh_degree=sc.sh_degree)
    none = lambda: None
    ms_r = timed_steps(raster, steps, warmup, flush, none)
    tr, _, _, gt = build_tracker(cfg, dev, use_graph=False)
    rt = ref_tracking.RefTracker(scd, camd, bg, torch.eye(4), S.make_camera)
    rt.set_frame(gt)
    ms_t = timed_steps(rt.step, steps, warmup, flush, none)
    out = {"raster_fwd_bwd_iters_per_s": 1e3 * steps / ms_r, "tracking_iters_per_s": 1e3 * steps / ms_t,
           "what": "unmodified reference CUDA sources built for sm_100a (oracle/Makefile); tracking iteration = "
                   "reference wrapper allocations + eager Huber/L2 loss + torch Adam + torch pose update; mapping "
                   "iteration = config C4 (8-view window) with per-view activations, eager L1 loss, one autograd "
                   "backward, torch Adam"}
    try:
        out["mapping_iters_per_s"] = ref_mapping_leg(dev, flush)
    except Exception as ex:  # noqa: BLE001 - context only
        out["mapping_error"] = repr(ex)[:200]
    return out


def second_order_leg(cfg, dev, steps, warmup, flush, with_reference=True):
    """One second-order (sketched Gauss-Newton / LM) tracking iteration, stack_dim 16 x sketch_dim 64 (the shipped
    configs/mono/tum/base_config.yaml RGN.second_order): here = mgs_track_second_order replayed from a CUDA graph
    (the new sketch of every iteration is copied into the graph's buffers inside the timed region); reference = the same
    iteration assembled from the reference's pieces (oracle.ref_tracking.RefTracker.second_order_step: reference forward,
    16 x the reference's renderSketchJacobian kernel with its [P, 64, ...] buffers, float64 lstsq), on the same GPU."""
    from monogs_b200.tracking import PoseTracker
    from monogs_b200 import scene as S
    from oracle import ref_gpu, ref_tracking
    stacks, d = 16, 64
    tr, scd, cam, gt = build_tracker(cfg, dev, use_graph=True)
    for _ in range(40):                       # first-order phase (first_order.max_iter: 40)
        tr.step()
    tr.begin_second_order(1e-3)
    g = torch.Generator(device=dev).manual_seed(0)
    sketches = [PoseTracker.gen_sketch_args(cam.H, cam.W, stacks, d, device=dev, generator=g) for _ in range(4)]
    k = [0]

    def ours():
        if k[0] % 10 == 0:                    # second_order.max_iter: 10 iterations per frame
            tr.begin_second_order(1e-3, use_first_order_best=False)
        idx, w = sketches[k[0] % 4]
        k[0] += 1
        tr.set_sketch(idx, w, d)
        tr.second_order_step()
    ms = timed_steps(ours, steps, warmup, flush, lambda: None)
    out = {"iters_per_s": 1e3 * steps / ms, "ms_per_iter": ms / steps, "stack_dim": stacks, "sketch_dim": d,
           "lambda": float(tr.lm_state[0]), "step_norm": float(tr.lm_state[34]), "overflow": tr.check_capacity()[1],
           "api": "PoseTracker.set_sketch + second_order_step (mgs_track_second_order, CUDA graph)"}
    if with_reference and ref_gpu.available():
        sc, camc = S.make_config(cfg)
        camd = S.Camera(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in camc])
        rt = ref_tracking.RefTracker(scd, camd, torch.zeros(3, device=dev), torch.eye(4), S.make_camera)
        rt.set_frame(gt)
        for _ in range(3):
            rt.step()
        rt.begin_second_order(1e-3)
        n_ref, j = 3, [0]

        def ref():
            idx, w = sketches[j[0] % 4]
            j[0] += 1
            rt.second_order_step(idx, w, d)
        ms_ref = timed_steps(ref, n_ref, 1, flush, lambda: None)
        out["ref_cuda_iters_per_s"] = 1e3 * n_ref / ms_ref
        out["ref_cuda_ms_per_iter"] = ms_ref / n_ref
        out["speedup_vs_ref_cuda"] = (ms_ref / n_ref) / (ms / steps)
        out["ref_what"] = ("reference forward + 16 x _C.rasterize_gaussians_backward_sketch_jacobian (each with its zero-filled "
                           "[P,64,...] gradient buffers, rasterize_points.cu:308-318) + eager sketched objective + lstsq")
    return out


def ref_mapping_leg(dev, flush, P=200_000, window=8, iters=6, warmup=2):
    """Config C4 through oracle/ref_mapping.py (the reference CUDA rasterizer inside the reference-style iteration)."""
    from oracle import ref_gpu, ref_mapping
    from monogs_b200 import scene as S
    sc, cam = S.make_config("C2", P=P)
    bg = torch.zeros(3, device=dev)
    inv_sig = lambda x: torch.log(x / (1 - x))
    raw = {"xyz": sc.means3D.to(dev), "f_dc": sc.shs.to(dev).contiguous(),
           "opacity": inv_sig(sc.opacities.clamp(1e-4, 1 - 1e-4)).to(dev), "scaling": torch.log(sc.scales).to(dev),
           "rotation": sc.rotations.to(dev)}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}
    cams, gts = [], []
    for i in range(window):
        tau = torch.tensor(S.GT_TAU) * (i - window / 2) * 0.5
        c = S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, S.se3_exp(tau).float())
        c = S.Camera(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in c])
        with torch.no_grad():
            fw = ref_gpu.forward(c, bg, raw["xyz"], torch.sigmoid(raw["opacity"]), shs=raw["f_dc"], scales=torch.exp(raw["scaling"]),
                                 rotations=torch.nn.functional.normalize(raw["rotation"]), sh_degree=0)
        cams.append(c)
        gts.append((fw["color"] * 0.9 + 0.05).contiguous())
    rm = ref_mapping.RefMapper(raw, lrs, cams, gts, bg)
    ms = timed_steps(rm.step, iters, warmup, flush, lambda: None)
    return 1e3 * iters / ms


def mapping_leg(dev, rank, world, iters=40, warmup=5, P=200_000, window=8, engine=True, exchange="nccl", cfg="C2",
                ref_params=None, keep_params=None, n_extra=0, use_graph=True, scaling="strong", profile=False):
    """Config C4: 8-keyframe mapping window, views round-robin over ranks, one allreduce of the gradient arena,
    fused Adam on every rank (exchange="peer": reduce-scatter + Adam + parameter all-gather as ONE kernel over NVLink peer
    mappings, mgs_peer_reduce_adam).  engine=True: monogs_b200.mapping.MapEngine (mgs_map_*: raw parameters, activations once
    per iteration, no host sync); engine=False: the same iteration through the drop-in autograd surface (MappingStep).
    Returns iterations/s (max over ranks)."""
    import torch.distributed as dist
    from monogs_b200 import scene as S, ops
    from monogs_b200.mapping import MappingStep, MapEngine, MapView
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    sc, cam = S.make_config(cfg, P=P)
    deg = sc.sh_degree
    bg = torch.zeros(3, device=dev)
    inv_sig = lambda x: torch.log(x / (1 - x))
    raw = {"xyz": sc.means3D.to(dev), "f_dc": sc.shs[:, :1].to(dev).contiguous(),
           "opacity": inv_sig(sc.opacities.clamp(1e-4, 1 - 1e-4)).to(dev), "scaling": torch.log(sc.scales).to(dev),
           "rotation": sc.rotations.to(dev)}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}
    if sc.shs.size(1) > 1:      # config C5: SH degree 3 (gaussian_model.py:281: the rest coefficients train at lr / 20)
        raw["f_rest"], lrs["f_rest"] = sc.shs[:, 1:].to(dev).contiguous(), 2.5e-3 / 20
    shs_all = sc.shs.to(dev).contiguous()
    n_window = window
    window = window + n_extra          # the reference renders two random keyframes from outside the window beside it
    views = []
    for i in range(window):
        tau = torch.tensor(S.GT_TAU) * (i - window / 2) * 0.5
        c = S.make_camera(cam.W, cam.H, cam.fx, cam.fy, cam.cx, cam.cy, S.se3_exp(tau).float())
        c = S.Camera(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in c])
        rs = GaussianRasterizationSettings(c.H, c.W, c.tanfovx, c.tanfovy, bg, 1.0, c.viewmatrix, c.projmatrix,
                                           c.projmatrix_raw, deg, c.campos, False, False)
        with torch.no_grad():
            gt = GaussianRasterizer(rs)(means3D=raw["xyz"], means2D=torch.zeros_like(raw["xyz"]), shs=shs_all,
                                        opacities=torch.sigmoid(raw["opacity"]), scales=torch.exp(raw["scaling"]),
                                        rotations=torch.nn.functional.normalize(raw["rotation"]))[0]
        views.append((rs, (gt * 0.9 + 0.05).contiguous()))
    if engine:
        del shs_all
        eng = MapEngine({k: v.clone() for k, v in raw.items()}, lrs, sh_degree=deg, W=cam.W, H=cam.H, capacity_headroom=1.5,
                        exchange=exchange, use_graph=use_graph)
        mviews = [MapView(rs, gt) for rs, gt in views]
        mine = eng.set_window(mviews, n_window=n_window)      # placement: window round-robin, extras to the least-loaded ranks
        run = lambda: eng.step()                              # the installed window, replayed from a CUDA graph (NCCL path)
        nbytes = eng.arena.nbytes()
    else:
        params = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
        opt = ops.FusedAdam([{"params": [p], "lr": lrs[k]} for k, p in params.items()], lr=0.0, eps=1e-15)

        def view_loss(p, view):      # gaussian_renderer/__init__.py:79-168 activations + get_loss_mapping_rgb
            rs, gt = view
            image, _, depth, _, _ = GaussianRasterizer(rs)(
                means3D=p["xyz"], means2D=torch.zeros_like(p["xyz"]), shs=p["f_dc"], opacities=torch.sigmoid(p["opacity"]),
                scales=torch.exp(p["scaling"]), rotations=torch.nn.functional.normalize(p["rotation"]))
            return ops.mapping_loss(image, None, gt)

        def reg(p):                  # slam_backend.py:244-246
            s = torch.exp(p["scaling"])
            return 10 * (s - s.mean(dim=1, keepdim=True)).abs().mean()
        ms = MappingStep(params, opt, view_loss, reg)
        run = lambda: ms.step(views)
        nbytes = ms.arena.nbytes()
    for _ in range(warmup):
        run()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    from monogs_b200.mapping import assign_views
    per_rank = assign_views(window, world, n_window)
    out = {"iters_per_s": 1e3 * iters / float(t), "views_per_s": 1e3 * iters * window / float(t), "n_gpus": world, "window": n_window,
           "extra_views": n_extra, "gaussians": sc.means3D.size(0), "config": cfg,
           "views_per_rank": [len(x) for x in per_rank], "allreduce_bytes": nbytes if world > 1 else 0,
           "scaling": scaling,
           # NCCL path: the whole iteration is one replayed graph; fused exchanges: views + window-pose step are the graph,
           # mgs_peer_reduce_adam (host-side step / epoch numbers) is launched behind it
           "cuda_graph": (True if exchange == "nccl" else "compute part; exchange launched eagerly") if (engine and use_graph) else False,
           "api": "MapEngine (mgs_map_*)" if engine else "drop-in autograd surface (MappingStep)",
           "what": f"{n_window}-keyframe window{' + %d extra keyframes' % n_extra if n_extra else ''}: per-view render+mapping "
                   "loss+full backward, isotropic regulariser, one sum-allreduce of the gradient arena, fused Adam on every rank"}
    if engine and profile:
        # which kernel bounds this rank's iteration: library-side CUDA events over three eager iterations (rank 0's views)
        import ctypes as C
        from monogs_b200 import _lib as L
        eng.use_graph = False
        lib = L.lib()
        lib.mgs_profile_enable(1)
        for _ in range(3):
            eng.step()
        buf = C.create_string_buffer(8192)
        lib.mgs_profile_report(buf, 8192)
        lib.mgs_profile_enable(0)
        ker = {}
        for ln in buf.value.decode().strip().splitlines():
            name, k, ms = ln.split()
            ker[name] = round(1e3 * float(ms) / 3, 1)
        top = max(ker, key=ker.get)
        out["kernels_us_per_iteration_rank0"] = ker
        out["bound_by"] = {"kernel": top, "us_per_iteration": ker[top], "sum_of_kernels_us": round(sum(ker.values()), 1),
                           "iteration_us": round(float(t) / iters * 1e3, 1)}
    if engine:
        # cross-GPU parity inside the timed job: every rank ends with the same parameters (bit for bit: the checksums are
        # exact float64 sums of the same float32 values), and the fused exchanges end where the NCCL path ends
        flat = torch.cat([p.detach().reshape(-1) for p in eng.params.values()])
        chk = torch.stack([flat.double().sum(), flat.double().abs().max(), flat.double().pow(2).sum()])
        if world > 1:
            allc = [torch.zeros_like(chk) for _ in range(world)]
            dist.all_gather(allc, chk)
            out["params_identical_across_ranks"] = all(bool(torch.equal(c, allc[0])) for c in allc)
        out["param_checksum"] = [float(x) for x in chk]
        if ref_params is not None:
            out["max_rel_diff_vs_nccl_path"] = float((flat - ref_params).abs().max() / ref_params.abs().max())
        if keep_params is not None:
            keep_params.append(flat.clone())
        out["overflow"] = eng.check_capacity()[1]
        out["exchange"] = ("mgs_peer_reduce_adam: gradient reduce-scatter + Adam + parameter all-gather in one kernel over "
                           "NVLink peer mappings" if exchange == "peer" else "NCCL all_reduce of the gradient arena, then fused Adam")
        if exchange == "nvls":
            out["exchange"] = ("mgs_peer_reduce_adam over the NVSwitch multicast mapping: multimem.ld_reduce of the gradient "
                               "slices (summed in the switch) + Adam + multimem.st of the parameters, one kernel")
        if exchange != "nccl":
            out["peer_timeout_epoch"] = eng.block.error_epoch(None)
            out["allreduce_bytes"] = 0
            out["peer_bytes_in_out_per_rank"] = 2 * nbytes * (world - 1) // world
    return out


def e2e_dropin_leg(cfg, dev, steps, warmup, gt_host):
    """The drop-in autograd surface MonoGS calls (GaussianRasterizer + fused loss + FusedAdam + pose update), host
    target frame copied in and loss read back every step."""
    from monogs_b200 import scene as S, ops
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    sc, cam = S.make_config(cfg)
    scd = S.Scene(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in sc])
    bg = torch.zeros(3, device=dev)
    T = torch.eye(4, device=dev)
    proj = cam.projmatrix_raw.to(dev)
    theta = torch.zeros(3, device=dev, requires_grad=True); rho = torch.zeros(3, device=dev, requires_grad=True)
    a = torch.ones(1, device=dev, requires_grad=True); b = torch.zeros(1, device=dev, requires_grad=True)   # camera_utils.py:56-61
    opt = ops.FusedAdam([{"params": [theta], "lr": 3e-3}, {"params": [rho], "lr": 1e-3}, {"params": [a], "lr": 2e-2},
                         {"params": [b], "lr": 2e-2}], eps=1e-15)
    gt_dev = torch.empty(3, cam.H, cam.W, device=dev)
    view, full = T.t().contiguous(), (T.t() @ proj).contiguous()
    m2d = torch.zeros_like(scd.means3D)

    def step():
        nonlocal view, full
        gt_dev.copy_(gt_host, non_blocking=True)
        rs = GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, bg, 1.0, view, full, proj, 0, view, False, False)
        image, _, depth, opacity, _ = GaussianRasterizer(rs)(means3D=scd.means3D, means2D=m2d, shs=scd.shs,
                                                            opacities=scd.opacities, scales=scd.scales,
                                                            rotations=scd.rotations, theta=theta, rho=rho)
        loss = ops.tracking_loss_huber(image, opacity, gt_dev, a, b)
        opt.zero_grad()
        loss.backward()
        opt.step()
        _, view, full = ops.pose_update(T, rho, theta, proj_raw=proj, want_matrices=True)
        return loss.item()
    for _ in range(warmup): step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps): step()
    torch.cuda.synchronize()
    return steps / (time.perf_counter() - t0)


def main():
    args = parse()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)
    from monogs_b200.tracking import KERNELS_PER_ITER
    flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)   # > 126 MB L2
    tr, scd, cam, gt = build_tracker(args.config, dev, use_graph=True, rank=rank)
    step_count = [0]

    def step():
        # 100 pose iterations per frame (BASELINE.json config 2): re-arm the frame every 100 steps
        if step_count[0] % 100 == 0:
            tr.set_frame(gt, gt_depth=tr.bench_gt_depth)
        step_count[0] += 1
        tr.step()
    clocks = ClockSampler(local)
    if rank == 0: clocks.start()
    ms = timed_steps(step, args.steps, args.warmup, flush, barrier)
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([ms], device=dev)
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t)
    value = world * args.steps / (ms * 1e-3)
    # L2-warm variant (no flush), informational
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps): step()
    e1.record(); torch.cuda.synchronize()
    warm_ms = e0.elapsed_time(e1)
    n_rendered, overflow = tr.check_capacity()
    # ---- e2e: host frame in, loss + tangent out, every step, through the C-ABI tracker ----
    gt_host = gt.cpu().pin_memory()

    # PoseTracker.bind_host: one graph launch per iteration = [frame <- staging] -> {kernels || next frame's H2D into the
    # staging buffer on a second stream} -> loss + tangent D2H.  Every step moves its own 3.7 MB frame host -> device and
    # its loss + tangent device -> host inside the timed region.  e2e: step_host_lagged() (iteration k is launched before
    # the host waits for the results of k-1, two result slots); e2e_sync: a full synchronize after every launch.
    loss_host = torch.zeros(2).pin_memory(); tau_host = torch.zeros(2, 6).pin_memory()
    depth_host = tr.bench_gt_depth.cpu().pin_memory() if tr.bench_gt_depth is not None else None
    tr.bind_host(gt_host, loss_host, tau_host, depth_host)
    sink = [0.0]

    def arm():
        if step_count[0] % 100 == 0:
            tr.set_frame(tr._stage, gt_depth=tr._stage_depth)     # new frame: fresh Adam state (slam_frontend.py:453)
        step_count[0] += 1

    loss_np, tau_np = loss_host.numpy(), tau_host.numpy()     # views of the pinned result slots (no torch indexing per step)

    def e2e_lagged():
        arm()
        slot = tr.step_host_lagged()
        if slot is not None: sink[0] += float(loss_np[slot]) + float(tau_np[slot, 0])   # the host reads the result

    def e2e_sync_step():
        arm()
        tr.step_host()
        sink[0] += float(loss_np[0]) + float(tau_np[0, 0])

    def run_e2e(fn):
        for _ in range(args.warmup): fn()
        torch.cuda.synchronize(); barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps): fn()
        torch.cuda.synchronize(); barrier()
        te = torch.tensor([time.perf_counter() - t0], device=dev)
        if world > 1: dist.all_reduce(te, op=dist.ReduceOp.MAX)
        return world * args.steps / float(te)
    # host-timed and short (K steps of ~0.27 ms): one hiccup of the host (the GPU boxes are shared) moves a single run by
    # 10 %, so each variant is run five times (K steps each); the median is reported and every run is kept in the line
    e2e_runs = {}

    def median3(name, fn):
        e2e_runs[name] = [round(run_e2e(fn), 1) for _ in range(5)]
        return sorted(e2e_runs[name])[2]
    # host warm-up: the timed loops above leave the host thread mostly waiting, and the first few hundred milliseconds of a
    # host-driven loop then run slower than its steady state (observed: five consecutive 55 ms runs climbing from 3130 to
    # 3700 it/s); ~0.5 s of untimed steps first
    for _ in range(max(8 * args.steps, 1000)): e2e_sync_step()
    torch.cuda.synchronize()
    e2e_sync = median3("sync_every_step", e2e_sync_step)
    e2e = {"value": median3("lagged", e2e_lagged), "unit": "iters/s", "runs": "median of 5 runs of K steps", "all_runs": e2e_runs,
           "h2d_bytes_per_step": gt_host.numel() * 4 + (depth_host.numel() * 4 if depth_host is not None else 0),
           "d2h_bytes_per_step": 28, "value_sync_every_step": e2e_sync,
           "api": "monogs_b200.tracking.PoseTracker.bind_host/step_host_lagged over the C-ABI: one CUDA-graph launch per "
                  "iteration, the next frame's H2D copy beside the kernels, the host waits for iteration k-1 after launching k "
                  "(value_sync_every_step: step_host(), a stream synchronize after every launch)"}
    line = {"metric": METRIC, "value": value, "unit": "iters/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOADS[args.config], "gaussians": scd.means3D.size(0), "visible": int((tr.radii > 0).sum()),
                       "tile_instances": n_rendered, "instance_capacity": tr.R_cap, "overflow": overflow,
                       "iters_per_frame": 100,
                       "loss": ("L1 rgb + depth, alpha 0.95 (slam_utils.py:102-113)" if args.config == "C3" else
                                "Huber(0.01)+L2 per-pixel residual (slam_frontend.py:596-598)"),
                       "l2": "256 MiB buffer written between steps, outside the per-step event pairs",
                       "multi_gpu": "replicas only (tracking does not shard); see mapping"},
            "clocks": clk, "e2e": e2e, "gpu_launches": tr.kernels_per_iter * args.steps,
            "value_l2_warm": world * args.steps / (warm_ms * 1e-3)}
    if not args.no_extras:
        if world == 1:
            try:
                line["e2e_dropin"] = {"value": e2e_dropin_leg(args.config, dev, min(args.steps, 100), 5, gt_host), "unit": "iters/s",
                                      "api": "diff_gaussian_rasterization.GaussianRasterizer autograd surface + fused loss/Adam/pose ops"}
            except Exception as ex:  # the drop-in leg must not take the bench line down
                line["e2e_dropin"] = {"error": repr(ex)[:200]}
            prof, R, p_vis, R_cap = kernel_profile(args.config, dev, min(args.steps, 50))
            P = scd.means3D.size(0)
            ab = algorithmic_bytes(P, p_vis, R, R_cap, cam.W, cam.H, segments=tr.tile_segments)
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
            peak, src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
            for k, v in prof.items():
                if k in ab:
                    v["alg_bytes"] = ab[k]; v["gbs"] = ab[k] / (v["avg_us"] * 1e-6) / 1e9
            top = max(prof, key=lambda k: prof[k]["avg_us"])
            step_us = sum(v["avg_us"] for v in prof.values())
            # dram bytes / warp instructions / lanes of one launch: parsed from the committed ncu table of THIS code (C2 only);
            # a table captured before the kernel's source last changed is reported as stale, not used
            row, problem = profile_row(top + "_kernel") if args.config == "C2" else (None, "the ncu table is captured at C2")
            usable = row is not None and problem is None
            traffic = (row["rdMB"] + row["wrMB"]) * 1e6 if usable else None
            line["roofline"] = {"kernel": top, "bound": "hbm", "achieved": prof[top].get("gbs"), "peak": peak, "unit": "GB/s",
                                "frac": (prof[top].get("gbs") or 0) / peak, "traffic": traffic,
                                "traffic_source": PROFILE_FILE + (f" ({problem})" if problem else ""), "peak_source": src,
                                "share_of_step": prof[top]["avg_us"] / step_us,
                                "note": "HBM roofline of the dominant kernel.  The blend kernels are bound by warp-instruction "
                                        "issue and the fp32 pipe, not by HBM (one 48-B record feeds up to 256 pixel evaluations), "
                                        "so this fraction is small by construction; issue_roofline is the binding one"}
            if usable:
                inst = row["Minst"] * 1e6
                rate = inst / (prof[top]["avg_us"] * 1e-6)
                line["issue_roofline"] = {"kernel": row["kernel"], "warp_inst_per_launch": inst, "achieved_warp_inst_per_s": rate,
                                          "peak_warp_inst_per_s": ISSUE_PEAK_WARP_INST_PER_S, "frac": rate / ISSUE_PEAK_WARP_INST_PER_S,
                                          "lanes_active_of_32": row["lanes"], "lane_frac": rate / ISSUE_PEAK_WARP_INST_PER_S * row["lanes"] / 32.0,
                                          "issue_active_pct_ncu": row["issue%"], "fma_pipe_pct_ncu": row["fma%"],
                                          "source": PROFILE_FILE + ": instruction count and lanes from ncu, time measured live"}
            line["kernels"] = {k: {"avg_us": round(v["avg_us"], 2), "gbs": round(v.get("gbs", 0), 1)} for k, v in prof.items()}
            iter_bytes = sum(ab[k] for k in prof if k in ab)
            line["iteration_roofline"] = {"alg_bytes": iter_bytes, "achieved_gbs": iter_bytes / (ms / args.steps * 1e-3) / 1e9,
                                          "frac": iter_bytes / (ms / args.steps * 1e-3) / 1e9 / peak}
            try:
                line["ref_cuda"] = ref_cuda_leg(args.config, dev, min(args.steps, 50), 5, flush)
            except Exception as ex:
                line["ref_cuda"] = {"error": repr(ex)[:200]}
            if args.config in ("C1", "C2"):
                try:
                    line["second_order"] = second_order_leg(args.config, dev, min(args.steps, 100), 5, flush)
                except Exception as ex:
                    line["second_order"] = {"error": repr(ex)[:300]}
            cpu_oracle_rate(args.config, 8, 8)       # untimed: builds the scene + tile lists, warms the allocator
            v, info = cpu_oracle_rate(args.config, 1, 1)
            line["cpu_baseline"] = {"value": v, "unit": "iters/s", "cores": os.cpu_count(), "kind": "port", "sample": _sample_text(info)}
        nccl_params = []
        try:
            line["mapping"] = mapping_leg(dev, rank, world, keep_params=nccl_params, profile=True)
        except Exception as ex:
            line["mapping"] = {"error": repr(ex)[:300]}
        try:    # the reference-shaped iteration: the window plus two random keyframes from outside it (slam_backend.py:215-242)
            line["mapping_window_plus_2"] = mapping_leg(dev, rank, world, n_extra=2, iters=20, warmup=3)
        except Exception as ex:
            line["mapping_window_plus_2"] = {"error": repr(ex)[:300]}
        try:    # weak scaling: two views per rank whatever the number of ranks (compare views_per_s across N)
            line["mapping_weak"] = mapping_leg(dev, rank, world, window=2 * world, iters=20, warmup=3, scaling="weak")
        except Exception as ex:
            line["mapping_weak"] = {"error": repr(ex)[:300]}
        if world > 1:
            refp = nccl_params[0] if nccl_params else None
            try:
                line["mapping_peer"] = mapping_leg(dev, rank, world, exchange="peer", ref_params=refp)
            except Exception as ex:
                line["mapping_peer"] = {"error": repr(ex)[:300]}
            try:
                line["mapping_nvls"] = mapping_leg(dev, rank, world, exchange="nvls", ref_params=refp)
            except Exception as ex:
                line["mapping_nvls"] = {"error": repr(ex)[:300]}
        if world == 1:
            try:
                line["mapping_dropin"] = mapping_leg(dev, rank, world, iters=10, warmup=3, engine=False)
            except Exception as ex:
                line["mapping_dropin"] = {"error": repr(ex)[:300]}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
