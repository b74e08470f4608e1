"""mgs_peer_reduce_adam (gradient reduce-scatter + Adam + parameter all-gather over peer mappings) against the step it
replaces — sum of the ranks' gradient arenas followed by the multi-tensor Adam (mgs_adam_step, itself tested against
torch.optim.Adam in test_ops_gpu.py).  On one GPU the "ranks" are blocks of one process driven from separate streams
(PeerBlock.local_group): flags, slices and the rank-ordered sum are the same code as across GPUs; the cross-GPU run is
scripts/check_peer_exchange.py under torchrun.  Bar: bit-exact (same element arithmetic, sum in rank order)."""
import ctypes as C
import json
import os
import subprocess
import sys

import pytest
import torch

from tests import util as U

pytestmark = pytest.mark.gpu
BETAS, EPS = (0.9, 0.999), 1e-15
LRS = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}
SHAPES = lambda P: {"xyz": (P, 3), "f_dc": (P, 1, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}


def _layout(P):
    from monogs_b200.mapping import GradArena
    protos = [torch.empty(s, device="cuda") for s in SHAPES(P).values()]
    return GradArena(protos, extra=1)


def _run_ranks(blocks, xs, step, epoch):
    """One exchange: every rank's two launches go to its own stream (the ranks wait for each other on the device)."""
    from monogs_b200 import _lib as L
    lib = L.lib()
    streams = [torch.cuda.Stream() for _ in blocks]
    torch.cuda.synchronize()
    for b, x, s in zip(blocks, xs, streams):
        L.check(lib.mgs_peer_reduce_adam(C.byref(x), step, epoch, BETAS[0], BETAS[1], EPS, C.c_void_p(s.cuda_stream)),
                "mgs_peer_reduce_adam")
    torch.cuda.synchronize()
    for b, s in zip(blocks, streams):
        assert b.error_epoch(C.c_void_p(s.cuda_stream)) == 0, "a rank timed out waiting for its peers"


@pytest.mark.parametrize("world,P", [(1, 1001), (2, 1001), (3, 777), (8, 5000), (8, 3)])
def test_peer_reduce_adam_matches_sum_then_adam(world, P):
    from monogs_b200 import ops
    from monogs_b200.peer import PeerBlock
    lay = _layout(P)
    n = lay.numel
    blocks = PeerBlock.local_group(n, world, torch.device("cuda", 0))
    g = torch.Generator().manual_seed(3)
    p0 = torch.randn(n, generator=g).cuda()
    ms = [torch.zeros(n, device="cuda") for _ in range(world)]
    vs = [torch.zeros(n, device="cuda") for _ in range(world)]
    xs = [b.exchange(ms[r], vs[r], lay.segment_ends, list(LRS.values()) + [-1.0]) for r, b in enumerate(blocks)]
    for b in blocks:
        b.params.copy_(p0)
    # the replaced step on a private copy: rank-ordered fp32 sum, then the multi-tensor Adam per parameter tensor
    ref_p = p0.clone()
    ref_views = lay.views_over(ref_p)
    ref_g = torch.zeros(n, device="cuda")
    ref_gv = lay.views_over(ref_g)
    for v, gv in zip(ref_views, ref_gv):
        v.grad = gv
    opt = ops.FusedAdam([{"params": [v], "lr": lr} for v, lr in zip(ref_views, LRS.values())], lr=0.0, betas=BETAS, eps=EPS)
    for step in range(1, 4):
        grads = [(torch.randn(n, generator=g) * (10.0 ** float(torch.randint(-6, 1, (1,), generator=g)))).cuda()
                 for _ in range(world)]
        for b, gr in zip(blocks, grads):
            b.grads.copy_(gr)
        s = grads[0].clone()
        for gr in grads[1:]:
            s += gr
        ref_g.copy_(s)
        opt.step()
        _run_ranks(blocks, xs, step, step)
        for r, b in enumerate(blocks):
            # parameters: every rank's arena holds every owner's slice; padding and the loss slot are not parameters
            for v_ref, v in zip(ref_views, lay.views_over(b.params)):
                assert torch.equal(v, v_ref), (step, r)
            assert torch.equal(b.loss_out[:1], s[lay.numel - 4: lay.numel - 3]), (step, r)
            assert torch.equal(b.grads, grads[r]), "the exchange must not write gradient arenas"
            lo, hi = b.slice_of(r)   # moments: live on the owner of the slice only
            for k, (v_ref, mv, vv) in enumerate(zip(ref_views, lay.views_over(ms[r]), lay.views_over(vs[r]))):
                st = opt.state[v_ref]
                beg = lay.segment_ends[k - 1] if k else 0
                a, e = max(lo, beg) - beg, min(hi, beg + v_ref.numel()) - beg
                if e > a:
                    assert torch.equal(mv.reshape(-1)[a:e], st["exp_avg"].reshape(-1)[a:e])
                    assert torch.equal(vv.reshape(-1)[a:e], st["exp_avg_sq"].reshape(-1)[a:e])
    for b in blocks:
        b.close()


def test_peer_wait_times_out_instead_of_hanging():
    """One rank of a two-rank group runs alone: its waits give up after the timeout and the block's error word records the epoch."""
    from monogs_b200 import _lib as L
    from monogs_b200.peer import PeerBlock
    lay = _layout(64)
    blocks = PeerBlock.local_group(lay.numel, 2, torch.device("cuda", 0))
    m, v = torch.zeros(lay.numel, device="cuda"), torch.zeros(lay.numel, device="cuda")
    x = blocks[0].exchange(m, v, lay.segment_ends, list(LRS.values()) + [-1.0], timeout_ms=1500)
    s = torch.cuda.current_stream()
    L.check(L.lib().mgs_peer_reduce_adam(C.byref(x), 1, 7, BETAS[0], BETAS[1], EPS, C.c_void_p(s.cuda_stream)), "peer")
    assert blocks[0].error_epoch(C.c_void_p(s.cuda_stream)) == 7
    for b in blocks:
        b.close()


def test_peer_argument_errors():
    from monogs_b200 import _lib as L
    from monogs_b200.peer import PeerBlock
    lib = L.lib()
    lay = _layout(64)
    b = PeerBlock.local_group(lay.numel, 1, torch.device("cuda", 0))[0]
    m, v = torch.zeros(lay.numel, device="cuda"), torch.zeros(lay.numel, device="cuda")
    x = b.exchange(m, v, lay.segment_ends, list(LRS.values()) + [-1.0])
    assert lib.mgs_peer_reduce_adam(None, 1, 1, 0.9, 0.999, EPS, None) == -1
    assert lib.mgs_peer_reduce_adam(C.byref(x), 0, 1, 0.9, 0.999, EPS, None) == -1      # step starts at 1
    assert lib.mgs_peer_reduce_adam(C.byref(x), 1, 0, 0.9, 0.999, EPS, None) == -1      # epoch starts at 1
    x.segment_end[x.n_segments - 1] -= 4                                                 # segments no longer cover the arena
    assert lib.mgs_peer_reduce_adam(C.byref(x), 1, 1, 0.9, 0.999, EPS, None) == -1
    assert b"segment" in lib.mgs_last_error_string()
    x.segment_end[x.n_segments - 1] += 4
    x.world = 9
    assert lib.mgs_peer_reduce_adam(C.byref(x), 1, 1, 0.9, 0.999, EPS, None) == -1
    b.close()


def test_map_engine_peer_exchange_matches_allreduce_path():
    """MapEngine(exchange="peer") on one rank against the default engine: same losses and parameters after several
    iterations up to the order of the float atomics in the backward, same statistics; densify re-creates the block."""
    from monogs_b200.mapping import MapEngine
    from tests.test_map_engine_gpu import _raw_params, _views
    sc, cam = U.small_case("dense")
    p0 = {k: v.detach().clone() for k, v in _raw_params(sc).items()}
    views = _views(sc, cam, 3, False, False)
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}
    engs = [MapEngine({k: v.clone() for k, v in p0.items()}, lrs, sh_degree=sc.sh_degree, W=cam.W, H=cam.H, fast_math=False,
                      exchange=ex) for ex in ("nccl", "peer")]
    for it in range(4):
        tot = [float(e.step(views)[0]) for e in engs]
        assert abs(tot[0] - tot[1]) <= 1e-5 * abs(tot[0]), it
    for k in engs[0].params:
        assert U.rel_err(engs[1].params[k], engs[0].params[k]) <= 1e-4, k
        st0, st1 = engs[0].opt.state[engs[0].params[k]], engs[1].opt.state[engs[1].params[k]]
        assert st0["step"] == st1["step"] == 4
        assert U.rel_err(st1["exp_avg"], st0["exp_avg"]) <= 1e-3, k
    assert torch.equal(engs[0].denom, engs[1].denom)
    peer = engs[1]
    grads = peer.xyz_gradient_accum / peer.denom
    thr = float(grads[~grads.isnan()].median())
    peer.sync_stats()
    out = peer.densify_and_prune(thr, 0.05, 6.0, 20.0, generator=torch.Generator(device="cuda").manual_seed(1))
    assert peer.P == out["counts"]["kept"] + out["counts"]["cloned"] + out["counts"]["children"]
    assert peer.block.numel == peer.arena.numel and peer.opt.state[peer.params["xyz"]]["step"] == 4
    l0 = float(peer.step(views)[0])
    for _ in range(5):
        l1 = float(peer.step(views)[0])
    assert l1 < l0 and peer.block.error_epoch(None) == 0


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs of one NVLink domain")
def test_peer_exchange_across_two_gpus(tmp_path):
    """The same exchange between two processes on two GPUs (CUDA IPC mappings, NVLink): scripts/check_peer_exchange.py
    under torchrun.  Unicast kernel: bit-exact against the rank-ordered sum + multi-tensor Adam.  Multicast (NVLS) kernel,
    where the box offers it: identical bits on both ranks, within 1e-5 of the rank-ordered result."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = tmp_path / "report.json"
    subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                    "127.0.0.1", "--master-port", "29533", os.path.join(root, "scripts", "check_peer_exchange.py"),
                    "--no-mapping", "--out", str(out)], check=True, timeout=600, cwd=root)
    r = json.load(open(out))
    assert r["parity_bit_exact"]
    for v in r["exchange_us"].values():
        assert v["peer_timeout_epoch"] == 0
    if "nvls_unavailable" not in r:
        assert r["nvls_ranks_bit_identical"] and r["nvls_max_rel_diff_vs_rank_ordered_sum"] <= 1e-5
