"""Cross-GPU check and timing of mgs_peer_reduce_adam (run under torchrun, one rank per GPU of one node):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/check_peer_exchange.py [--out gpurun_out/peer_exchange.json]

1. parity: seeded per-rank gradient arenas; the peer exchange against the rank-ordered sum of the all-gathered arenas
   followed by the multi-tensor Adam — bit-exact parameters on every rank, bit-exact moments on the slice owners.
2. timing (CUDA events, max over ranks): the exchange alone, peer kernel vs NCCL all_reduce + Adam, at the arena sizes
   of configs C4 (200k Gaussians, SH 0) and C5 (2M Gaussians, SH 3).
3. the C4 mapping iteration through MapEngine with either exchange (bench.mapping_leg).
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

BETAS, EPS = (0.9, 0.999), 1e-15


def layout(P, M, dev):
    from monogs_b200.mapping import GradArena
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "opacity": 0.05, "scaling": 1e-3, "rotation": 1e-3}
    if M > 1:
        shapes["f_rest"], lrs["f_rest"] = (P, M - 1, 3), 2.5e-3 / 20
    protos = [torch.empty(s, device=dev) for s in shapes.values()]
    return GradArena(protos, extra=1), list(lrs.values()) + [-1.0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--no-mapping", action="store_true")
    ap.add_argument("--c5", action="store_true", help="also the C5 stress mapping iteration (2M Gaussians, 1920x1080, SH 3), "
                                                      "one view per rank, with either exchange")
    ap.add_argument("--no-nvls", action="store_true", help="skip the NVSwitch multicast variant")
    ap.add_argument("--timing-only", action="store_true", help="skip the parity pass (kernel-variant sweeps)")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from monogs_b200 import _lib as L, ops
    from monogs_b200.peer import PeerBlock
    lib = L.lib()
    stream = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    report = {"world": world, "variant": os.environ.get("MGS_PEER_VARIANT", "0")}

    if not args.timing_only:
        parity(report, rank, world, dev, lib, L, ops, PeerBlock, stream)
        if not args.no_nvls:
            try:
                parity(report, rank, world, dev, lib, L, ops, PeerBlock, stream, backend="symm")
            except Exception as ex:  # no multicast on this box: say so and go on with the unicast kernel
                report["nvls_unavailable"] = repr(ex)[:300]
                args.no_nvls = True
    timing(report, rank, world, dev, lib, L, ops, PeerBlock, stream, nvls=not args.no_nvls)
    exchanges = ("nccl", "peer") if args.no_nvls else ("nccl", "peer", "nvls")
    if not args.no_mapping:
        import bench
        report["mapping_C4"] = {ex: bench.mapping_leg(dev, rank, world, exchange=ex)["iters_per_s"] for ex in exchanges}
    if args.c5:
        import bench
        report["mapping_C5_one_view_per_rank"] = {
            ex: bench.mapping_leg(dev, rank, world, iters=6, warmup=2, P=None, window=world, exchange=ex, cfg="C5")["iters_per_s"]
            for ex in exchanges}
    if rank == 0:
        print(json.dumps(report))
        if args.out:
            with open(args.out, "w") as f:
                json.dump(report, f, indent=1)
    dist.destroy_process_group()


def parity(report, rank, world, dev, lib, L, ops, PeerBlock, stream, backend="ipc"):
    """backend="ipc": unicast TMA kernel, bit-exact.  backend="symm": the multicast kernel; the switch's summation order is
    its own, so parameters are compared by max |difference| / max |value| per tensor and must agree BETWEEN ranks bit for bit."""
    lay, lrs = layout(10_007, 1, dev)
    n = lay.numel
    blk = PeerBlock(n, dev, backend=backend).connect()
    worst = 0.0
    g = torch.Generator().manual_seed(17)
    p0 = torch.randn(n, generator=g).to(dev)
    blk.params.copy_(p0)
    m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    x = blk.exchange(m, v, lay.segment_ends, lrs)
    ref_p, ref_g = p0.clone(), torch.zeros(n, device=dev)
    rv, rg = lay.views_over(ref_p), lay.views_over(ref_g)
    for a, b in zip(rv, rg):
        a.grad = b
    opt = ops.FusedAdam([{"params": [a], "lr": lr} for a, lr in zip(rv, lrs[:-1])], lr=0.0, betas=BETAS, eps=EPS)
    ok = True
    for step in range(1, 5):
        gr = torch.Generator().manual_seed(100 * step + rank)
        mine = torch.randn(n, generator=gr).to(dev) * 10.0 ** (-(step % 4))
        blk.grads.copy_(mine)
        allg = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allg, mine)
        s = allg[0].clone()
        for t in allg[1:]:
            s += t
        ref_g.copy_(s)
        opt.step()
        torch.cuda.synchronize()
        dist.barrier()
        L.check(lib.mgs_peer_reduce_adam(C.byref(x), step, step, BETAS[0], BETAS[1], EPS, stream()), "mgs_peer_reduce_adam")
        torch.cuda.synchronize()
        if backend == "symm":
            for a, b in zip(rv, lay.views_over(blk.params)):
                worst = max(worst, float((a - b).abs().max() / a.abs().max()))
            worst = max(worst, float((blk.loss_out[0] - s[n - 4]).abs() / s[n - 4].abs().clamp_min(1e-30)))
            # every rank must hold the same bits: compare the arena with rank 0's
            flat = torch.cat([t_.reshape(-1) for t_ in lay.views_over(blk.params)])
            ref0 = flat.clone()
            dist.broadcast(ref0, src=0)
            ok &= bool(torch.equal(flat, ref0))
            ref_p.copy_(blk.params)      # follow the switch's sums so that the steps stay comparable one at a time
            continue
        for a, b in zip(rv, lay.views_over(blk.params)):
            ok &= bool(torch.equal(a, b))
        ok &= bool(torch.equal(blk.loss_out[:1], s[n - 4:n - 3]))
        lo, hi = blk.slice_of(rank)
        for k, (a, mv, vv) in enumerate(zip(rv, lay.views_over(m), lay.views_over(v))):
            beg = lay.segment_ends[k - 1] if k else 0
            i0, i1 = max(lo, beg) - beg, min(hi, beg + a.numel()) - beg
            if i1 > i0:
                ok &= bool(torch.equal(mv.reshape(-1)[i0:i1], opt.state[a]["exp_avg"].reshape(-1)[i0:i1]))
                ok &= bool(torch.equal(vv.reshape(-1)[i0:i1], opt.state[a]["exp_avg_sq"].reshape(-1)[i0:i1]))
    ok &= blk.error_epoch(stream()) == 0
    t = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if backend == "symm":
        w = torch.tensor([worst], device=dev)
        dist.all_reduce(w, op=dist.ReduceOp.MAX)
        report["nvls_ranks_bit_identical"] = bool(int(t))
        report["nvls_max_rel_diff_vs_rank_ordered_sum"] = float(w)
    else:
        report["parity_bit_exact"] = bool(int(t))
    blk.close()


def timing(report, rank, world, dev, lib, L, ops, PeerBlock, stream, nvls=True):
    def timed(fn, iters=20, warm=5):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(); dist.barrier()
        tt = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt) * 1e3   # us

    # NVLink reference: a plain peer copy of 256 MB (rank 0 pulls from the next device), the denominator of the
    # exchange's link bandwidth
    if rank == 0 and torch.cuda.device_count() > 1:
        src = torch.empty(64 << 20, device=torch.device("cuda", 1), dtype=torch.float32)
        dst = torch.empty(64 << 20, device=dev, dtype=torch.float32)
        for _ in range(3): dst.copy_(src)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): dst.copy_(src)
        e1.record(); torch.cuda.synchronize()
        report["p2p_copy_gbs"] = 10 * src.numel() * 4 / (e0.elapsed_time(e1) * 1e-3) / 1e9
        del src, dst
    dist.barrier()
    report["exchange_us"] = {}
    for name, P, M in (("C4", 200_000, 1), ("C5", 2_000_000, 16)):
        lay, lrs = layout(P, M, dev)
        n = lay.numel
        blk = PeerBlock(n, dev).connect()
        blk.params.normal_(); blk.grads.normal_()
        m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        x = blk.exchange(m, v, lay.segment_ends, lrs)
        state = {"e": 0}

        def peer():
            state["e"] += 1
            L.check(lib.mgs_peer_reduce_adam(C.byref(x), state["e"], state["e"], BETAS[0], BETAS[1], EPS, stream()), "peer")
        pp, gg = torch.randn(n, device=dev), torch.randn(n, device=dev)
        pv, gv = lay.views_over(pp), lay.views_over(gg)
        for a, b in zip(pv, gv):
            a.grad = b
        opt2 = ops.FusedAdam([{"params": [a], "lr": lr} for a, lr in zip(pv, lrs[:-1])], lr=0.0, betas=BETAS, eps=EPS)

        def nccl():
            dist.all_reduce(gg)
            opt2.step()
        t_peer, t_nccl = timed(peer), timed(nccl)
        t_nvls = None
        if nvls:
            blk2 = PeerBlock(n, dev, backend="symm").connect()
            blk2.params.normal_(); blk2.grads.normal_()
            x2 = blk2.exchange(m, v, lay.segment_ends, lrs)
            st2 = {"e": 0}

            def nvls_fn():
                st2["e"] += 1
                L.check(lib.mgs_peer_reduce_adam(C.byref(x2), st2["e"], st2["e"], BETAS[0], BETAS[1], EPS, stream()), "nvls")
            t_nvls = timed(nvls_fn)
            assert blk2.error_epoch(stream()) == 0
            del blk2, x2
        report["exchange_us"][name] = {"arena_bytes": 4 * n, "peer_reduce_adam": t_peer, "nccl_allreduce_plus_adam": t_nccl, "nvls_reduce_adam": t_nvls,
                                       "peer_timeout_epoch": blk.error_epoch(stream()),
                                       # every direction of a rank's links carries the gradient slices it pulls in (read responses) or serves, plus
                                       # the parameter slices it pushes / receives: 2 x (world-1)/world x arena bytes each way
                                       "peer_link_gbs_per_direction": 2 * 4 * n * (world - 1) / world / (t_peer * 1e-6) / 1e9}
        blk.close()
        del pp, gg, m, v, opt2



if __name__ == "__main__":
    main()
