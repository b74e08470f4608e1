"""profiles/<tag>_ncu_full_peer_kernel.txt from gpurun_out/<tag>_full_peer{,_c4}.ncu-rep (ncu --set full of
scripts/prof_peer.py, world = 1 on one GPU).  Usage: python scripts/summarise_peer_profile.py <tag> <commit>"""
import csv, subprocess, sys
tag, commit = sys.argv[1], sys.argv[2]
WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__grid_size', 'launch__registers_per_thread',
        'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum']
with open(f"profiles/{tag}_ncu_full_peer_kernel.txt", "w") as out:
    out.write(f"# ncu --set full --import-source on --clock-control none -k regex:peer_reduce_adam ; python scripts/prof_peer.py <cfg> 2 ; commit {commit}\n"
              "# world = 1 on one GPU (a multi-rank group can not run under kernel replay): the local-HBM side of the kernel.\n\n")
    for name, rep in (("C5 arena (472 MB)", f"gpurun_out/{tag}_full_peer.ncu-rep"), ("C4 arena (11.2 MB)", f"gpurun_out/{tag}_full_peer_c4.ncu-rep")):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(raw.splitlines())); h = rows[0]; r = rows[2]
        out.write(f"== {name}: {r[h.index('Kernel Name')]}\n")
        for w in WANT:
            if w in h: out.write(f"  {w:62s} {r[h.index(w)]:>18s} {rows[1][h.index(w)]}\n")
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(src.splitlines())); h = rows[1]; ci = {c: i for i, c in enumerate(h)}; k = ci['# Samples']
        data = [x for x in rows[2:] if len(x) > k]; tot = sum(float(x[k] or 0) for x in data)
        stalls = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
        agg = sorted(((sum(float(x[ci[s]] or 0) for x in data), s) for s in stalls), reverse=True)[:6]
        out.write("  warp stall samples: " + ", ".join(f"{s} {100 * v / tot:.0f}%" for v, s in agg) + "\n\n")
print(open(f"profiles/{tag}_ncu_full_peer_kernel.txt").read())
