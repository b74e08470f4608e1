"""Turns the raw ncu CSV / .ncu-rep files of scripts/prof_round.sh (gpurun_out/r01c_*) into the tracked text summaries
under profiles/.  Usage: python scripts/summarise_profiles.py <tag> <commit>"""
import collections
import csv
import subprocess
import sys

tag, commit = sys.argv[1], sys.argv[2]


def rows_of(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    h = rows[0]
    return [dict(zip(h, r)) for r in rows[1:]]


def short(name):
    name = name.replace("void ", "").replace("mgs::", "")
    return name.split("(")[0]


def launch_list(src, dst_txt, dst_csv, header):
    rows = rows_of(src)
    per = collections.OrderedDict()
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum": continue
        v = float(r["Metric Value"].replace(",", ""))
        us = v / 1e3 if r["Metric Unit"] in ("ns", "nsecond") else v
        per.setdefault(short(r["Kernel Name"]), []).append(us)
    tot = sum(sum(v) for v in per.values())
    with open(dst_txt, "w") as f:
        f.write(header)
        f.write(f"{'kernel':60s} {'n':>4s} {'total_us':>9s} {'avg_us':>8s} {'share':>6s}\n")
        for k, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"{k:60s} {len(v):4d} {sum(v):9.1f} {sum(v) / len(v):8.1f} {sum(v) / tot:6.3f}\n")
        f.write(f"{'total':60s} {'':4s} {tot:9.1f}\n")
    with open(dst_csv, "w") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel", "us"])
        for r in rows:
            if r["Metric Name"] == "gpu__time_duration.sum":
                v = float(r["Metric Value"].replace(",", ""))
                w.writerow([r["ID"], short(r["Kernel Name"]), v / 1e3 if r["Metric Unit"] in ("ns", "nsecond") else v])


def metric_table(src, f, title):
    rows = rows_of(src)
    per = collections.OrderedDict()
    for r in rows:
        per.setdefault(short(r["Kernel Name"]), collections.defaultdict(list))[r["Metric Name"]].append(
            float(r["Metric Value"].replace(",", "")))
    cols = [("gpu__time_duration.sum", "us", 1e-3), ("smsp__inst_executed.sum", "Minst", 1e-6),
            ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 1), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 1),
            ("smsp__thread_inst_executed_per_inst_executed.ratio", "lanes", 1), ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma%", 1),
            ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%", 1), ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%", 1),
            ("dram__bytes_read.sum", "rdMB", None), ("dram__bytes_write.sum", "wrMB", None), ("lts__t_bytes.sum", "L2MB", None),
            ("smsp__inst_executed_op_local_ld.sum", "lld", 1), ("smsp__inst_executed_op_local_st.sum", "lst", 1)]
    units = {}
    for r in rows: units[r["Metric Name"]] = r["Metric Unit"]
    f.write(title)
    f.write(f"{'kernel':44s} {'n':>3s} " + " ".join(f"{c[1]:>8s}" for c in cols) + "\n")
    for k, m in per.items():
        n = len(m["gpu__time_duration.sum"])
        vals = []
        for name, _, sc in cols:
            v = sum(m[name]) / max(len(m[name]), 1) if name in m else float("nan")
            if sc is None:
                u = units.get(name, "byte")
                sc = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
            if name == "gpu__time_duration.sum" and units.get(name) in ("us", "usecond"): sc = 1
            vals.append(v * sc)
        f.write(f"{k:44s} {n:3d} " + " ".join(f"{v:8.2f}" for v in vals) + "\n")
    f.write("\n")


def full_table(rep, f, title, keep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, u = rows[0], rows[1]
    f.write(title)
    names = [short(r[h.index("Kernel Name")]) for r in rows[2:]]
    f.write("metric | unit | " + " | ".join(names) + "\n")
    for i, name in enumerate(h):
        if name in keep:
            f.write(f"{name} | {u[i]} | " + " | ".join(r[i] for r in rows[2:]) + "\n")
    f.write("\n")


KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size", "launch__waves_per_multiprocessor",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"] + [
        f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio" for s in
        ("barrier", "short_scoreboard", "long_scoreboard", "wait", "not_selected", "math_pipe_throttle", "branch_resolving", "mio_throttle", "lg_throttle")]

hdr = (f"# ncu --metrics gpu__time_duration.sum --clock-control none  python scripts/prof_step.py 4\n"
       f"# config C2 (200k Gaussians, 640x480), the 4 un-graphed fused tracking iterations; commit {commit}\n"
       "# per-launch times are cold-cache and serialised: compare SHARES, not absolutes\n")
launch_list(f"gpurun_out/{tag}_launches_trk.csv", f"profiles/{tag}_launches_tracking_c2.txt", f"profiles/{tag}_launches_tracking_c2.csv", hdr)
hdr = (f"# ncu --metrics gpu__time_duration.sum --clock-control none  python scripts/prof_map.py 1\n"
       f"# config C4 (8-keyframe mapping window, 200k Gaussians, 640x480) through MapEngine: setup + 2 iterations; commit {commit}\n"
       "# per-launch times are cold-cache and serialised: compare SHARES, not absolutes\n")
launch_list(f"gpurun_out/{tag}_launches_map.csv", f"profiles/{tag}_launches_mapping_c4.txt", f"profiles/{tag}_launches_mapping_c4.csv", hdr)
import glob, hashlib, os
shas = " ".join(f"{os.path.basename(p)}={hashlib.sha1(open(p, 'rb').read()).hexdigest()[:12]}"
                for p in sorted(glob.glob("monogs_b200/csrc/*.cu") + glob.glob("monogs_b200/csrc/*.cuh")))
with open(f"profiles/{tag}_kernel_metrics.txt", "w") as f:
    # bench.py (profile_row) refuses numbers from this table for a kernel whose sources changed after the capture
    f.write(f"# source-sha1: {shas}\n")
    f.write(f"# ncu --metrics <13 metrics> --clock-control none; commit {commit}; averages over the launches of each kernel\n"
            "# us = gpu__time_duration, Minst = warp instructions, issue% = smsp__issue_active, occ% = sm__warps_active, lanes = threads per\n"
            "# instruction, fma/alu/xu% = pipe utilisation, rd/wr MB = DRAM bytes, L2MB = lts__t_bytes, lld/lst = local-memory instructions\n\n")
    metric_table(f"gpurun_out/{tag}_metrics_trk.csv", f, "## tracking iteration, config C2 (python scripts/prof_step.py 3)\n")
    metric_table(f"gpurun_out/{tag}_metrics_map.csv", f, "## mapping iteration, config C4 (python scripts/prof_map.py 1)\n")
with open(f"profiles/{tag}_ncu_full_blend_kernels.txt", "w") as f:
    f.write(f"# ncu --set full --clock-control none --import-source on; commit {commit}\n\n")
    full_table(f"gpurun_out/{tag}_full_track.ncu-rep", f, "## blend_track_kernel (tracking, C2): python scripts/prof_step.py 4, -k regex:blend_track -s 2 -c 1\n", KEEP)
    full_table(f"gpurun_out/{tag}_full_map.ncu-rep", f, "## mapping blend kernels (C4, MapEngine fast_math): python scripts/prof_map.py 1, -k regex:blend -s 10 -c 2\n", KEEP)
print("written")
