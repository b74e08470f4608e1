python bench.py > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err
for c in C1 C3 C5; do python bench.py --no-extras --config $c > gpurun_out/r02_bench_1gpu_$c.json 2>/dev/null; done
python - <<PY
import json
for f in ("", "_C1", "_C3", "_C5"):
    d = json.loads([l for l in open("gpurun_out/r02_bench_1gpu%s.json" % f) if l.startswith("{")][0])
    print(f or "C2", round(d["value"], 1), round(d["e2e"]["value"], 1), d["gpu_launches"], d["config"]["overflow"])
PY
