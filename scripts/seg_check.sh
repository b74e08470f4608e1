timeout 600 python -m pytest tests/test_tile_segments_gpu.py -x -q 2>&1 | tail -30
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py --no-extras > gpurun_out/bench_seg.json 2> gpurun_out/bench_seg.err; tail -3 gpurun_out/bench_seg.err
python -c "import json; d=json.load(open('gpurun_out/bench_seg.json')); print('track %.1f warm %.1f e2e %.1f launches %d' % (d['value'], d['value_l2_warm'], d['e2e']['value'], d['gpu_launches'])); print(d['config'])"
for c in C1 C3; do python bench.py --no-extras --config $c 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$c %.1f launches %d overflow %s' % (d['value'], d['gpu_launches'], d['config']['overflow']))"; done
