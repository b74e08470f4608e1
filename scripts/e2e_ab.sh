run() { python bench.py --no-extras --config ${CFG:-C2} 2>/dev/null | python -c "import json,sys; d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); print('$1 track %.1f warm %.1f e2e %.1f sync %.1f' % (d['value'], d['value_l2_warm'], d['e2e']['value'], d['e2e']['value_sync_every_step']), d['e2e'].get('all_runs'))"; }
for i in 1 2 3 4; do run rep$i; done
