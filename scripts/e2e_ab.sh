run() { python bench.py --no-extras 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1 track %.1f warm %.1f e2e %.1f sync %.1f' % (d['value'], d['value_l2_warm'], d['e2e']['value'], d['e2e']['value_sync_every_step']))"; }
MGS_TILE_SEGMENTS=1 MGS_PDL=1 run seg1_pdl1
MGS_TILE_SEGMENTS=1 MGS_PDL=0 run seg1_pdl0
MGS_TILE_SEGMENTS=0 MGS_PDL=1 run seg0_pdl1
MGS_TILE_SEGMENTS=0 MGS_PDL=0 run seg0_pdl0
