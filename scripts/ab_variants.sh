# A/B of build variants: variants/<name>.so (builds with MGS_NVCC_EXTRA=-D<macro>=<value>, copied from monogs_b200/lib/) are
# copied over the product library in turn and the tracking (+ with WITH_MAP=1 the mapping) rates are taken:
#   gpurun -- 'VARIANTS="base coop2" bash scripts/ab_variants.sh'   (rebuild the default library afterwards)
for m in $VARIANTS; do
  cp variants/$m.so monogs_b200/lib/libmonogs_b200.so
  echo "== $m"
  python bench.py --no-extras 2>/dev/null | python -c "import json,sys; d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); print('track %.1f warm %.1f e2e %.1f' % (d['value'], d['value_l2_warm'], d['e2e']['value']))"
  if [ -n "$WITH_MAP" ]; then python scripts/map_rate.py 1 | tail -1; python scripts/map_rate.py 8 | tail -1; fi
done 2>&1 | tee gpurun_out/ab_variants.log
