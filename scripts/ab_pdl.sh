# A/B of the programmatic-dependent-launch trigger placement: variants/pdl_<mask>.so are builds with -DMGS_PDL_MASK=<mask>
#   for m in 0 2047 39; do MGS_NVCC_EXTRA="-DMGS_PDL_MASK=$m" bash monogs_b200/csrc/build.sh && cp monogs_b200/lib/libmonogs_b200.so variants/pdl_$m.so; done
#   gpurun -- 'MASKS="0 2047 39" PDLS="0 1" bash scripts/ab_pdl.sh'   (rebuild the default library afterwards)
# (bits: common.cuh PDL_*); each is copied over the product library and the tracking / mapping rates are taken.
run() {
  python bench.py --no-extras > gpurun_out/ab_tmp.json 2> gpurun_out/ab_tmp.err
  python -c "import json; d=json.load(open('gpurun_out/ab_tmp.json')); print('track %.1f warm %.1f e2e %.1f' % (d['value'], d['value_l2_warm'], d['e2e']['value']))"
}
for m in $MASKS; do
  cp variants/pdl_$m.so monogs_b200/lib/libmonogs_b200.so
  for pdl in $PDLS; do
    echo "== mask $m MGS_PDL=$pdl"
    MGS_PDL=$pdl run
    MGS_PDL=$pdl python scripts/map_rate.py 1 | tail -1
    MGS_PDL=$pdl python scripts/map_rate.py 8 | tail -1
  done
done 2>&1 | tee gpurun_out/ab_pdl_masks.log
