TAG=r02 bash scripts/prof_round.sh > gpurun_out/prof_round.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:"tile_sort_segments|preprocess_kernel|track_update" -s 9 -c 3 -o gpurun_out/r02_full_binning -f python scripts/prof_step.py 5 > /dev/null 2>&1
bash scripts/final_1gpu.sh
