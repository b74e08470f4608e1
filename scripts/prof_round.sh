set -x
python scripts/prof_step.py 4 > gpurun_out/plain_step.log 2>&1 || exit 1
python scripts/prof_map.py 1 > gpurun_out/plain_map.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG:-r02}_launches_trk.csv python scripts/prof_step.py 4 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG:-r02}_launches_map.csv python scripts/prof_map.py 1 > /dev/null 2>&1
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,lts__t_bytes.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum
ncu --metrics $M --clock-control none --csv --log-file gpurun_out/${TAG:-r02}_metrics_trk.csv python scripts/prof_step.py 3 > /dev/null 2>&1
ncu --metrics $M --clock-control none --csv --log-file gpurun_out/${TAG:-r02}_metrics_map.csv python scripts/prof_map.py 1 > /dev/null 2>&1
ncu --set full --import-source on --clock-control none -k regex:blend_track -s 2 -c 1 -o gpurun_out/${TAG:-r02}_full_track -f python scripts/prof_step.py 4 > /dev/null 2>&1
ncu --set full --import-source on --clock-control none -k regex:blend -s 10 -c 2 -o gpurun_out/${TAG:-r02}_full_map -f python scripts/prof_map.py 1 > /dev/null 2>&1
ls -la gpurun_out/${TAG:-r02}*
