set -x
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) > gpurun_out/final_pytest.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1
TAG=r02 bash scripts/prof_round.sh > gpurun_out/prof_round.log 2>&1
bash scripts/final_1gpu.sh > gpurun_out/final_1gpu.log 2>&1
cat gpurun_out/final_pytest.log gpurun_out/final_smoke.log gpurun_out/final_1gpu.log
