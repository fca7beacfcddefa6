# round 2, call 5: whole suite and bench after colour-head-on-candidates, sort key bits
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc $?"; tail -n 15 gpurun_out/r2h_pytest.log
python bench.py --breakdown --no-configs --no-cpu-baseline > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench_breakdown.txt; cat gpurun_out/r2h_bench_breakdown.txt; tail -c 1800 gpurun_out/r2h_bench.json
