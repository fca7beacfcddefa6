# round 2, call 6: rANS decoder (index array as a template parameter, running pointers, 256-bin search windows)
set -x
python -m pytest tests/test_gpu_entropy.py tests/test_gpu_pipeline.py tests/test_gpu_faults.py tests/test_gpu_single.py -m gpu -x -q > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc $?"; tail -n 4 gpurun_out/r2i_pytest.log
python tools/rans_bench.py --reps 5 > gpurun_out/r2i_rans_segs.txt 2>&1; cat gpurun_out/r2i_rans_segs.txt
python bench.py --breakdown --no-configs --no-cpu-baseline > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench_breakdown.txt; head -n 8 gpurun_out/r2i_bench_breakdown.txt; tail -n 1 gpurun_out/r2i_bench_breakdown.txt
