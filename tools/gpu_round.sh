# round 2, call 26: lane encoder with 4-round prefetch chunks (no spill): coder tests, coder timing
set -x
python -m pytest tests/test_gpu_entropy.py tests/test_gpu_faults.py tests/test_gpu_pipeline.py -m gpu -x -q > gpurun_out/r2z_pytest.log 2>&1; echo "pytest rc $?"; tail -n 2 gpurun_out/r2z_pytest.log
python tools/rans_bench.py --reps 5 > gpurun_out/r2z_rans_segs.txt 2>&1; cat gpurun_out/r2z_rans_segs.txt
