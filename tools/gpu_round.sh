# round 2, call 23: octet kernel with tiles handed out on demand (drawer warp of its own)
set -x
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_pipeline.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2w_pytest.log 2>&1; echo "pytest rc $?"; tail -n 5 gpurun_out/r2w_pytest.log
timeout 120 python tools/octet_trace.py 2>&1 | head -n 3
timeout 300 python tools/inflight_bench.py --workers 1,2 --steps 20 --reps 6 2>&1 | grep workers
timeout 300 python tools/inflight_bench.py --workers 2 --steps 20 --reps 6 --stagger-ms 2.5 --e2e 2>&1 | grep workers
