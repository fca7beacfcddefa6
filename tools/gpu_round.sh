# round 2, call 4: MMA rate probe with commits, octet stage-order A/B (timing + output difference), conv tests and bench with
# the new T selection (order 0 and order 1)
set -x
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I demo-learned-point-cloud-compression_b200/csrc tools/umma_rate_probe.cu -o /tmp/umma_rate_probe && timeout 120 /tmp/umma_rate_probe > gpurun_out/r2g_umma_rate_probe.txt 2>&1; tail -n 30 gpurun_out/r2g_umma_rate_probe.txt
python tools/octet_trace.py --exp > gpurun_out/r2g_octet_order_exp.txt 2>&1; cat gpurun_out/r2g_octet_order_exp.txt
python -m pytest tests/test_gpu_conv.py tests/test_gpu_pipeline.py -m gpu -x -q > gpurun_out/r2g_pytest_conv.log 2>&1; echo "pytest rc $?"; tail -n 3 gpurun_out/r2g_pytest_conv.log
PCC_OCTET_ORDER=1 python -m pytest tests/test_gpu_conv.py tests/test_gpu_pipeline.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2g_pytest_order1.log 2>&1; echo "pytest order1 rc $?"; tail -n 3 gpurun_out/r2g_pytest_order1.log
python bench.py --breakdown --no-configs --no-cpu-baseline > gpurun_out/r2g_bench_order0.json 2> gpurun_out/r2g_bench_order0_breakdown.txt; head -n 12 gpurun_out/r2g_bench_order0_breakdown.txt; tail -n 2 gpurun_out/r2g_bench_order0_breakdown.txt
PCC_OCTET_ORDER=1 python bench.py --breakdown --no-configs --no-cpu-baseline > gpurun_out/r2g_bench_order1.json 2> gpurun_out/r2g_bench_order1_breakdown.txt; head -n 12 gpurun_out/r2g_bench_order1_breakdown.txt; tail -n 2 gpurun_out/r2g_bench_order1_breakdown.txt
du -sh gpurun_out
