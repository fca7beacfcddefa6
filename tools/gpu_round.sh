# round 2, call 11: several GOPs in flight on one GPU (feasibility)
set -x
timeout 300 python tools/inflight_bench.py --workers 1,2,3 --steps 24 > gpurun_out/r2n_inflight.txt 2>&1; cat gpurun_out/r2n_inflight.txt | tail -n 12
timeout 300 python tools/inflight_bench.py --workers 1,2,3 --steps 24 --e2e > gpurun_out/r2n_inflight_e2e.txt 2>&1; cat gpurun_out/r2n_inflight_e2e.txt | tail -n 12
