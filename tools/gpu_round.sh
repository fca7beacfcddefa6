# one gpurun call that validates the tree as the driver will run it: whole GPU suite, smoke(), both bench arms with the
# driver's flags, an ncu launch list of two bench steps.  (Earlier calls of round 2 used this file for their own scripts;
# the commands of each are quoted in profiles/README.md.)
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest rc $?"; tail -n 3 gpurun_out/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc $?"; tail -n 2 gpurun_out/final_smoke.log
python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/final_reference_arm.json 2> gpurun_out/final_reference_arm.err; echo "ref rc $?"; cut -c1-200 gpurun_out/final_reference_arm.json
python bench.py --gpus 1 --steps 20 --warmup 3 --breakdown > gpurun_out/final_bench.json 2> gpurun_out/final_bench_breakdown.txt; echo "bench rc $?"; tail -n 2 gpurun_out/final_bench_breakdown.txt
python - <<PY
import json
d=json.loads(open("gpurun_out/final_bench.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_step","inflight","gpu_launches","steps","warmup")}); print(d["e2e"]); print(d["sequential"]); print(d["latency_ms"]["encode_p50"], d["latency_ms"]["decode_p50"], d["clocks"]); print(d["roofline"]["frac"], d["roofline"]["traffic"])
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/final_ncu_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-configs --inflight 1 > gpurun_out/final_ncu_bench.log 2>&1; echo "ncu list rc $?"
