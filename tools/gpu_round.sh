# round 2, call 24: whole suite and bench after the shorter sort keys of lex_order
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc $?"; tail -n 3 gpurun_out/r2x_pytest.log
python bench.py --breakdown --no-configs --no-cpu-baseline > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench_breakdown.txt; echo "bench rc $?"; tail -n 1 gpurun_out/r2x_bench_breakdown.txt
python - <<PY
import json
d=json.loads(open("gpurun_out/r2x_bench.json").read().strip().splitlines()[-1])
print(round(d["value"],1), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), "seq", round(d["sequential"]["value"],1), round(d["sequential"]["e2e"]["value"],1), d["latency_ms"]["encode_p50"], d["latency_ms"]["decode_p50"])
PY
