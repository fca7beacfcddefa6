# round 2, call 13: variance of the two-in-flight arms (five runs)
set -x
for r in 1 2 3 4 5; do python bench.py --no-configs --no-cpu-baseline > gpurun_out/r2p_bench_run$r.json 2> gpurun_out/r2p_bench_run$r.err; echo "bench rc $?"; tail -n 2 gpurun_out/r2p_bench_run$r.err; python - <<PY
import json
d=json.loads(open("gpurun_out/r2p_bench_run$r.json").read().strip().splitlines()[-1])
print("RUN $r", round(d["value"],1), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), "seq", round(d["sequential"]["value"],1), round(d["sequential"]["e2e"]["value"],1), d["clocks"])
PY
done
