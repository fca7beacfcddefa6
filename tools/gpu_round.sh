set -x
N=$(nvidia-smi -L | wc -l)
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2z_bench_n$N.json 2> gpurun_out/r2z_bench_n$N.err; echo "n$N rc $?"
python - <<PY
import json,glob
for f in glob.glob("gpurun_out/r2z_bench_n$N.json"):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d["n_gpus"], round(d["value"],1), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), "seq", round(d["sequential"]["value"],1), d["rank_step_ms"], {k:d["configs"]["cfg4"][k] for k in ("mpoints_s","frames_per_s","sequence_sha256","gather_seconds")})
PY
