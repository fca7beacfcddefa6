# round 2, call 21 (8 GPUs): per-rank core pinning off (default: on)
set -x
PCC_BENCH_PIN=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 8 --steps 20 --warmup 3 --no-configs > gpurun_out/r2s_bench_n8_nopin.json 2> gpurun_out/r2s_bench_n8_nopin.err; echo "n8 nopin rc $?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r2s_bench_n8_nopin.json").read().strip().splitlines()[-1])
print("NOPIN", round(d["value"],1), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3), "seq", d["sequential"], d["rank_step_ms"])
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 8 --steps 20 --warmup 3 --no-configs > gpurun_out/r2s_bench_n8_pin.json 2> gpurun_out/r2s_bench_n8_pin.err; echo "n8 pin rc $?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r2s_bench_n8_pin.json").read().strip().splitlines()[-1])
print("PIN", round(d["value"],1), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3), "seq", d["sequential"], d["rank_step_ms"])
PY
lscpu -e | head -36
