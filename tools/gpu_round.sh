# round 2, call 18: two callers in phase vs out of phase; persistent conv grids capped below 148 SMs
set -x
for st in 0 2.5; do
  timeout 300 python tools/inflight_bench.py --workers 2 --steps 20 --reps 8 --stagger-ms $st 2>&1 | grep workers
  timeout 300 python tools/inflight_bench.py --workers 2 --steps 20 --reps 8 --stagger-ms $st --e2e 2>&1 | grep workers
done
for cap in 140 132; do
  PCC_CONV_SMS=$cap timeout 300 python tools/inflight_bench.py --workers 1,2 --steps 20 --reps 5 2>&1 | grep "workers" | sed "s/^/cap $cap: /"
done
