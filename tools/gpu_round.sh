# round 2, call 25: ncu --set full (with source) of the lane-rANS decoder and encoder on the cfg-2-shaped stream
set -x
ncu --set full --clock-control none --import-source on -k regex:k_rans_decode -s 1 -c 1 -f -o gpurun_out/r2y_rans_decode python tools/rans_bench.py --reps 1 --segs 16384 > gpurun_out/r2y_rans_decode_ncu.log 2>&1; echo "rc $?"
ncu --set full --clock-control none --import-source on -k regex:k_rans_encode -s 1 -c 1 -f -o gpurun_out/r2y_rans_encode python tools/rans_bench.py --reps 1 --segs 16384 > gpurun_out/r2y_rans_encode_ncu.log 2>&1; echo "rc $?"
ls -la gpurun_out/*.ncu-rep
