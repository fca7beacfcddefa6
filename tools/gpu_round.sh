# one gpurun call: no-alloc test alone + whole suite, ncu --set full of the non-conv families, compute-sanitizer logs
set -x
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2e_pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc $?"
python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc $?"
tail -3 gpurun_out/r2e_pytest_noalloc_alone.log gpurun_out/r2e_pytest.log
ncu --set full --clock-control none -k regex:"$(python tools/ncu_families.py regex)" -c 400 -f -o gpurun_out/r2e_families python tools/ncu_families.py run --capture > gpurun_out/r2e_families_ncu.log 2>&1; echo "ncu families rc $?"
timeout 1200 compute-sanitizer --tool memcheck --log-file gpurun_out/r2e_sanitizer_memcheck.log python -m pytest tests/test_gpu_coords.py tests/test_gpu_entropy.py tests/test_gpu_octree.py tests/test_gpu_conv.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r2e_sanitizer_memcheck_pytest.log 2>&1; echo "memcheck rc $?"
tail -5 gpurun_out/r2e_sanitizer_memcheck.log gpurun_out/r2e_sanitizer_memcheck_pytest.log
timeout 900 compute-sanitizer --tool memcheck --log-file gpurun_out/r2e_sanitizer_memcheck_smoke.log python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2e_sanitizer_memcheck_smoke_out.log 2>&1; echo "memcheck smoke rc $?"
tail -5 gpurun_out/r2e_sanitizer_memcheck_smoke.log gpurun_out/r2e_sanitizer_memcheck_smoke_out.log
timeout 900 compute-sanitizer --tool racecheck --log-file gpurun_out/r2e_sanitizer_racecheck_smoke.log python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2e_sanitizer_racecheck_smoke_out.log 2>&1; echo "racecheck smoke rc $?"
tail -5 gpurun_out/r2e_sanitizer_racecheck_smoke.log gpurun_out/r2e_sanitizer_racecheck_smoke_out.log
