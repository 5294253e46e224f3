# GPU tests, the no-alloc test alone in a fresh process, the HBM-side kernels event-timed, a training bench line,
# then the HBM-side kernels under ncu --set full
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r02z9_pytest_gpu.log 2>&1; tail -2 gpurun_out/r02z9_pytest_gpu.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02z9_pytest_no_alloc_alone.log 2>&1; tail -2 gpurun_out/r02z9_pytest_no_alloc_alone.log
python tools/perf_hbm.py > gpurun_out/r02z9_perf_hbm.log 2>&1; cat gpurun_out/r02z9_perf_hbm.log
python bench.py --no-cpu-baseline --no-infer --no-parity --steps 300 --warmup 20 > gpurun_out/r02z9_bench_train.json 2> gpurun_out/r02z9_bench_train.err
grep -h -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' gpurun_out/r02z9_bench_train.json
NWB_PERF_ITERS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'sample_|composite_' -c 40 \
  -o gpurun_out/r02z9_hbm -f python tools/perf_hbm.py > gpurun_out/r02z9_ncu.log 2>&1
tail -3 gpurun_out/r02z9_ncu.log
