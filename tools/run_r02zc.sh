set -x
python -m pytest tests/test_render_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
B="--no-cpu-baseline --no-infer --no-parity --steps 300 --warmup 20"
for i in 1 2; do
  (cd _ab_base && python bench.py $B) > gpurun_out/r02zc_ab_base_$i.json 2> gpurun_out/r02zc_ab_base_$i.err
  python bench.py $B > gpurun_out/r02zc_ab_new_$i.json 2> gpurun_out/r02zc_ab_new_$i.err
done
tail -3 gpurun_out/r02zc_ab_new_1.err
grep -H -o '"ms_per_step": [0-9.]*' gpurun_out/r02zc_ab_*.json
python tools/perf_hbm.py > gpurun_out/r02zc_perf_hbm.log 2>&1; cat gpurun_out/r02zc_perf_hbm.log
