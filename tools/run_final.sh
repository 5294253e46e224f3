# round-end validation on one GPU: tests, smoke, default bench, reference arm, HBM-side kernels, ncu launch list of the bench
# usage: bash tools/run_final.sh <tag>
set -x
T=$1
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${T}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/${T}_bench_default.json 2> gpurun_out/${T}_bench_default.err; tail -c 600 gpurun_out/${T}_bench_default.json
python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err; tail -c 400 gpurun_out/${T}_bench_reference.json
python tools/perf_hbm.py > gpurun_out/${T}_perf_hbm.log 2>&1; cat gpurun_out/${T}_perf_hbm.log
python tools/graph_timeline.py > gpurun_out/${T}_graph_timeline.log 2>&1; cat gpurun_out/${T}_graph_timeline.log
S="--steps 8 --warmup 3 --no-cpu-baseline --no-infer --no-parity"
python bench.py $S > gpurun_out/${T}_bench_short.json 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py $S > gpurun_out/${T}_ncu_launches.log 2>&1
tail -2 gpurun_out/${T}_ncu_launches.log
