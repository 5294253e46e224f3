# last run of the round on one GPU: every GPU test, smoke, default bench, reference arm, step timeline
set -x
T=$1
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -2 gpurun_out/${T}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/${T}_bench_default.json 2> gpurun_out/${T}_bench_default.err; tail -c 300 gpurun_out/${T}_bench_default.json
python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err; tail -c 200 gpurun_out/${T}_bench_reference.json
python tools/graph_timeline.py > gpurun_out/${T}_graph_timeline.log 2>&1; cat gpurun_out/${T}_graph_timeline.log
S="--steps 8 --warmup 3 --no-cpu-baseline --no-infer --no-parity"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py $S > gpurun_out/${T}_ncu_launches.log 2>&1
