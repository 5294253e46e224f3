# same-box A/B of the working tree against _ab_base (an export of HEAD): GPU tests first, then alternating bench runs
# usage: bash tools/run_ab.sh <tag> [pairs]
set -x
T=$1; P=${2:-2}
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -4 gpurun_out/${T}_pytest_gpu.log
B="--no-cpu-baseline --no-infer --no-parity --steps 300 --warmup 20"
for i in $(seq 1 $P); do
  (cd _ab_base && python bench.py $B) > gpurun_out/${T}_ab_base_$i.json 2> gpurun_out/${T}_ab_base_$i.err
  python bench.py $B > gpurun_out/${T}_ab_new_$i.json 2> gpurun_out/${T}_ab_new_$i.err
done
tail -3 gpurun_out/${T}_ab_new_1.err
grep -H -o '"ms_per_step": [0-9.]*\|"gpu_launches": [0-9]*' gpurun_out/${T}_ab_*.json
