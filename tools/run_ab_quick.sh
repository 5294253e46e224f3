# usage: bash tools/run_ab_quick.sh <tag> "<pytest targets>" [pairs]   -- a few GPU tests, then alternating A/B bench runs
T=$1; P=${3:-2}
python -m pytest $2 -m gpu -x -q 2>&1 | tail -2
B="--no-cpu-baseline --no-infer --no-parity --steps 300 --warmup 20"
for i in $(seq 1 $P); do
  (cd _ab_base && python bench.py $B) > gpurun_out/${T}_ab_base_$i.json 2> gpurun_out/${T}_ab_base_$i.err
  python bench.py $B > gpurun_out/${T}_ab_new_$i.json 2> gpurun_out/${T}_ab_new_$i.err
done
tail -2 gpurun_out/${T}_ab_new_1.err
grep -H -o '"ms_per_step": [0-9.]*' gpurun_out/${T}_ab_*.json
