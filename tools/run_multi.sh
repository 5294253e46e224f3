# usage: bash tools/run_multi.sh N tag     (inside gpurun --gpus N): P2P check, gather check, bench p2p (+infer), bench nccl
N=$1; TAG=$2
T="timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
$T --master-port 29511 tools/p2p_check.py > gpurun_out/${TAG}_p2p_check_${N}gpu.log 2>&1; grep "launch 5\|floats\|captured\|CHECK" gpurun_out/${TAG}_p2p_check_${N}gpu.log
$T --master-port 29512 tools/gather_check.py > gpurun_out/${TAG}_gather_check_${N}gpu.log 2>&1; grep "world\|CHECK" gpurun_out/${TAG}_gather_check_${N}gpu.log
$T --master-port 29513 bench.py --gpus $N > gpurun_out/${TAG}_bench_${N}gpu_p2p.json 2> gpurun_out/${TAG}_bench_${N}gpu_p2p.err; python tools/bench_sum.py p2p < gpurun_out/${TAG}_bench_${N}gpu_p2p.json

python - <<'P'
import json,sys,glob
for f in sorted(glob.glob("gpurun_out/*_bench_*gpu_p2p.json")):
    d=json.loads(open(f).read().strip().splitlines()[-1]); i=d.get("infer") or {}
    print(f, "train", round(d["value"]), "ms", round(d["ms_per_step"],4), "infer", round(i.get("value",0)), i.get("ms_per_image"))
P
