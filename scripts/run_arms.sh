#!/bin/bash
# Both arms of bench.py back to back at N = 1 and N = $1 (default 2), the way the driver runs them; prints a summary.
N=${1:-2}; STEPS=${2:-10}
python bench.py --impl reference --steps $STEPS --warmup 3 > gpurun_out/b1_ref.json 2> gpurun_out/b1_ref.err; echo ref1 rc=$?
python bench.py --steps $STEPS --warmup 3 > gpurun_out/b1_ours.json 2> gpurun_out/b1_ours.err; echo ours1 rc=$?
if [ "$N" -gt 1 ]; then
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus $N --steps $STEPS --warmup 3 > gpurun_out/b${N}_ref.json 2> gpurun_out/b${N}_ref.err; echo ref$N rc=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps $STEPS --warmup 3 > gpurun_out/b${N}_ours.json 2> gpurun_out/b${N}_ours.err; echo ours$N rc=$?
fi
python - <<PY
import json
for f in ["b1_ref", "b1_ours"] + (["b${N}_ref", "b${N}_ours"] if $N > 1 else []):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        blk = lambda k: (d.get(k, {}).get("value"), d.get(k, {}).get("error") or d.get(k, {}).get("unavailable"))
        print(f, d.get("value"), d.get("ms_per_step"), "e2e", (d.get("e2e") or {}).get("value"), "| resnet50", blk("resnet50"), "| dcgan", blk("dcgan"), d.get("unavailable"), d.get("comm_check"))
    except Exception as e:
        print(f, "FAILED", e)
PY
