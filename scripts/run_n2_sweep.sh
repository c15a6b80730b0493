# N = 2 A/B of the benchmark under different collective settings (run on a 2-GPU box)
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/n2_$name.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$name', d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d.get('comm_check',{}).get('max_abs_err_vs_nccl'), d['final_loss'])
except Exception as e:
    print('$name FAILED', e)"
}
python bench.py --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('n1', d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['final_loss'])"
for cfg in "$@"; do run $(echo $cfg | tr ',' ' '); done
