set -x
run() { name=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "rc=$?"; grep -v "^\[W\|^$\|^W1107" gpurun_out/$name.err | grep -i "error\|Traceback\|rank" | head -8; python -c "
import json,sys
d=json.load(open('gpurun_out/$name.json'))
print('$name', d['value'], d['ms_per_step'], d['scaling'], d['config'].get('sharded_parity'), d['config'].get('keys_per_rank'), d.get('e2e',{}).get('value'), d.get('incremental',{}).get('value'))"; }
export APM_BENCH_WATCHDOG=300 APM_BENCH_SCALE_KEYS=1

run t_cfg4 --workload cfg4 --scale 0.02 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e

