# the 8-rank jobs of round 2 (one box, 8 x B200): default workload, cfg4 and cfg5 as key-sharded jobs
run() { name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; grep "^\[rank" gpurun_out/$name.err | head -8; python - <<PY
import json
try:
    d=json.load(open('gpurun_out/$name.json'))
    print('$name', 'value %.4g' % d['value'], 'ms %.2f' % d['ms_per_step'], d['scaling'], d['config'].get('sharded_parity'), 'e2e', d.get('e2e',{}).get('value'), 'inc', d.get('incremental',{}).get('value'), d['clocks'])
except Exception as e:
    print('$name: no line', e)
PY
}
export APM_BENCH_WATCHDOG=500
run bench_r2_8gpu --steps 5 --warmup 3 --no-cpu-baseline
run bench_r2_8gpu_cfg4 --workload cfg4 --steps 3 --warmup 2 --no-cpu-baseline --no-e2e
run bench_r2_8gpu_cfg5 --workload cfg5_job --steps 3 --warmup 2 --no-cpu-baseline
