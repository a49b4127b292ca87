# the 8-rank jobs of round 2 (one box, 8 x B200): default workload, cfg4 and cfg5 as key-sharded jobs
NP=${NP:-8}
run() { name=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $NP "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; grep "^\[rank" gpurun_out/$name.err | tr ']' '\n' | grep -o "rank [0-9]*.*step" | head -8; python - <<PY
import json
try:
    d=json.load(open('gpurun_out/$name.json'))
    print('$name', 'value %.4g' % d['value'], 'ms %.2f' % d['ms_per_step'], d['scaling'], d['config'].get('sharded_parity'), 'e2e', d.get('e2e',{}).get('value'), 'inc', d.get('incremental',{}).get('value'), d['clocks'])
except Exception as e:
    print('$name: no line', e)
PY
}
export APM_BENCH_WATCHDOG=500
for w in "$@"; do
  case $w in
    cfg3) run bench_r2_${NP}gpu --steps 5 --warmup 3 --no-cpu-baseline ;;
    cfg3hash) run bench_r2_${NP}gpu_hash --placement hash --steps 5 --warmup 3 --no-cpu-baseline --no-e2e ;;
    cfg4) run bench_r2_${NP}gpu_cfg4_job --workload cfg4 --steps 3 --warmup 2 --no-cpu-baseline --no-e2e ;;
    cfg5) run bench_r2_${NP}gpu_cfg5_job --workload cfg5_job --steps 3 --warmup 2 --no-cpu-baseline ;;
  esac
done
