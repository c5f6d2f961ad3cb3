# usage: mg_ab.sh <world> <out.log> "ENV=..." ...   -- A/B of the multi-GPU bench under environment overrides
W=$1; out=$2; shift 2
: > $out
port=29520
for cfg in "$@"; do
  echo "== $cfg" >> $out
  port=$((port+1))
  envs=""; extra=""
  for tok in $cfg; do case $tok in --*) extra="$extra ${tok/=/ }";; *) envs="$envs $tok";; esac; done
  env $envs timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port $port bench.py --gpus $W --steps 40 --warmup 8 $extra 2>>$out.err | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['ms_per_step'], d['config'].get('step_ms_min_med_max'), d['config'].get('ms_per_step_by_rank'), 'e2e', d['e2e']['ms_per_step'], 'ok', d.get('multi_gpu_check',{}).get('ok'))
" >> $out
done
