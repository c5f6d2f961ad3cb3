# Scaling round on one 8-GPU box: ours (sync), graph-captured DDP baseline, async mode at N = 1, 2, 4, 8, plus the
# numerics check at 8.  Each bench line lands in gpurun_out/scale/<impl>_n<N>.json
mkdir -p gpurun_out/scale
port=29600
one() {  # impl-tag N extra-args...
  tag=$1; N=$2; shift 2
  port=$((port+1))
  if [ "$N" = "1" ]; then
    timeout 300 python bench.py --gpus 1 --steps 50 --warmup 10 "$@" > gpurun_out/scale/${tag}_n$N.json 2> gpurun_out/scale/${tag}_n$N.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 50 --warmup 10 "$@" > gpurun_out/scale/${tag}_n$N.json 2> gpurun_out/scale/${tag}_n$N.err
  fi
  echo "$tag n=$N exit=$?" >> gpurun_out/scale/status.log
}
: > gpurun_out/scale/status.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29599 scripts/multigpu_check.py --sweep > gpurun_out/scale/mgcheck_w8.log 2>&1; echo "mgcheck exit=$?" >> gpurun_out/scale/status.log
for N in 8 4 2 1; do one ours $N; done
for N in 8 1; do one ddp_graph $N --impl ddp; done     # (N = 2, 4 of the baseline: profiles/scale_r2, unchanged code)
for N in 8 2; do one async $N --mode async; done
one wide_syncbn 8 --wide --sync-bn
one ddp_graph_wide_syncbn 8 --impl ddp --wide --sync-bn
one wide 1 --wide
cp gpurun_out/multigpu_check_w8.json gpurun_out/scale/ 2>/dev/null
