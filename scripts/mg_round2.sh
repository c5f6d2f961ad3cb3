W=${1:-2}
timeout 600 python -m pytest tests/test_multigpu.py -q -x > gpurun_out/pytest_mg$W.log 2>&1; echo exit=$? >> gpurun_out/pytest_mg$W.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29509 scripts/multigpu_check.py > gpurun_out/mgcheck_w$W.log 2>&1; echo exit=$? >> gpurun_out/mgcheck_w$W.log
bash scripts/mg_ab.sh $W gpurun_out/mg_ab_w$W.log "X=1 --mode=async" "DCGAN_ASYNC_DEFER_G=0 --mode=async" "DCGAN_COMM_BLOCKS=32 --mode=async" "X=0"
