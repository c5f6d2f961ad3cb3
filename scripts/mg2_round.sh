set -x
W=${1:-2}
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $W --steps 40 --warmup 8 "${@:3}" > gpurun_out/$2.log 2>&1; echo exit=$? >> gpurun_out/$2.log; }
timeout 600 python -m pytest tests/test_multigpu.py -q -x > gpurun_out/pytest_mg$W.log 2>&1; echo exit=$? >> gpurun_out/pytest_mg$W.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29509 scripts/multigpu_check.py > gpurun_out/mgcheck_w$W.log 2>&1; echo exit=$? >> gpurun_out/mgcheck_w$W.log
run 29511 bench_n${W}_ours
DCGAN_COMM_SPLIT_G=0 run 29512 bench_n${W}_ours_nosplit
run 29513 bench_n${W}_ddp --impl ddp
run 29514 bench_n${W}_async --mode async
run 29515 bench_n${W}_nccl --comm nccl
