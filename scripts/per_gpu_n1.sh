# One single-GPU bench per GPU of the box, all at once: is a GPU slower on its own, or only inside the 8-rank job?
mkdir -p gpurun_out/pergpu
for i in 0 1 2 3 4 5 6 7; do
  CUDA_VISIBLE_DEVICES=$i timeout 200 python bench.py --steps 40 --warmup 8 > gpurun_out/pergpu/gpu$i.json 2> gpurun_out/pergpu/gpu$i.err &
done
wait
nvidia-smi topo -m > gpurun_out/pergpu/topo.txt 2>&1
nvidia-smi --query-gpu=index,pci.bus_id,clocks.max.sm,clocks.max.mem,power.limit,memory.total --format=csv > gpurun_out/pergpu/gpus.csv 2>&1
