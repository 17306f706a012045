# weak-scaling bench on N GPUs of one box (torchrun, one rank per GPU); N from $1
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --align-pairs 0 > gpurun_out/g17_bench_${N}gpu.json 2> gpurun_out/g17_bench_${N}gpu.err; echo "rc=$?" >> gpurun_out/g17_bench_${N}gpu.err
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline --align-pairs 0 > gpurun_out/g17_bench_1gpu_same_box.json 2>> gpurun_out/g17_bench_${N}gpu.err
nvidia-smi topo -m > gpurun_out/g17_topo.txt 2>&1
