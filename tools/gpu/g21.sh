# default torchrun bench on N GPUs (alignment leg sharded over the ranks)
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/g21_bench_${N}gpu.json 2> gpurun_out/g21_bench_${N}gpu.err; echo "rc=$?" >> gpurun_out/g21_bench_${N}gpu.err
