# 2-GPU weak-scaling run of the bench, launched the way the driver does
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/g7_smi.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 --align-pairs 0 > gpurun_out/g7_bench2.json 2> gpurun_out/g7_bench2.err; echo "rc=$?" >> gpurun_out/g7_bench2.err
python bench.py --gpus 1 --steps 20 --warmup 3 --align-pairs 0 --no-cpu-baseline > gpurun_out/g7_bench1.json 2> gpurun_out/g7_bench1.err; echo "rc=$?" >> gpurun_out/g7_bench1.err
