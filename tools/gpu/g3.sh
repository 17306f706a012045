mkdir -p gpurun_out
df -h /tmp > gpurun_out/g3_env.txt; free -g >> gpurun_out/g3_env.txt; nproc >> gpurun_out/g3_env.txt
python bench.py > gpurun_out/g3_bench.json 2> gpurun_out/g3_bench.err; echo "rc=$?" >> gpurun_out/g3_bench.err
