mkdir -p gpurun_out
for k in 0 3 7; do
python bench.py --seed-offset $k --no-cpu-baseline --align-pairs 0 --no-pipelined > gpurun_out/g27_bench_o$k.json 2> gpurun_out/g27_bench_o$k.err
done
