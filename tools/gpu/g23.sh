mkdir -p gpurun_out
for m in 1 2 3 4; do
ORFX_HEAVY_MULT=$m python bench.py --no-cpu-baseline --align-pairs 0 --no-pipelined > gpurun_out/g23_bench_m$m.json 2> gpurun_out/g23_bench_m$m.err
done
