mkdir -p gpurun_out
for v in rs8 rs6; do
ORFX_LIB=$PWD/orfanage_b200/liborfx_$v.so python bench.py --no-cpu-baseline --align-pairs 0 --no-pipelined > gpurun_out/g32_bench_$v.json 2> gpurun_out/g32_bench_$v.err
done
