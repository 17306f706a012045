mkdir -p gpurun_out
for v in s8 s16; do for k in 0 3 7; do
ORFX_LIB=$PWD/orfanage_b200/liborfx_$v.so python bench.py --seed-offset $k --no-cpu-baseline --align-pairs 0 --no-pipelined > gpurun_out/g26_bench_${v}_o$k.json 2> gpurun_out/g26_bench_${v}_o$k.err
done; done
