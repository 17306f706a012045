# final build: default bench line + launch list of the short command
mkdir -p gpurun_out
python bench.py > gpurun_out/g33_bench.json 2> gpurun_out/g33_bench.err; echo "rc=$?" >> gpurun_out/g33_bench.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --align-pairs 20000 --no-pipelined"
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/g33_launches.csv $CMD > gpurun_out/g33_ncu1.log 2>&1
