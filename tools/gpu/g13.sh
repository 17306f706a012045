mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --align-pairs 0"
ncu --set full --clock-control none --import-source on -k regex:'select_kernel|rescue_score_kernel|heavy_rescue_kernel' -s 9 -c 3 -o gpurun_out/g13_prof $CMD > gpurun_out/g13_ncu.log 2>&1
ls -la gpurun_out | tail -3
