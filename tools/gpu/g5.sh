mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/g5_plain.json 2> gpurun_out/g5_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/g5_launches.csv $CMD > gpurun_out/g5_ncu1.log 2>&1
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'codon_scan_kernel|base_planes_kernel|rescue_score_kernel|fill_rows_kernel|select_kernel|count_rows_kernel' -s 18 -c 6 -o gpurun_out/g5_prof $CMD > gpurun_out/g5_ncu2.log 2>&1
ls -la gpurun_out
