# tests + CLI sweep + default bench + launch list + ncu --set full of every kernel of one step and of the alignment kernel
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/g24_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g24_pytest.log
bash tests/sweep_parity.sh orfanage_b200/bin/orfanage /tmp/sw 1 2 > gpurun_out/g24_sweep.log 2>&1; echo "sweep rc=$?" >> gpurun_out/g24_sweep.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/g24_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/g24_smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/g24_bench_reference.json 2> gpurun_out/g24_bench_reference.err
python bench.py > gpurun_out/g24_bench.json 2> gpurun_out/g24_bench.err; echo "rc=$?" >> gpurun_out/g24_bench.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --align-pairs 20000"
$CMD > gpurun_out/g24_plain.json 2> gpurun_out/g24_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/g24_launches.csv $CMD > gpurun_out/g24_ncu1.log 2>&1
$CMD > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'codon_scan_kernel|base_planes.*|rescue_score_kernel|fill_rows_kernel|select_kernel|count_rows_kernel|heavy_.*|tx_prep_kernel' -s 30 -c 12 -o gpurun_out/g24_prof $CMD > gpurun_out/g24_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'align_kernel' -s 1 -c 1 -o gpurun_out/g24_prof_align $CMD > gpurun_out/g24_ncu3.log 2>&1
ls -la gpurun_out
