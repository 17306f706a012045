# codon-scan prefetch: parity, bench, scan kernel profile; then 4-GPU scaling if asked
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x -k "masks or bit_for_bit or overhang or full_size" > gpurun_out/g18_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g18_pytest.log
python bench.py --no-cpu-baseline --align-pairs 0 > gpurun_out/g18_bench.json 2> gpurun_out/g18_bench.err; echo "rc=$?" >> gpurun_out/g18_bench.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --align-pairs 0"
ncu --set full --clock-control none --import-source on -k regex:'codon_scan_kernel' -s 3 -c 1 -o gpurun_out/g18_prof $CMD > gpurun_out/g18_ncu.log 2>&1
