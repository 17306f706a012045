mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/g9_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g9_pytest.log
bash tests/sweep_parity.sh orfanage_b200/bin/orfanage /tmp/sw 5 6 > gpurun_out/g9_sweep.log 2>&1; echo "sweep rc=$?" >> gpurun_out/g9_sweep.log
python bench.py --align-pairs 0 --no-cpu-baseline > gpurun_out/g9_bench.json 2> gpurun_out/g9_bench.err; echo "rc=$?" >> gpurun_out/g9_bench.err
