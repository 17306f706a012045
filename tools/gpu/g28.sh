mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/g28_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g28_pytest.log
bash tests/sweep_parity.sh orfanage_b200/bin/orfanage /tmp/sw 3 > gpurun_out/g28_sweep.log 2>&1; echo "sweep rc=$?" >> gpurun_out/g28_sweep.log
python bench.py --no-cpu-baseline --align-pairs 0 > gpurun_out/g28_bench.json 2> gpurun_out/g28_bench.err; echo "rc=$?" >> gpurun_out/g28_bench.err
