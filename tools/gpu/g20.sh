mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/g20_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g20_pytest.log
python bench.py --no-cpu-baseline --align-pairs 0 > gpurun_out/g20_bench.json 2> gpurun_out/g20_bench.err; echo "rc=$?" >> gpurun_out/g20_bench.err
