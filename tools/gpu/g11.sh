mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/g11_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g11_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/g11_bench.json 2> gpurun_out/g11_bench.err; echo "rc=$?" >> gpurun_out/g11_bench.err
