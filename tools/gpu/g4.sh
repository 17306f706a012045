mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/g4_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g4_pytest.log
python bench.py > gpurun_out/g4_bench.json 2> gpurun_out/g4_bench.err; echo "rc=$?" >> gpurun_out/g4_bench.err
