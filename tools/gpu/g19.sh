mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x -k "masks or bit_for_bit or overhang or full_size or golden" > gpurun_out/g19_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g19_pytest.log
python bench.py --no-cpu-baseline --align-pairs 0 > gpurun_out/g19_bench.json 2> gpurun_out/g19_bench.err; echo "rc=$?" >> gpurun_out/g19_bench.err
