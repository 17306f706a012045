mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x -k "cleaning or overhang or degenerate" > gpurun_out/g29_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g29_pytest.log
