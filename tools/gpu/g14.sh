mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "full_size" > gpurun_out/g14_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g14_pytest.log
python tools/gpu/rows_hist.py > gpurun_out/g14_rows.log 2>&1
