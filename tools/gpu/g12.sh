mkdir -p gpurun_out
python tools/align_probe.py 50000 > gpurun_out/g12_probe.log 2>&1
python tools/align_probe.py 50000 4000 >> gpurun_out/g12_probe.log 2>&1
python tools/align_probe.py 50000 600 >> gpurun_out/g12_probe.log 2>&1
python -m pytest tests/test_gpu_align.py -m gpu -q -x > gpurun_out/g12_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g12_pytest.log
