mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/g8_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g8_pytest.log
bash tests/sweep_parity.sh orfanage_b200/bin/orfanage /tmp/sw 1 2 3 4 > gpurun_out/g8_sweep.log 2>&1; echo "sweep rc=$?" >> gpurun_out/g8_sweep.log
