mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/g30_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g30_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/g30_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/g30_smoke.log
