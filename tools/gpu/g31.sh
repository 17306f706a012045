mkdir -p gpurun_out
bash tests/sweep_parity.sh orfanage_b200/bin/orfanage /tmp/sw 4 5 6 7 8 9 > gpurun_out/g31_sweep.log 2>&1; echo "sweep rc=$?" >> gpurun_out/g31_sweep.log
