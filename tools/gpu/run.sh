#!/bin/bash
# One parameterised driver for a gpurun call (replaces the one-off scripts of round 1).  Everything lands in gpurun_out/<tag>_*.
#   usage: tools/gpu/run.sh <tag> <step> [<step> ...]
#   steps: tests[:<pytest -k expr>]   pytest -m gpu (whole suite, or a -k selection)
#          smoke                      __graft_entry__.smoke()
#          bench[:<extra args>]       python bench.py <extra args>  -> <tag>_bench.json
#          mbench:<N>[:<extra args>]  the same under torchrun on N GPUs
#          launches[:<extra args>]    ncu launch list (gpu__time_duration) of a short bench command
#          full:<kernel regex>        ncu --set full of the matching kernels of a short bench command -> .ncu-rep
#          sanitize:<tool>:<k expr>   compute-sanitizer --tool <memcheck|racecheck|synccheck|initcheck> over a pytest selection
#          sancase:<tool>             compute-sanitizer over tools/gpu/sanitize_case.py
#          sh:<command>               anything else
TAG=$1; shift
mkdir -p gpurun_out
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-pipelined --align-pairs 20000"
for step in "$@"; do
  kind=${step%%:*}; arg=""; [ "$step" != "$kind" ] && arg=${step#*:}
  case $kind in
    tests)
      if [ -n "$arg" ]; then python -m pytest tests -m gpu -q -x -k "$arg" > gpurun_out/${TAG}_pytest.log 2>&1
      else python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; fi
      echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log; tail -15 gpurun_out/${TAG}_pytest.log ;;
    smoke)
      python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${TAG}_smoke.log; tail -3 gpurun_out/${TAG}_smoke.log ;;
    bench)
      label=$(echo "$arg" | grep -o "c4\|reference" | head -1); [ -n "$label" ] && label="_$label"
      python bench.py $arg > gpurun_out/${TAG}_bench${label}.json 2> gpurun_out/${TAG}_bench${label}.err; echo "rc=$?" >> gpurun_out/${TAG}_bench${label}.err
      tail -5 gpurun_out/${TAG}_bench${label}.err; head -c 1500 gpurun_out/${TAG}_bench${label}.json; echo ;;
    mbench)     # mbench:<N>[:<extra args>]  bench.py under torchrun on N GPUs -> <tag>_bench_<N>gpu[_<label>].json
      n=${arg%%:*}; rest=""; [ "$arg" != "$n" ] && rest=${arg#*:}
      label=$(echo "$rest" | grep -o "c4" | head -1); [ -n "$label" ] && label="_$label"
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n $rest \
        > gpurun_out/${TAG}_bench_${n}gpu${label}.json 2> gpurun_out/${TAG}_bench_${n}gpu${label}.err; echo "rc=$?" >> gpurun_out/${TAG}_bench_${n}gpu${label}.err
      tail -4 gpurun_out/${TAG}_bench_${n}gpu${label}.err; head -c 600 gpurun_out/${TAG}_bench_${n}gpu${label}.json; echo ;;
    launches)
      ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${TAG}_launches.csv $SHORT $arg > gpurun_out/${TAG}_ncu_launches.log 2>&1
      echo "launches rc=$?" ;;
    full)
      ncu --set full --clock-control none --import-source on -k "regex:$arg" -c 40 -o gpurun_out/${TAG}_full -f $SHORT > gpurun_out/${TAG}_ncu_full.log 2>&1
      echo "full rc=$?" ;;
    alignfull)   # ncu --set full of the alignment kernel alone (bench.py --only-align on <arg> pairs)
      ncu --set full --clock-control none --import-source on -k "regex:align_kernel" -c 2 -o gpurun_out/${TAG}_alignfull -f python bench.py --only-align --no-cpu-baseline --align-pairs ${arg:-20000} > gpurun_out/${TAG}_ncu_alignfull.log 2>&1
      echo "alignfull rc=$?" ;;
    sanitize)
      tool=${arg%%:*}; sel=${arg#*:}
      compute-sanitizer --tool $tool --error-exitcode 9 python -m pytest tests -m gpu -q -x -k "$sel" > gpurun_out/${TAG}_sanitize_${tool}.log 2>&1
      echo "sanitize $tool rc=$?" >> gpurun_out/${TAG}_sanitize_${tool}.log; tail -6 gpurun_out/${TAG}_sanitize_${tool}.log ;;
    sancase)    # sancase:<tool>  compute-sanitizer over tools/gpu/sanitize_case.py (every kernel once or twice, results checked)
      compute-sanitizer --tool $arg --error-exitcode 9 python tools/gpu/sanitize_case.py > gpurun_out/${TAG}_sanitize_${arg}.log 2>&1
      echo "sanitize $arg rc=$?" >> gpurun_out/${TAG}_sanitize_${arg}.log; tail -8 gpurun_out/${TAG}_sanitize_${arg}.log ;;
    sh)
      bash -c "$arg" ;;
    *) echo "unknown step $step" ;;
  esac
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > gpurun_out/${TAG}_smi.csv 2>&1
