#!/bin/bash
# One parameterised runner for everything that goes through `gpurun` (replaces the per-round launchers).
#   gpurun --timeout 1500 -- 'bash scripts/gpu.sh <tag> <task> [<task> ...]'
# Every task writes gpurun_out/<tag>_<task>.log (and its artefacts); tasks run in order, a failing task does not stop
# the rest.  ncu tasks first run the same command without ncu (profiling recipe) and are skipped if that run fails.
#   tests[:<-k expr>]   pytest -m gpu (whole suite, or one -k expression) in its own process
#   smoke               __graft_entry__.smoke()
#   bench[:<args>]      python bench.py <args>          (JSON line -> <tag>_bench.json)
#   refarm[:<args>]     python bench.py --impl reference <args>
#   byshape[:<B>]       per-launch-shape table of one eager score step (DITSEP_BENCH_DUMP)
#   conv[:<B>]          scripts/bench_conv.py: conv_tc vs cuDNN per census shape
#   hbm[:<B>]           scripts/bench_hbm.py: HBM-bound kernels, GB/s
#   probe               scripts/umma_probe.cu (nvcc here, run there)
#   launches[:<B>]      ncu launch list of two eager score steps  -> <tag>_launches.csv
#   ncufull:<regex>[:<B>]  ncu --set full of up to 3 launches of the kernels matching <regex> -> <tag>_full.ncu-rep
#   py:<script and args>   python <script and args>
tag=$1; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,power.limit --format=csv,noheader > gpurun_out/${tag}_gpu.txt 2>&1
for task in "$@"; do
  name=${task%%:*}; arg=""; [[ "$task" == *:* ]] && arg=${task#*:}
  log=gpurun_out/${tag}_${name}.log
  echo "=== $task" | tee -a gpurun_out/${tag}_summary.log
  case $name in
    tests)
      if [ -n "$arg" ]; then timeout 1500 python -m pytest tests -m gpu -q -s -k "$arg" --timeout 600 > $log 2>&1
      else timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > $log 2>&1; fi
      tail -4 $log | tee -a gpurun_out/${tag}_summary.log; grep -E "FAILED|Error|rel-L2|SI-SDR" $log | head -40 >> gpurun_out/${tag}_summary.log ;;
    smoke)
      timeout 900 python -c "import __graft_entry__ as g; g.smoke()" > $log 2>&1; tail -8 $log | tee -a gpurun_out/${tag}_summary.log ;;
    bench)
      timeout 1500 python bench.py $arg > gpurun_out/${tag}_bench.json 2> $log; tail -c 3000 gpurun_out/${tag}_bench.json | tee -a gpurun_out/${tag}_summary.log; tail -3 $log ;;
    refarm)
      timeout 1500 python bench.py --impl reference $arg > gpurun_out/${tag}_refarm.json 2> $log; cat gpurun_out/${tag}_refarm.json | tee -a gpurun_out/${tag}_summary.log ;;
    byshape)
      DITSEP_BENCH_DUMP=gpurun_out/${tag}_byshape.txt timeout 900 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-library-baseline --batch ${arg:-32} > $log 2>&1
      head -40 gpurun_out/${tag}_byshape.txt | tee -a gpurun_out/${tag}_summary.log ;;
    conv)
      timeout 900 python scripts/bench_conv.py ${arg:-32} > $log 2>&1; tail -60 $log | tee -a gpurun_out/${tag}_summary.log ;;
    hbm)
      timeout 900 python scripts/bench_hbm.py ${arg:-32} > $log 2>&1; tail -40 $log | tee -a gpurun_out/${tag}_summary.log ;;
    probe)
      timeout 300 ./scripts/umma_probe > $log 2>&1; cat $log | tee -a gpurun_out/${tag}_summary.log ;;
    launches)
      export DITSEP_GRAPH=0
      timeout 600 python scripts/profile_step.py ${arg:-32} 2 > $log 2>&1 &&
      timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_launches.csv python scripts/profile_step.py ${arg:-32} 2 >> $log 2>&1
      unset DITSEP_GRAPH; tail -2 $log ;;
    ncufull)
      rx=${arg%%:*}; b=32; [[ "$arg" == *:* ]] && b=${arg#*:}
      export DITSEP_GRAPH=0
      timeout 600 python scripts/profile_step.py $b 1 > $log 2>&1 &&
      timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$rx -s 0 -c 3 -o gpurun_out/${tag}_full -f python scripts/profile_step.py $b 1 >> $log 2>&1
      unset DITSEP_GRAPH; tail -2 $log ;;
    py)
      timeout 1500 python $arg > $log 2>&1; tail -40 $log | tee -a gpurun_out/${tag}_summary.log ;;
    *) echo "unknown task $task" ;;
  esac
done
