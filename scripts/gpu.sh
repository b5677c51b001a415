#!/bin/bash
# The one GPU-box runner (run under gpurun from the repo root):  scripts/gpu.sh <stage> [<stage> ...]
# Every stage writes under gpurun_out/ and prints a one-line verdict; a stage's failure does not stop the next one.
#   tests [pytest args]   GPU test suite (-m gpu)             smoke        __graft_entry__.smoke()
#   bench:<workload>      python bench.py --workload <w>      ref          bench.py --impl reference (CPU arm)
#   launches:<workload>   ncu launch list of the bench command (after the plain run exited 0)
#   ncu:<regex>:<script>  ncu --set full of kernels matching <regex> in `python <script>` (after the plain run exited 0)
#   py:<script>[:args]    python <script> args   (args separated by ',')
mkdir -p gpurun_out
STEPS=${STEPS:-5}; WARMUP=${WARMUP:-3}
for stage in "$@"; do
  kind=${stage%%:*}; rest=${stage#*:}; [ "$rest" == "$stage" ] && rest=""
  case $kind in
    tests)
      timeout ${T:-1500} python -m pytest tests -m gpu -q -x $rest > gpurun_out/t_gpu.log 2>&1; echo "[tests] rc=$?"; tail -4 gpurun_out/t_gpu.log ;;
    smoke)
      timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "[smoke] rc=$?"; tail -2 gpurun_out/smoke.log ;;
    bench)
      w=${rest:-v1080}
      timeout ${T:-900} python bench.py --steps $STEPS --warmup $WARMUP --workload $w $BENCH_ARGS > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "[bench $w] rc=$?"
      python scripts/bench_summary.py gpurun_out/bench_$w.json ;;
    ref)
      timeout 900 python bench.py --impl reference --steps 2 --warmup 1 ${rest:+--workload $rest} > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "[ref] rc=$?"; cat gpurun_out/bench_ref.json | cut -c1-400 ;;
    launches)
      w=${rest:-v1080}
      timeout 600 python bench.py --steps 2 --warmup 1 --workload $w --no-cpu-baseline --no-ref-gpu > gpurun_out/plain_$w.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c ${NCU_COUNT:-4000} --csv --log-file gpurun_out/launches_$w.csv \
          python bench.py --steps 2 --warmup 1 --workload $w --no-cpu-baseline --no-ref-gpu > gpurun_out/ncu_launches_$w.log 2>&1
      echo "[launches $w] rc=$?"; python scripts/ncu/launch_list_summary.py gpurun_out/launches_$w.csv 2>/dev/null | head -30 ;;
    ncu)
      regex=${rest%%:*}; script=${rest#*:}; args=${script#*:}; [ "$args" == "$script" ] && args=""; script=${script%%:*}
      timeout 300 python $script ${args//,/ } > gpurun_out/ncu_plain.log 2>&1 &&
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:$regex -c ${NCU_COUNT:-3} -f -o gpurun_out/prof_$(basename $script .py) \
          python $script ${args//,/ } > gpurun_out/ncu_full.log 2>&1
      echo "[ncu $regex] rc=$?"; tail -3 gpurun_out/ncu_full.log ;;
    py)
      script=${rest%%:*}; args=${rest#*:}; [ "$args" == "$rest" ] && args=""
      timeout ${T:-600} python $script ${args//,/ } 2>&1 | tail -${TAIL:-40}; echo "[py $script] rc=${PIPESTATUS[0]}" ;;
    *) echo "unknown stage $stage" ;;
  esac
done
