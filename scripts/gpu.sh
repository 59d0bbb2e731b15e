#!/usr/bin/env bash
# One GPU visit, composed from tasks:   bash scripts/gpu.sh <tag> <task> [<task> ...]
#   tests[:<pytest -k expr>]  parity tests (-m gpu) -> gpurun_out/<tag>_pytest.log
#   smoke                     __graft_entry__.smoke()
#   bench[:<extra args>]      python bench.py --steps 50 --warmup 5 <extra args> -> <tag>_bench.json
#   launches[:<extra args>]   ncu launch list (gpu__time_duration) of a short bench -> <tag>_launches.csv + summary
#   ncu:<kernel regex>[:<skip>[:<bench args>]]  ncu --set full of ONE launch of that kernel -> <tag>_<regex>.ncu-rep + raw/source csv
#   sanitize:<tool>[:<pytest -k expr>]          compute-sanitizer --tool <memcheck|racecheck|synccheck|initcheck> over the GPU tests
#   py:<script args>          python <script args>   (log -> <tag>_py.log)
# Every task runs under its own timeout; a failing task does not stop the later ones.
set -u
TAG=${1:?tag}; shift
OUT=gpurun_out; mkdir -p $OUT
SHORT="--steps 3 --warmup 3 --no-cpu --no-extras --emitters 0"
for TASK in "$@"; do
  KIND=${TASK%%:*}; ARG=""; [ "$TASK" != "$KIND" ] && ARG=${TASK#*:}
  echo "=== $TASK"
  case $KIND in
    tests)
      timeout 1500 python -m pytest tests -m gpu -q ${ARG:+-k "$ARG"} > $OUT/${TAG}_pytest.log 2>&1
      echo "pytest exit $?"; tail -12 $OUT/${TAG}_pytest.log ;;
    smoke)
      timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -5 $OUT/${TAG}_smoke.log ;;
    bench)
      N=$(ls $OUT/${TAG}_bench*.json 2>/dev/null | wc -l)
      timeout 900 python bench.py --steps 50 --warmup 5 $ARG > $OUT/${TAG}_bench$N.json 2> $OUT/${TAG}_bench$N.err
      echo "bench exit $?"; cat $OUT/${TAG}_bench$N.json; tail -4 $OUT/${TAG}_bench$N.err ;;
    launches)
      python bench.py --steps 2 --warmup 3 --no-cpu $ARG > $OUT/${TAG}_plain.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $OUT/${TAG}_launches.csv \
        python bench.py --steps 2 --warmup 3 --no-cpu $ARG > $OUT/${TAG}_ncu_launches.log 2>&1
      echo "ncu launches exit $?"; python scripts/ncu_launch_summary.py $OUT/${TAG}_launches.csv | tee $OUT/${TAG}_launches_summary.txt ;;
    ncu)
      IFS=: read -r RX SKIP BARGS <<< "$ARG"; SKIP=${SKIP:-3}; BARGS=${BARGS:-$SHORT}
      NAME=$(echo "$RX" | tr -c 'A-Za-z0-9_' '_')
      python bench.py $BARGS > $OUT/${TAG}_plain_$NAME.log 2>&1 &&
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RX" -s $SKIP -c 1 -f -o $OUT/${TAG}_$NAME \
        python bench.py $BARGS > $OUT/${TAG}_ncu_$NAME.log 2>&1
      echo "ncu $RX exit $?"
      ncu -i $OUT/${TAG}_$NAME.ncu-rep --page raw --csv > $OUT/${TAG}_${NAME}_raw.csv 2>/dev/null
      ncu -i $OUT/${TAG}_$NAME.ncu-rep --page source --csv > $OUT/${TAG}_${NAME}_source.csv 2>/dev/null
      python scripts/ncu_report_summary.py $OUT/${TAG}_${NAME}_raw.csv $OUT/${TAG}_${NAME}_source.csv | tee $OUT/${TAG}_${NAME}_ncu.txt ;;
    sanitize)
      IFS=: read -r TOOL KEXPR <<< "$ARG"
      timeout 2400 compute-sanitizer --tool ${TOOL:-memcheck} --error-exitcode 7 --log-file $OUT/${TAG}_sanitize_${TOOL}.log \
        python -m pytest tests -m gpu -q -x ${KEXPR:+-k "$KEXPR"} > $OUT/${TAG}_sanitize_${TOOL}_pytest.log 2>&1
      echo "sanitize $TOOL exit $?"; tail -3 $OUT/${TAG}_sanitize_${TOOL}_pytest.log; tail -5 $OUT/${TAG}_sanitize_${TOOL}.log ;;
    py)
      timeout 1500 python $ARG > $OUT/${TAG}_py.log 2>&1; echo "py exit $?"; tail -30 $OUT/${TAG}_py.log ;;
    *) echo "unknown task $TASK" ;;
  esac
done
ls -la $OUT | tail -12
