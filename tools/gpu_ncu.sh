#!/bin/bash
# development aid: the round's ncu evidence -- launch list of the bench command + one --set full capture of each
# traversal kernel (after the plain command has exited 0)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec < /dev/null
CMD="python bench.py --steps 1 --warmup 3 --skip-micro --skip-cpu --skip-frame --skip-configs"
BENCH_E2E_PASSES=1 timeout 300 $CMD > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err || { echo "plain bench failed"; tail gpurun_out/ncu_plain.err; exit 1; }
BENCH_E2E_PASSES=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
STREAMS=1 timeout 200 python tools/diag_step.py 1 > gpurun_out/plain_step.log 2>&1 || { echo "plain step failed"; exit 1; }
for k in k_trace_fast k_shadow_fast k_shade; do
  STREAMS=1 timeout 500 ncu --set full --clock-control none --import-source on -k regex:$k -s 12 -c 1 -o gpurun_out/r2_$k -f python tools/diag_step.py 1 > gpurun_out/ncu_$k.log 2>&1
  echo "$k rc=$?"; tail -n 2 gpurun_out/ncu_$k.log
done
ls -la gpurun_out
