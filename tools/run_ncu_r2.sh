#!/bin/bash
# Round-2 ncu captures (one GPU): each command first runs plainly and must exit 0, then under ncu.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
ncu --query-metrics > gpurun_out/r2_ncu_query_metrics.txt 2>&1
EXTRA="smsp__inst_executed.sum,smsp__inst_issued.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_uniform.sum"
for c in "262144 plummer" "262144 disk" "1048576 plummer" "1048576 disk" "20000 plummer" "20000 disk"; do
  tag=$(echo $c | tr ' ' '_')
  python tools/ncu_case.py $c > gpurun_out/r2_ncu_plain_$tag.log 2>&1 &&
  ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:force_fast_kernel -s 2 -c 1 -o gpurun_out/r2_prof_$tag -f python tools/ncu_case.py $c > gpurun_out/r2_ncu_full_$tag.log 2>&1
  echo "$tag rc=$?"
done
python bench.py --steps 2 --warmup 3 > gpurun_out/r2_bench_plain_for_launchlist.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/r2_ncu_launches.log 2>&1
echo "launch list rc=$?"
for m in strict fp64; do python bench.py --workload c3 --mode $m --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_c3_$m.log 2>&1; echo "bench $m rc=$?"; done
python bench.py --workload c2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_c2.log 2>&1; echo "c2 rc=$?"
python bench.py --workload c3 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_c3.log 2>&1; echo "c3 rc=$?"
