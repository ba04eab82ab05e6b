#!/bin/bash
# One GPU call that produces the round's evidence under gpurun_out/ (bench lines, launch list, full captures).
TAG=${1:-r02}
KF='regex:sweep3|contract_|resp_kernel|solve_kernel|scalars_kernel|lb_kernel|lb_finish|fill_kernel|init_params'
./tools/bin/fp64_peak > gpurun_out/${TAG}_fp64_peak.jsonl 2>&1
python bench.py --workload c3 --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_c3.json 2> gpurun_out/${TAG}_bench_c3.err
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_c4.json 2> gpurun_out/${TAG}_bench_c4.err
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
for WL in c3 c4; do
  CMD="python bench.py --workload $WL --steps 2 --warmup 3 --no-e2e --no-cpu"
  $CMD > gpurun_out/plain_$WL.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 400 --csv \
      --log-file gpurun_out/${TAG}_launches_$WL.csv $CMD > gpurun_out/ncu_l_$WL.log 2>&1
  $CMD > gpurun_out/plain2_$WL.log 2>&1 && ncu --set full --clock-control none --import-source on -k 'regex:sweep3|contract_mma' -s 6 -c 2 \
      -o gpurun_out/${TAG}_top_$WL $CMD > gpurun_out/ncu_f_$WL.log 2>&1
done
ls -la gpurun_out | tail -20
