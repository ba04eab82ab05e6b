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
done
# full captures, one launch per kernel; summaries are made on the box, because gpurun brings back at most 64 MiB
# (no --import-source: with the CUB headers in the translation unit a report grows to 40 MB)
CMD="python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/plain2_c4.log 2>&1 && ncu --set full --clock-control none -k 'regex:sweep3' -s 2 -c 1 \
    -o gpurun_out/${TAG}_sweep_c4 $CMD > gpurun_out/ncu_f_c4.log 2>&1
$CMD > gpurun_out/plain3_c4.log 2>&1 && ncu --set full --clock-control none -k 'regex:contract_' -s 2 -c 1 \
    -o gpurun_out/${TAG}_contract_c4 $CMD > gpurun_out/ncu_g_c4.log 2>&1
CMD="python bench.py --workload c3 --steps 2 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/plain2_c3.log 2>&1 && ncu --set full --clock-control none -k 'regex:sweep3' -s 2 -c 1 \
    -o gpurun_out/${TAG}_sweep_c3 $CMD > gpurun_out/ncu_f_c3.log 2>&1
for R in gpurun_out/${TAG}_*.ncu-rep; do python tools/ncu_summary.py $R > ${R%.ncu-rep}.txt 2>/dev/null; done
python tools/ncu_lines.py gpurun_out/${TAG}_sweep_c4.ncu-rep sweep3_kernelILi5ELb0 40 > gpurun_out/${TAG}_sweep_c4_lines.txt 2>/dev/null
while [ $(du -sm gpurun_out | cut -f1) -gt 55 ]; do rm -f $(ls -S gpurun_out/*.ncu-rep | head -1); done
ls -la gpurun_out | tail -20
