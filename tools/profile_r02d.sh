#!/bin/bash
# End-of-round evidence (r02d): micro-benchmarks, bench lines, launch lists, full captures of the plain and the bound sweep.
TAG=r02d
KF='regex:sweep3|contract_|resp_kernel|solve_kernel|scalars_kernel|lb_kernel|lb_finish|fill_kernel|init_params'
./tools/bin/fp64_peak > gpurun_out/${TAG}_fp64_peak.jsonl 2>&1
./tools/bin/dmma_peak > gpurun_out/${TAG}_dmma_peak.jsonl 2>&1
./tools/bin/issue_mix > gpurun_out/${TAG}_issue_mix.jsonl 2>&1
python bench.py --workload c3 --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_c3.json 2> gpurun_out/${TAG}_bench_c3.err
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_c4.json 2> gpurun_out/${TAG}_bench_c4.err
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
python tools/sweep_time.py 20000 5000 10 10 > gpurun_out/${TAG}_sweep_time_c4.json 2>&1
for WL in c3 c4; do
  CMD="python bench.py --workload $WL --steps 2 --warmup 3 --no-e2e --no-cpu"
  $CMD > gpurun_out/plain_$WL.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 400 --csv \
      --log-file gpurun_out/${TAG}_launches_$WL.csv $CMD > gpurun_out/ncu_l_$WL.log 2>&1
done
CMD="python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/plain2_c4.log 2>&1 && ncu --set full --clock-control none -k 'regex:sweep3' -s 2 -c 1 \
    -o gpurun_out/${TAG}_sweep_c4 $CMD > gpurun_out/ncu_f_c4.log 2>&1
$CMD > gpurun_out/plain3_c4.log 2>&1 && ncu --set full --clock-control none --kernel-name-base demangled -k 'regex:sweep3_kernel<5, *1>' -c 1 \
    -o gpurun_out/${TAG}_sweep_lb_c4 $CMD > gpurun_out/ncu_h_c4.log 2>&1
for R in gpurun_out/${TAG}_*.ncu-rep; do python tools/ncu_summary.py $R > ${R%.ncu-rep}.txt 2>/dev/null; done
while [ $(du -sm gpurun_out | cut -f1) -gt 55 ]; do rm -f $(ls -S gpurun_out/*.ncu-rep | head -1); done
ls -la gpurun_out | tail -24
