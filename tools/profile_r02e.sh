#!/bin/bash
# Last evidence of the round (r02e): the committed build after the chunk-list change -- tests, bench lines, launch list,
# one full capture of the bound sweep.
TAG=r02e
KF='regex:sweep3|contract_|resp_kernel|solve_kernel|scalars_kernel|lb_kernel|lb_finish|fill_kernel|init_params'
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_gpu_tests.log 2>&1
python bench.py --workload c3 --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_c3.json 2> gpurun_out/${TAG}_bench_c3.err
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_c4.json 2> gpurun_out/${TAG}_bench_c4.err
CMD="python bench.py --workload c4 --steps 2 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/plain_c4.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k "$KF" -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches_c4.csv $CMD > gpurun_out/ncu_l_c4.log 2>&1
$CMD > gpurun_out/plain3_c4.log 2>&1 && ncu --set full --clock-control none --kernel-name-base demangled \
    -k 'regex:sweep3_kernel<\(int\)5, \(bool\)1>' -c 1 -o gpurun_out/${TAG}_sweep_lb_c4 $CMD > gpurun_out/ncu_h_c4.log 2>&1
python tools/ncu_summary.py gpurun_out/${TAG}_sweep_lb_c4.ncu-rep > gpurun_out/${TAG}_sweep_lb_c4.txt 2>/dev/null
python tools/ncu_lines.py gpurun_out/${TAG}_sweep_lb_c4.ncu-rep sweep3_kernelILi5ELb1 30 > gpurun_out/${TAG}_sweep_lb_c4_lines.txt 2>/dev/null
rm -f gpurun_out/${TAG}_sweep_lb_c4.ncu-rep
tail -2 gpurun_out/${TAG}_gpu_tests.log; head -12 gpurun_out/${TAG}_sweep_lb_c4.txt; tail -3 gpurun_out/ncu_h_c4.log
