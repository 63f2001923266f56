#!/usr/bin/env bash
# final 1-GPU pass: whole GPU suite, smoke, default bench, fresh ncu --set full of the G=3 scan kernel
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r02t_tests.log 2>&1; tail -4 gpurun_out/r02t_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/r02t_bench.json 2> gpurun_out/r02t_bench.err; tail -c 500 gpurun_out/r02t_bench.json; tail -2 gpurun_out/r02t_bench.err
CMD="python bench.py --genomes 2000 --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/r02t_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:kssd_scan_kernel -s 12 -c 1 -f -o gpurun_out/r02t_scan_g3 $CMD > gpurun_out/r02t_ncu_scan.log 2>&1
echo "scan ncu rc=$?"
