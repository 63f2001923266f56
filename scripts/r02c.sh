#!/usr/bin/env bash
# round-2 third GPU pass: CLI wall clock, launch list of the default bench, ncu --set full of the changed kernels
timeout 900 python scripts/cli_e2e.py --genomes 1000 > gpurun_out/r02c_cli_e2e_1000.json 2> gpurun_out/r02c_cli.err; tail -c 1800 gpurun_out/r02c_cli_e2e_1000.json; tail -3 gpurun_out/r02c_cli.err
timeout 300 python scripts/cli_e2e.py --genomes 100 > gpurun_out/r02c_cli_e2e_100.json 2>> gpurun_out/r02c_cli.err; tail -c 1800 gpurun_out/r02c_cli_e2e_100.json
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/r02c_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02c_launches.csv $CMD > gpurun_out/r02c_ncu_list.log 2>&1
echo "launch list rc=$?"
for K in dist_tile_kernel row_work_kernel bucket_sort_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/r02c_$K $CMD > gpurun_out/r02c_ncu_$K.log 2>&1
  echo "$K rc=$?"
done
C4="python bench.py --config C4 --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline --extra genomes=4,sample_len=0"
$C4 > gpurun_out/r02c_plain_c4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:kssd_scan_kernel -s 8 -c 1 -f -o gpurun_out/r02c_scan_g4 $C4 > gpurun_out/r02c_ncu_scan_g4.log 2>&1
echo "scan g4 rc=$?"; tail -c 600 gpurun_out/r02c_plain_c4.log
