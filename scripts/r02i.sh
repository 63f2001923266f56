#!/usr/bin/env bash
timeout 900 python -m pytest tests/test_gpu_dist.py tests/test_dropin_cli.py -m gpu -q -k "set_union or union_and_sub or peer or packed or restricted" > gpurun_out/r02i_tests.log 2>&1; tail -5 gpurun_out/r02i_tests.log
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/r02i_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"genome_|bucket_|dist_tile|row_work|segment_scan|ref_sizes" -s 36 -c 24 --csv --log-file gpurun_out/r02i_launches.csv $CMD > gpurun_out/r02i_ncu_list.log 2>&1
echo "launch list rc=$?"
