set -x
mkdir -p gpurun_out
KRE='regex:kssd_|radix_|unique_|scan_tiles|split_hash|genome_off|make_pairs|ref_sizes|head_|set_u64|list_lookup|dist_tile'
CMD="python bench.py --genomes 400 --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/plain_400.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" -c 600 --csv --log-file gpurun_out/launches_r1a.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:kssd_scan_kernel -s 3 -c 1 -o gpurun_out/scan_r1a $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:dist_tile_kernel -s 3 -c 1 -o gpurun_out/dist_r1a $CMD > gpurun_out/ncu_full2.log 2>&1
echo "full2 rc=$?"
tail -3 gpurun_out/plain_400.log
