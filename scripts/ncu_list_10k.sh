TAG=${1:-x}
KRE='regex:kssd_|radix_|unique_|scan_tiles|split_hash|genome_off|make_pairs|ref_sizes|head_|set_u64|list_lookup|dist_tile'
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/plain_10k_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" -s 600 -c 220 --csv --log-file gpurun_out/launches_10k_$TAG.csv $CMD > gpurun_out/ncu_list_10k_$TAG.log 2>&1
echo "rc=$?"; tail -c 400 gpurun_out/plain_10k_$TAG.log
