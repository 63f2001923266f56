# usage: bash scripts/ncu_list_10k.sh <tag> [skip] [count]  -- launch list (gpu__time_duration) of one timed step of the 10k-genome bench
TAG=${1:-x}
SKIP=${2:-186}
COUNT=${3:-70}
KRE='regex:kssd_|radix_|unique_|scan_tiles|split_hash|genome_|bucket_|segment_|make_pairs|ref_sizes|head_|set_u64|list_lookup|dist_tile'
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/plain_10k_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" -s $SKIP -c $COUNT --csv --log-file gpurun_out/launches_10k_$TAG.csv $CMD > gpurun_out/ncu_list_10k_$TAG.log 2>&1
echo "rc=$?"; tail -c 400 gpurun_out/plain_10k_$TAG.log
