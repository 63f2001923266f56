# one gpurun call: GPU tests, 10k bench, scan prefetch variants (timed + dram bytes under ncu), launch list
TAG=${1:-m}
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
( time python bench.py --steps 5 --warmup 3 --e2e-steps 1 --no-cpu-baseline --phases ) > gpurun_out/${TAG}_bench.log 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python scripts/scan_variants.py > gpurun_out/${TAG}_variants.log 2>&1; echo "variants rc=$?"; cat gpurun_out/${TAG}_variants.log
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:kssd_scan --csv --log-file gpurun_out/${TAG}_variants_ncu.csv \
  python scripts/scan_variants.py --genomes 430 --steps 1 --warmup 0 > gpurun_out/${TAG}_variants_ncu.log 2>&1; echo "ncu variants rc=$?"
bash scripts/ncu_list_10k.sh $TAG
