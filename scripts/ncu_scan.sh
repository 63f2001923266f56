# usage: bash scripts/ncu_scan.sh <tag>   (ncu --set full capture of the scan kernel on a 400-genome bench run)
TAG=${1:-x}
mkdir -p gpurun_out
CMD="python bench.py --genomes 400 --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:kssd_scan_kernel -s 3 -c 1 -o gpurun_out/scan_$TAG $CMD > gpurun_out/ncu_scan_$TAG.log 2>&1
echo "ncu rc=$?"
tail -c 600 gpurun_out/plain_$TAG.log
