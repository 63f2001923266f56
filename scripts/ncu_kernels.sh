# usage: bash scripts/ncu_kernels.sh <tag> <kernel-regex>...   (one ncu --set full capture per kernel on a 2000-genome bench run)
TAG=$1; shift
mkdir -p gpurun_out
CMD="python bench.py --genomes ${GENOMES:-2000} --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.log; exit 1; }
for K in "$@"; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/${K}_$TAG $CMD > gpurun_out/ncu_${K}_$TAG.log 2>&1
  echo "$K rc=$?"
done
