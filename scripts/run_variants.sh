# usage: bash scripts/run_variants.sh <tag>   -- times the scan kernel with every library under scratch_variants/ and the default one
TAG=${1:-v}
mkdir -p gpurun_out
cp rabbitkssd_b200/libkssd_b200.so /tmp/libkssd_default.so
python ${SCRIPT:-scripts/scan_variants.py} --tag default > gpurun_out/${TAG}_variants.log 2>&1
for f in scratch_variants/*.so; do
  cp $f rabbitkssd_b200/libkssd_b200.so
  python ${SCRIPT:-scripts/scan_variants.py} --tag $(basename $f .so) >> gpurun_out/${TAG}_variants.log 2>&1
done
cp /tmp/libkssd_default.so rabbitkssd_b200/libkssd_b200.so
python ${SCRIPT:-scripts/scan_variants.py} --tag default_again >> gpurun_out/${TAG}_variants.log 2>&1
grep '^{' gpurun_out/${TAG}_variants.log
