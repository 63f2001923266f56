# one gpurun call: parity tests, 10k bench, ncu full capture of the scan kernel, launch list
TAG=${1:-n}
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_parity.py -m gpu -x -q ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
( time python bench.py --steps 5 --warmup 3 --e2e-steps 1 --no-cpu-baseline --phases ) > gpurun_out/${TAG}_bench.log 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; cat gpurun_out/${TAG}_bench.err | head -3
GENOMES=2000 bash scripts/ncu_kernels.sh $TAG kssd_scan_kernel
bash scripts/ncu_list_10k.sh $TAG
