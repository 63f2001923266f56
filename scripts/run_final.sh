# round-end style check on one GPU: full GPU test suite, smoke, default bench (with cpu baseline), reference arm
TAG=${1:-w}
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
( time python -c 'import __graft_entry__ as g; g.smoke()' ) > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; grep "smoke ok" gpurun_out/${TAG}_smoke.log
( time python bench.py ) > gpurun_out/${TAG}_bench.log 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/${TAG}_bench.err
( time python bench.py --impl reference --steps 2 --warmup 1 ) > gpurun_out/${TAG}_ref.log 2> gpurun_out/${TAG}_ref.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/${TAG}_ref.log
