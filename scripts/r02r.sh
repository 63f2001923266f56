#!/usr/bin/env bash
timeout 900 python -m pytest tests/test_gpu_dist.py tests/test_gpu_fullsize.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r02r_tests.log 2>&1; tail -3 gpurun_out/r02r_tests.log
timeout 300 python bench.py --config C3 > gpurun_out/r02r_C3.json 2> gpurun_out/r02r_C3.err
python -c "
import json; d=json.loads(open('gpurun_out/r02r_C3.json').read()); print('C3 ms', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['parity'].get('set_equal'))"
timeout 300 python bench.py --steps 10 --e2e-steps 1 --no-cpu-baseline > gpurun_out/r02r_bench.json 2> gpurun_out/r02r_bench.err
python -c "
import json; d=json.loads(open('gpurun_out/r02r_bench.json').read()); print('C2 step', d['ms_per_step'], 'frac', d['roofline']['frac'], 'alldist ms', d['alldist_kernel']['ms_per_launch'], d['parity']['sha256'][:12])"
