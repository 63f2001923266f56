#!/usr/bin/env bash
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_seams.py tests/test_gpu_fasta.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/r02q_tests.log 2>&1; tail -3 gpurun_out/r02q_tests.log
timeout 300 python bench.py --steps 10 --e2e-steps 1 --no-cpu-baseline > gpurun_out/r02q_bench.json 2> gpurun_out/r02q_bench.err
python -c "
import json; d=json.loads(open('gpurun_out/r02q_bench.json').read()); print('step', d['ms_per_step'], 'frac', d['roofline']['frac'], 'scan ms/launch', d['roofline']['ms_per_launch'], 'hits', d['hits'], d['parity']['sha256'][:12])"
timeout 300 python bench.py --config C4 --steps 5 --e2e-steps 1 --no-cpu-baseline --extra genomes=6,sample_len=0 > gpurun_out/r02q_c4.json 2> gpurun_out/r02q_c4.err
python -c "
import json; d=json.loads(open('gpurun_out/r02q_c4.json').read()); print('C4(6 genomes) frac', d['roofline']['frac'], d['value'])"
