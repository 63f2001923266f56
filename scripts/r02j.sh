#!/usr/bin/env bash
# tile height experiment (KSSD_B200_TILE_QMAX) on the default bench
for Q in 4 8 16 32; do
  KSSD_B200_TILE_QMAX=$Q timeout 300 python bench.py --steps 10 --e2e-steps 1 --no-cpu-baseline > gpurun_out/r02j_q$Q.json 2> gpurun_out/r02j_q$Q.err
  python -c "
import json; d=json.loads(open('gpurun_out/r02j_q$Q.json').read()); print('QMAX $Q step', round(d['ms_per_step'],4), 'alldist region ms', round(d['alldist_kernel']['ms_per_launch'],4))"
done
timeout 600 python -m pytest tests/test_dropin_cli.py -m gpu -q -k "union_and_sub" 2>&1 | tail -2
