#!/usr/bin/env bash
timeout 1200 python -m pytest tests/test_gpu_dist.py tests/test_gpu_parity.py tests/test_dropin_cli.py tests/test_gpu_fullsize.py -m gpu -q > gpurun_out/r02f_tests.log 2>&1; tail -4 gpurun_out/r02f_tests.log
timeout 600 python scripts/cli_trace.py 1000 > gpurun_out/r02f_cli_trace.txt 2>&1; grep -E "==|contexts|read / parse" gpurun_out/r02f_cli_trace.txt
timeout 300 python bench.py --steps 5 --no-cpu-baseline > gpurun_out/r02f_bench.json 2> gpurun_out/r02f_bench.err; tail -c 600 gpurun_out/r02f_bench.json
