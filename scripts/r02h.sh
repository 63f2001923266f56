#!/usr/bin/env bash
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r02h_tests.log 2>&1; tail -6 gpurun_out/r02h_tests.log
timeout 400 python bench.py > gpurun_out/r02h_bench.json 2> gpurun_out/r02h_bench.err; tail -c 900 gpurun_out/r02h_bench.json; tail -3 gpurun_out/r02h_bench.err
timeout 300 python bench.py --config C5 --no-cpu-baseline > gpurun_out/r02h_C5.json 2> gpurun_out/r02h_C5.err; tail -c 700 gpurun_out/r02h_C5.json
