#!/usr/bin/env bash
# round-2 first GPU pass: whole GPU suite (no -x: collect every failure), default bench, launch list of the bench
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02a_gputests.log 2>&1
tail -40 gpurun_out/r02a_gputests.log
timeout 400 python bench.py > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err
tail -c 3500 gpurun_out/r02a_bench.json; tail -5 gpurun_out/r02a_bench.err
