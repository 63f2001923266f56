#!/usr/bin/env bash
timeout 600 python bench.py --config C4 --extra shuffle=L5K10 > gpurun_out/r02p_C4_L5K10.json 2> gpurun_out/r02p_C4_L5K10.err; echo "L5K10 rc=$?"; tail -c 1500 gpurun_out/r02p_C4_L5K10.json; tail -3 gpurun_out/r02p_C4_L5K10.err
timeout 300 python bench.py --config C1 > gpurun_out/r02p_C1.json 2> gpurun_out/r02p_C1.err; echo "C1 rc=$?"; tail -c 1200 gpurun_out/r02p_C1.json; tail -3 gpurun_out/r02p_C1.err
