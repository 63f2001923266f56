#!/usr/bin/env bash
KSSD_B200_DEVICE=0 timeout 600 python scripts/cli_trace.py 1000 > gpurun_out/r02e_cli_trace.txt 2>&1; cat gpurun_out/r02e_cli_trace.txt
bash scripts/r02_multi.sh 2
