#!/usr/bin/env bash
KSSD_B200_DEVICE=0 timeout 600 python scripts/cli_trace.py 1000 > gpurun_out/r02m_cli_trace.txt 2>&1; grep -vE "set_params|sketch_t|saveSketches" gpurun_out/r02m_cli_trace.txt
