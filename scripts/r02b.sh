#!/usr/bin/env bash
# round-2 second GPU pass: the tests that failed in r02a, the reference arm, the other BASELINE configs
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fasta.py tests/test_dropin_cli.py -m gpu -q > gpurun_out/r02b_tests.log 2>&1
tail -15 gpurun_out/r02b_tests.log
timeout 600 python bench.py --impl reference > gpurun_out/r02b_ref.json 2> gpurun_out/r02b_ref.err; tail -c 1500 gpurun_out/r02b_ref.json; tail -3 gpurun_out/r02b_ref.err
for C in C3 C5 C4; do
  timeout 900 python bench.py --config $C > gpurun_out/r02b_$C.json 2> gpurun_out/r02b_$C.err
  echo "== $C rc=$?"; tail -c 2500 gpurun_out/r02b_$C.json; tail -5 gpurun_out/r02b_$C.err
done
