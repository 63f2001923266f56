#!/usr/bin/env bash
# usage: bash scripts/r02_multi.sh N   (under gpurun --gpus N): multi-GPU tests (N >= 2), default bench with phases, C3 / C5, CLI on N GPUs
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "2" ]; then
  timeout 900 python -m pytest tests/test_dropin_cli.py tests/test_gpu_dist.py -m gpu -q -k "multi_gpu or peer or replicated or restricted" > gpurun_out/r02m${N}_tests.log 2>&1
  tail -5 gpurun_out/r02m${N}_tests.log
fi
timeout 600 $TR --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --phases > gpurun_out/r02m${N}_bench.json 2> gpurun_out/r02m${N}_bench.err
echo "== bench rc=$?"; tail -c 2600 gpurun_out/r02m${N}_bench.json; grep -E "phases rank|trace rank|Error|error" gpurun_out/r02m${N}_bench.err | head -24
timeout 600 $TR --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --index-replicated --no-cpu-baseline > gpurun_out/r02m${N}_bench_replicated.json 2> gpurun_out/r02m${N}_bench_replicated.err
echo "== bench replicated rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r02m${N}_bench_replicated.json').read()); print('replicated index: ms_per_step', d['ms_per_step'], 'parity', d['parity'].get('set_equal'))"
for C in C3 C5; do
  timeout 900 $TR --master-port 29513 bench.py --gpus $N --config $C --steps 5 --warmup 3 > gpurun_out/r02m${N}_$C.json 2> gpurun_out/r02m${N}_$C.err
  echo "== $C rc=$?"; tail -c 1800 gpurun_out/r02m${N}_$C.json; tail -3 gpurun_out/r02m${N}_$C.err
done
timeout 600 python scripts/cli_e2e.py --genomes 1000 --repeat 1 > gpurun_out/r02m${N}_cli_e2e_1000.json 2> gpurun_out/r02m${N}_cli.err; tail -c 1500 gpurun_out/r02m${N}_cli_e2e_1000.json; tail -3 gpurun_out/r02m${N}_cli.err
