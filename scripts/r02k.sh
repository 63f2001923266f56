#!/usr/bin/env bash
# final 8-GPU pass: default bench (peer flag barrier, restricted index) with phases, A/B with the NCCL barrier, C3, C5
N=8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --phases > gpurun_out/r02k_bench.json 2> gpurun_out/r02k_bench.err
echo "== bench rc=$?"; tail -c 2200 gpurun_out/r02k_bench.json; grep -E "phases rank|rror" gpurun_out/r02k_bench.err | head -12
timeout 600 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --peer-nccl-barrier --no-cpu-baseline > gpurun_out/r02k_bench_ncclbarrier.json 2> gpurun_out/r02k_bench_ncclbarrier.err
echo "== nccl barrier rc=$?"; python -c "
import json; t=open('gpurun_out/r02k_bench_ncclbarrier.json').read(); d=json.loads(t[t.index('{'):]); print('nccl barrier: ms_per_step', d['ms_per_step'], d['parity'].get('set_equal'))"
for NN in 2 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NN --master-addr 127.0.0.1 --master-port 2952$NN bench.py --gpus $NN --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02k_bench_n$NN.json 2> gpurun_out/r02k_bench_n$NN.err
  python -c "
import json; t=open('gpurun_out/r02k_bench_n$NN.json').read(); d=json.loads(t[t.index('{'):]); print('N=$NN: ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['parity'].get('set_equal'))"
done
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02k_bench_n1.json 2> gpurun_out/r02k_bench_n1.err
python -c "
import json; d=json.loads(open('gpurun_out/r02k_bench_n1.json').read()); print('N=1: ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])"
for C in C3 C5; do
  timeout 900 $TR --master-port 29513 bench.py --gpus $N --config $C --steps 5 --warmup 3 > gpurun_out/r02k_$C.json 2> gpurun_out/r02k_$C.err
  echo "== $C rc=$?"; tail -c 700 gpurun_out/r02k_$C.json; tail -2 gpurun_out/r02k_$C.err
done
