# 2-GPU sanity of the final tree: the driver's own launch line for the default bench (no cpu baseline leg: rank 0 alone would run it)
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02z2_bench.json 2> gpurun_out/r02z2_bench.err
echo "rc=$?"; tail -c 3000 gpurun_out/r02z2_bench.json; tail -n 3 gpurun_out/r02z2_bench.err
