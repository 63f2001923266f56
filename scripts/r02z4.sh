# tokeniser scatter kernels through shared memory: tokeniser tests, every CLI test (FASTA inputs take the raw path), launch list
mkdir -p gpurun_out
python -m pytest tests/test_gpu_fastq.py tests/test_gpu_fasta.py -x -q -m gpu > gpurun_out/r02z4_tok.log 2>&1; echo "tok rc=$?"; tail -n 3 gpurun_out/r02z4_tok.log
python -m pytest tests/test_dropin_cli.py -x -q -m gpu > gpurun_out/r02z4_cli.log 2>&1; echo "cli rc=$?"; tail -n 3 gpurun_out/r02z4_cli.log
python scripts/fastq_device.py > gpurun_out/r02z4_fastq_device.json 2> gpurun_out/r02z4_fastq_device.err; echo "rc=$?"; cat gpurun_out/r02z4_fastq_device.json
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'fq_|kssd_scan|segment_scan' -c 400 --csv --log-file gpurun_out/r02z4_fastq_launches.csv python scripts/fastq_device.py > gpurun_out/r02z4_ncu.log 2>&1; echo "ncu rc=$?"
