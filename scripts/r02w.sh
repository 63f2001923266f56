# device FASTQ tokeniser: its tests, the FASTA tokeniser's (unchanged path), the CLI tests that take FASTQ
mkdir -p gpurun_out
python -m pytest tests/test_gpu_fastq.py tests/test_gpu_fasta.py -x -q -m gpu > gpurun_out/r02w_fastq.log 2>&1; echo "rc=$?" >> gpurun_out/r02w_fastq.log
python -m pytest tests/test_dropin_cli.py -x -q -m gpu -k "fastq or raw_fasta" > gpurun_out/r02w_cli.log 2>&1; echo "rc=$?" >> gpurun_out/r02w_cli.log
tail -n 30 gpurun_out/r02w_fastq.log; tail -n 30 gpurun_out/r02w_cli.log
