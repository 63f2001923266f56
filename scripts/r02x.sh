# wall clock of the FASTQ sketch path of the CLI (raw text on the device vs kseq on the host vs the reference)
mkdir -p gpurun_out
python scripts/cli_fastq.py --files 32 > gpurun_out/r02y_cli_fastq.json 2> gpurun_out/r02y_cli_fastq.err; echo "rc=$?"
cat gpurun_out/r02y_cli_fastq.json; tail -n 5 gpurun_out/r02y_cli_fastq.err
