# per-kernel times of the FASTQ tokeniser (launch list of one raw FASTQ session; the plain run first)
mkdir -p gpurun_out
python scripts/fastq_device.py > gpurun_out/r02z3_fastq_device.json 2> gpurun_out/r02z3_fastq_device.err; echo "rc=$?"; cat gpurun_out/r02z3_fastq_device.json
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'fq_|kssd_scan|segment_scan' -c 400 --csv --log-file gpurun_out/r02z3_fastq_launches.csv python scripts/fastq_device.py > gpurun_out/r02z3_ncu.log 2>&1; echo "ncu rc=$?"
tail -n 3 gpurun_out/r02z3_fastq_launches.csv
