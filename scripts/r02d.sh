#!/usr/bin/env bash
# CLI phase trace on one GPU + launch list of my kernels in the default bench
mkdir -p /dev/shm/kt && cd /dev/shm/kt && python - <<'PY'
import sys; sys.path[:0]=['/root/repo','/root/repo/oracle']
import oracle, os
os.makedirs('fa', exist_ok=True)
oracle.synth_fasta_files('fa', 20261106, 0, 1000, 10, 5000000, 0.04, 2)
open('list','w').write(''.join(f'fa/g{i}.fna\n' for i in range(1000)))
open('list100','w').write(''.join(f'fa/g{i}.fna\n' for i in range(100)))
PY
/root/repo/oracle/_ref/rabbit_kssd shuffle -k 10 -s 6 -l 3 -o L3K10.shuf > /dev/null 2>&1
for L in list100 list list; do
  for MODE in raw kseq; do
    rm -f $L.sketch*; E=""; [ $MODE = kseq ] && E="KSSD_B200_NO_RAW_FASTA=1"
    /usr/bin/time -f "$L $MODE wall %e s" env KSSD_B200_TRACE=1 $E /root/repo/rabbitkssd_b200/_dropin/rabbit_kssd_b200 sketch -q -L L3K10.shuf -i $L -o $L -t 16 2>&1 | grep -E "trace|wall|time of"
  done
done
cd /root/repo; rm -rf /dev/shm/kt
CMD="python bench.py --steps 1 --warmup 3 --e2e-steps 1 --no-cpu-baseline"
$CMD > gpurun_out/r02d_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"kssd_|genome_|bucket_|dist_tile|row_work|segment_scan|ref_sizes|filter_set" -s 186 -c 130 --csv --log-file gpurun_out/r02d_launches.csv $CMD > gpurun_out/r02d_ncu_list.log 2>&1
echo "launch list rc=$?"
