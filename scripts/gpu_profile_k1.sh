# ncu evidence for K1 (run under gpurun; one ncu "use" per call): launch list + one full capture of the persistent kernel
mkdir -p gpurun_out
CMD="python scripts/k1_once.py 24"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_k1.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:impala_loss -s 10 -c 2 -o gpurun_out/prof_k1 -f $CMD > gpurun_out/ncu2.log 2>&1
tail -n 3 gpurun_out/plain.log gpurun_out/ncu1.log gpurun_out/ncu2.log
grep impala_loss gpurun_out/launches_k1.csv | tail -8
