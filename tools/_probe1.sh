set -x
mkdir -p gpurun_out
K="python tools/kbench.py --only-rows 25600 --reps 10"
for h in 3584 4096 2048; do for f in plain add; do for i in stream tma; do
  $K --only-h $h --only-form $f --impl $i >> gpurun_out/probe1.jsonl 2>> gpurun_out/probe1.err
done; done; done
P="python tools/kbench.py --only-rows 25600 --reps 2"
for spec in "4096 add stream" "3584 add stream" "4096 plain stream" "4096 plain tma" "4096 add tma" "3584 tail stream"; do
  set -- $spec
  $P --only-h $1 --only-form $2 --impl $3 > gpurun_out/plain_$1_$2_$3.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:k2_ -s 3 -c 1 -o /tmp/r2_k2_$1_$2_$3 $P --only-h $1 --only-form $2 --impl $3 > gpurun_out/ncu_$1_$2_$3.log 2>&1
  ncu -i /tmp/r2_k2_$1_$2_$3.ncu-rep --page raw --csv > gpurun_out/r2_k2_$1_$2_$3_raw.csv 2>/dev/null
  ncu -i /tmp/r2_k2_$1_$2_$3.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/r2_k2_$1_$2_$3_source.csv.gz
done
du -sh gpurun_out; ls -la gpurun_out
