#!/bin/bash
# Round-2 measurement pass on one B200 (run under gpurun from the repo root): GPU tests, kernel ladder,
# ncu --set full captures of the kernels the bench launches (exported to CSV: the reports themselves are
# too large to bring back), the bench's launch list, and the bench lines of the three architectures.
# SKIP_KERNELS=1 leaves out the kernel ladder, the ncu captures and the launch list (tests and bench lines only).
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 > $O/gpu_tests.log 2>&1; echo "gpu tests rc=$?" | tee -a $O/gpu_tests.log
if [ -z "$SKIP_KERNELS" ]; then
timeout 900 python tools/kbench.py --r2 --reps 10 > $O/r2_kbench_ladder.jsonl 2> $O/kbench.err; echo "kbench rc=$?"
P="python tools/kbench.py --only-rows 25600 --reps 2"
for spec in "4096 add auto" "3584 tail auto" "4096 plain auto" "5120 add auto" "4096 add tma"; do
  set -- $spec
  $P --only-h $1 --only-form $2 --impl $3 > $O/plain_$1_$2_$3.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:k2_ -s 3 -c 1 -o /tmp/r2_k2_$1_$2_$3 $P --only-h $1 --only-form $2 --impl $3 > $O/ncu_$1_$2_$3.log 2>&1
  ncu -i /tmp/r2_k2_$1_$2_$3.ncu-rep --page raw --csv > $O/r2_k2_$1_$2_$3_full_raw.csv 2>/dev/null
  ncu -i /tmp/r2_k2_$1_$2_$3.ncu-rep --page source --csv 2>/dev/null | gzip > $O/r2_k2_$1_$2_$3_source.csv.gz
done
B="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --ncu-window"
$B > $O/bench_window_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 3000 --csv --log-file $O/r2_bench_launches_window.csv $B > $O/bench_window_ncu.log 2>&1
echo "launch list rc=$?"
fi
timeout 900 python bench.py > $O/r2_bench_n1.json 2> $O/bench_n1.err; echo "bench n1 rc=$?"
timeout 900 python bench.py --arch gemma-2-9b --no-extras --no-cpu-baseline > $O/r2_bench_gemma2_9b.json 2> $O/bench_gemma.err; echo "bench gemma rc=$?"
timeout 900 python bench.py --arch qwen2.5-32b --no-extras --no-cpu-baseline > $O/r2_bench_qwen25_32b.json 2> $O/bench_qwen.err; echo "bench qwen rc=$?"
du -sh $O
