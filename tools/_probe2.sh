set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_k2_gpu.py tests/test_k2_stream_gpu.py -x -q > gpurun_out/t_k2.log 2>&1; echo "k2 tests rc=$?" >> gpurun_out/t_k2.log; tail -3 gpurun_out/t_k2.log
K="python tools/kbench.py --reps 10"
rm -f gpurun_out/probe2.jsonl
for rows in 25600 262144; do for h in 4096 5120 3584; do for f in plain add tail; do for i in tma stream; do
  if [ $f = tail ] && [ $h != 3584 ]; then continue; fi
  if [ $i = stream ] && [ $h != 3584 ]; then continue; fi
  $K --only-rows $rows --only-h $h --only-form $f --impl $i >> gpurun_out/probe2.jsonl 2>> gpurun_out/probe2.err
done; done; done; done
cat gpurun_out/probe2.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['kernel'],d['impl'],d['rows'],d['H'],round(d['us'],1),round(d.get('dram_gbs',0)))"
