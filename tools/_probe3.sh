mkdir -p gpurun_out; rm -f gpurun_out/probe4.jsonl
K="python tools/kbench.py --reps 10 --impl stream"
for c in 4 7; do for f in plain add tail; do
  echo "chunks=$c" >> gpurun_out/probe4.jsonl
  IA_K2_STREAM_CHUNKS=$c $K --only-rows 25600 --only-h 3584 --only-form $f >> gpurun_out/probe4.jsonl 2>> gpurun_out/probe4.err
done; done
for c in 5 10; do for f in plain add; do
  echo "chunks=$c" >> gpurun_out/probe4.jsonl
  IA_K2_STREAM_CHUNKS=$c $K --only-rows 25600 --only-h 5120 --only-form $f >> gpurun_out/probe4.jsonl 2>> gpurun_out/probe4.err
done; done
for r in 8192 262144; do for f in plain add; do
  echo "chunks=8" >> gpurun_out/probe4.jsonl
  $K --only-rows $r --only-h 4096 --only-form $f >> gpurun_out/probe4.jsonl 2>> gpurun_out/probe4.err
done; done
python - <<'PY'
import json
for l in open('gpurun_out/probe4.jsonl'):
    if l.startswith('chunks'): print(l.strip(), end=' '); continue
    d=json.loads(l); print(d['kernel'],d['impl'],d['rows'],d['H'],round(d['us'],1),round(d.get('dram_gbs',0)))
PY
