mkdir -p gpurun_out/final
python bench.py > gpurun_out/final/r2_bench_gpt2.json 2> gpurun_out/final/gpt2.err; echo "gpt2 rc=$?"
python bench.py --dropout 0.1 --no-cpu-baseline > gpurun_out/final/r2_bench_gpt2_dropout.json 2> gpurun_out/final/drop.err; echo "drop rc=$?"
python bench.py --no-cpu-baseline > gpurun_out/final/r2_bench_gpt2_nodropout_same_box.json 2> gpurun_out/final/nodrop.err; echo "nodrop rc=$?"
for m in gpt-nano gpt-mini gpt2-medium gpt2-xl; do
  python bench.py --model $m --no-cpu-baseline > gpurun_out/final/r2_bench_$m.json 2> gpurun_out/final/$m.err; echo "$m rc=$?"
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final/r2_bench_reference_arm.json 2> gpurun_out/final/ref.err; echo "ref rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/final/*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], round(d.get('ms_per_step',0),3), round(d.get('value',0)), (d.get('e2e') or {}).get('value'), (d.get('roofline') or {}).get('frac'), (d.get('clocks') or {}).get('sm_mhz'), d.get('gpu_launches'), (d.get('extra') or {}).get('mfu'))
    except Exception as e:
        print(f, 'ERR', e)
PY
