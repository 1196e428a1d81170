# GPU box: stage times of the default engine and of every tuning build under variants/
show() { python bench.py --steps 3 --warmup 3 --no-cpu --batch ${BATCH:-60} 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']; n=d['config']['pairs_per_step_per_gpu']; print({k: round(1e3*v/n,1) for k,v in s.items() if k in (${STAGES:-'cost','paths','wta'})}, 'fps', round(d['value']))"; }
echo -n "default: "; show
for f in variants/libb200stereo_*.so; do echo -n "$f: "; B200S_LIB=$PWD/$f show; done
