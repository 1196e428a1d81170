show() { python bench.py --steps 3 --warmup 3 --no-cpu --batch ${BATCH:-60} 2>&1 | grep -E "b200s|metric" | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('[b200s]'): print(l.strip()); continue
    d=json.loads(l); s=d['stage_ms_per_step']; n=d['config']['pairs_per_step_per_gpu']; print({k: round(1e3*v/n,1) for k,v in s.items() if k in ('cost','paths','wta')}, 'fps', round(d['value']))"; }
for vt in 512 288; do for b in 60 72 90; do echo "VT=$vt B=$b"; B200S_VERBOSE=1 B200S_VERT_VT=$vt BATCH=$b show; done; done
echo auto; B200S_VERBOSE=1 BATCH=0 show
