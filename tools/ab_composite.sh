for i in 1 2; do
NB_COMPOSITE_LC=0 python bench.py --no-cpu-baseline --steps 100 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('LC=0', d['value'], d['ms_per_step'], d['render']['value'])"
NB_COMPOSITE_LC=1 python bench.py --no-cpu-baseline --steps 100 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('LC=1', d['value'], d['ms_per_step'], d['render']['value'])"
done
