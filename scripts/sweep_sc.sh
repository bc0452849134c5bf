for sc in 1 2 4; do
  echo "== SC=$sc"
  MDT_LATENT_SC=$sc python bench.py --steps 2000 --warmup 50 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(d['value'], d['ms_per_step'], d['e2e']['value'])
pk=d.get('profile',{}) if isinstance(d.get('profile'),dict) else {}
"
done
MDT_LATENT_SC=2 GRAPH=1 timeout 200 python scripts/tc05_phases.py tf32x3 2>&1 | grep -E "go \+" | cut -c1-62
