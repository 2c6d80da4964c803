cd /root/repo
for p in 1 2 3 4; do
echo "== VSLAMB200_PARTS=$p"
VSLAMB200_PARTS=$p python bench.py --no-cpu --steps 4 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',round(d['value']),'ms/step',round(d['ms_per_step'],2),'e2e',round(d['e2e']['value']))"
done
