for i in 1 2; do
for v in 0 1; do
FGDM_PDL=$v timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --c3-prompts 0 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('FGDM_PDL=$v', 'value', round(d['value'], 4), 'e2e', round(d['e2e']['value'], 4), 'sm_mhz', d['clocks']['sm_mhz'], 'checksum', d['e2e']['image_checksum'])
"
done
done
