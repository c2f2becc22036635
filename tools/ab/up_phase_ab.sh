# A/B of the sub-pixel Upsample on one box: FGDM_UP_PHASE=0/1 and the row threshold FGDM_UP_PHASE_MIN_ROWS
run() {
env $1 timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --c3-prompts 0 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', 'value', round(d['value'], 4), 'e2e', round(d['e2e']['value'], 4), 'sm_mhz', d['clocks']['sm_mhz'], 'checksum', d['e2e']['image_checksum'], 'exec TF/img', round(d['roofline_step']['executed_tflop_per_image'], 2))
"
}
for i in 1 2; do
run FGDM_UP_PHASE_MIN_ROWS=4096
run FGDM_UP_PHASE_MIN_ROWS=1024
done
