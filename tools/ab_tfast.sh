run() { timeout 200 python bench.py --steps 20 --warmup 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1','value',round(d['value'],1),'e2e',round(d['e2e']['value'],1))
kb=d.get('kernel_breakdown') or d.get('breakdown') or {}
print('   ',{k:round(v['ms_per_step'],3) for k,v in kb.items()})
"; }
VFI_T_FAST=0 run tfast0
VFI_T_FAST=1 run tfast1
