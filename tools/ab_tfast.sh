#!/bin/bash
# A/B of the tile orders of the gathers / blend (vfi_set_option "t_fast"): bash tools/ab_tfast.sh [values...]
run() { timeout 200 python bench.py --steps 20 --warmup 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('t_fast=$1','value',round(d['value'],1),'e2e',round(d['e2e']['value'],1))
print('   ',{k:round(v['ms_per_step'],3) for k,v in (d.get('kernel_breakdown') or {}).items()})
"; }
for v in ${@:-0 1 2}; do VFI_T_FAST=$v run $v; done
