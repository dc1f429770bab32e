#!/bin/bash
# A/B of one library option through bench.py's VFI_<OPTION> knobs: bash tools/ab_opt.sh <option> <values...>   (e.g. gc_prefetch 0 1)
OPT=$1; shift
run() { env VFI_${OPT^^}=$1 timeout 200 python bench.py --steps 20 --warmup 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$OPT=$1','value',round(d['value'],1),'e2e',round(d['e2e']['value'],1))
print('   ',{k:round(v['ms_per_step'],3) for k,v in (d.get('kernel_breakdown') or {}).items()})
"; }
for v in "$@"; do run $v; done
