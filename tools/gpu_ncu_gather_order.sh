#!/bin/bash
# ncu --set full of the persistent fused gather (second step of a 30-pair 1080p run) in both tile orders:
#   gpurun -- 'bash tools/gpu_ncu_gather_order.sh'   ->  gpurun_out/gc_order{0,1}.ncu-rep ; summarise with tools/ncu_summary.py
PAIRS=${PAIRS:-30}
python tools/profile_step.py 2 $PAIRS > gpurun_out/plain_gc_order.log 2>&1 || { cat gpurun_out/plain_gc_order.log; exit 1; }
for o in ${ORDERS:-1 0}; do
  VFI_OPT_T_FAST=$o timeout 280 ncu --set full --clock-control none --import-source on -k regex:gather_conv0a_p_kernel -s 1 -c 1 -f \
      -o gpurun_out/gc_order$o python tools/profile_step.py 2 $PAIRS > gpurun_out/ncu_gc_order$o.log 2>&1
  echo "order $o rc=$?"
done
ls -la gpurun_out/gc_order*.ncu-rep
