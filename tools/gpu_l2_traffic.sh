#!/bin/bash
# DRAM bytes per ResNet-50 step with the caches left alone between kernels (ncu --cache-control none): the tile order
# + L2 eviction hints against first-to-last / unhinted.  One pass each; prints the per-step totals of the last two steps.
mkdir -p gpurun_out
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
PL_L2_HINTS=0 PL_NET_ZIGZAG=0 timeout 900 ncu --metrics $M --clock-control none --cache-control none -c 1200 --csv \
    --log-file gpurun_out/l2_traffic_off.csv $B > gpurun_out/l2_traffic_off.log 2>&1
echo "off rc=$?"
timeout 900 ncu --metrics $M --clock-control none --cache-control none -c 1200 --csv \
    --log-file gpurun_out/l2_traffic_on.csv $B > gpurun_out/l2_traffic_on.log 2>&1
echo "on rc=$?"
python tools/l2_traffic_sum.py gpurun_out/l2_traffic_off.csv gpurun_out/l2_traffic_on.csv
