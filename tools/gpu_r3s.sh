#!/bin/bash
# r3s: --set full of the wide kernels at the final code: k_wide_secondary on teapot_scene (a full-size pass) and cornell_smoke, k_trace_batch
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1"
NCU="ncu --set full --clock-control none --import-source on"
C1="$B --workload teapot_scene --steps 1 --warmup 1"
$C1 > $O/plain_c1.log 2>&1 && $NCU -k regex:'k_wide_secondary' -s 3 -c 1 -o $O/r3s_c1 $C1 > $O/ncu_c1.log 2>&1
C4="$B --workload cornell_smoke --spp 64 --steps 1 --warmup 1"
$C4 > $O/plain_c4.log 2>&1 && $NCU -k regex:'k_wide_secondary' -s 3 -c 1 -o $O/r3s_c4 $C4 > $O/ncu_c4.log 2>&1
RB="$B --workload raybatch --rays 2097152 --steps 1 --warmup 1"
$RB > $O/plain_rb.log 2>&1 && $NCU -k regex:'k_trace_batch' -s 1 -c 1 -o $O/r3s_rb $RB > $O/ncu_rb.log 2>&1
for r in c1 c4 rb; do ncu -i $O/r3s_$r.ncu-rep --page raw --csv > $O/r3s_${r}_raw.csv 2>/dev/null; ncu -i $O/r3s_$r.ncu-rep --page source --csv > $O/r3s_${r}_source.csv 2>/dev/null; done
rm -f $O/*.ncu-rep
for r in c1 c4 rb; do python tools/sass_regions.py $O/r3s_${r}_source.csv 100 0 | awk '$3+0 > 1.5 || NR<3' > $O/r3s_${r}_regions.txt; done
rm -f $O/r3s_*_source.csv; ls -la $O | grep r3s
