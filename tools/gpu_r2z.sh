#!/bin/bash
# r2z: after removing the resumable two-child walk: gpu suite; --set full of k_wide_secondary on teapot_scene and k_trace_batch at HEAD
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2z_pytest_gpu.log 2>&1; tail -3 $O/r2z_pytest_gpu.log
B="python bench.py --no-cpu --e2e-steps 1"
C1="$B --workload teapot_scene --steps 1 --warmup 1"
$C1 > $O/plain_c1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_wide_secondary' -s 30 -c 2 -o $O/r2z_c1 $C1 > $O/ncu_c1.log 2>&1
RB="$B --workload raybatch --rays 2097152 --steps 1 --warmup 1"
$RB > $O/plain_rb.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_trace_batch|k_resolve_batch' -s 2 -c 2 -o $O/r2z_rb $RB > $O/ncu_rb.log 2>&1
for r in c1 rb; do ncu -i $O/r2z_$r.ncu-rep --page raw --csv > $O/r2z_${r}_raw.csv 2>/dev/null; done
ncu -i $O/r2z_c1.ncu-rep --page source --csv > $O/r2z_c1_source.csv 2>/dev/null
rm -f $O/*.ncu-rep; ls -la $O | grep r2z
