#!/bin/bash
# r2v: --set full + SASS page of the camera-path kernel on the headline scene (one pass of 128 spp), at HEAD
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --spp 128 --steps 1 --warmup 1"
$B > $O/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_wave2_trace' -s 0 -c 1 -o $O/r2v_c2 $B > $O/ncu_c2.log 2>&1
ncu -i $O/r2v_c2.ncu-rep --page raw --csv > $O/r2v_c2_raw.csv 2>/dev/null
ncu -i $O/r2v_c2.ncu-rep --page source --csv > $O/r2v_c2_source.csv 2>/dev/null
rm -f $O/*.ncu-rep; ls -la $O | grep r2v
