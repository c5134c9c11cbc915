#!/bin/bash
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1"
NCU="ncu --set full --clock-control none --import-source on"
C1="$B --workload teapot_scene --spp 16 --steps 1 --warmup 1"
$C1 > $O/plain_c1.log 2>&1 && $NCU -k regex:'k_wave3_trace' -s 12 -c 2 -o $O/r2e_c1 $C1 > $O/ncu_c1.log 2>&1; tail -2 $O/ncu_c1.log
C2="$B --workload shiny_teapot --spp 32 --steps 1 --warmup 1"
$C2 > $O/plain_c2.log 2>&1 && $NCU -k regex:'k_wave3_trace' -s 3 -c 1 -o $O/r2e_c2 $C2 > $O/ncu_c2.log 2>&1; tail -2 $O/ncu_c2.log
