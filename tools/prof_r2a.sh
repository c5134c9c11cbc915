#!/bin/bash
# Round-2 baseline: bench lines at HEAD for the configs whose rays really traverse (C1, C5, ray batch) and --set full
# captures of the incoherent-walk kernels (VERDICT r1 "next round" item 1).  Run under gpurun on one B200.
O=gpurun_out
mkdir -p $O
run() { echo "== $*" >&2; "$@"; }
B="python bench.py --no-cpu --e2e-steps 1"
run $B --workload teapot_scene --steps 2 --warmup 1 > $O/r2a_teapot_scene.json 2> $O/r2a_teapot_scene.err
run $B --workload raybatch --rays 8388608 --steps 2 --warmup 1 > $O/r2a_raybatch.json 2> $O/r2a_raybatch.err
run $B --workload big_mesh --bvh sah --spp 16 --steps 1 --warmup 1 > $O/r2a_big_mesh.json 2> $O/r2a_big_mesh.err
NCU="ncu --set full --clock-control none --import-source on"
C1="$B --workload teapot_scene --spp 16 --steps 1 --warmup 1"
run $C1 > $O/plain_c1.log 2>&1 && run $NCU -k regex:'k_wave2_secondary|k_wave2_trace' -s 24 -c 4 -o $O/r2a_c1 $C1 > $O/ncu_c1.log 2>&1
RB="$B --workload raybatch --rays 2097152 --steps 1 --warmup 1"
run $RB > $O/plain_rb.log 2>&1 && run $NCU -k regex:k_trace_batch -s 1 -c 1 -o $O/r2a_rb $RB > $O/ncu_rb.log 2>&1
C5="$B --workload big_mesh --bvh sah --spp 4 --steps 1 --warmup 1"
run $C5 > $O/plain_c5.log 2>&1 && run $NCU -k regex:'k_wave2_trace' -s 6 -c 2 -o $O/r2a_c5 $C5 > $O/ncu_c5.log 2>&1
ls -la $O
