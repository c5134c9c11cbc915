#!/bin/bash
# r2n (1 GPU): launch list of the default bench, --set full captures of the round-2 kernels at HEAD (each after its own command
# has exited 0 without ncu), raw pages exported on the box, memcheck of the sanitizer case
O=gpurun_out; mkdir -p $O
run() { echo "== $*" >&2; "$@"; }
B="python bench.py --no-cpu --e2e-steps 1"
run python bench.py --steps 2 --warmup 1 --no-cpu > $O/plain_default.log 2>&1 && \
  run ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2n_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > $O/ncu_launches.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
C1="$B --workload teapot_scene --spp 16 --steps 1 --warmup 1"
run $C1 > $O/plain_c1.log 2>&1 && run $NCU -k regex:'k_wide_secondary|k_wave2_trace|k_wave2_shade|k_wave3_resolve' -s 40 -c 8 -o $O/r2n_c1 $C1 > $O/ncu_c1.log 2>&1
RB="$B --workload raybatch --rays 2097152 --steps 1 --warmup 1"
run $RB > $O/plain_rb.log 2>&1 && run $NCU -k regex:'k_trace_batch|k_resolve_batch' -s 2 -c 2 -o $O/r2n_rb $RB > $O/ncu_rb.log 2>&1
C5="$B --workload big_mesh --bvh lbvh_gpu --indexed --spp 4 --steps 1 --warmup 1"
run $C5 > $O/plain_c5.log 2>&1 && run $NCU -k regex:'k_wave2_trace|k_wide_secondary' -s 6 -c 3 -o $O/r2n_c5 $C5 > $O/ncu_c5.log 2>&1
for r in c1 rb c5; do
  [ -f $O/r2n_$r.ncu-rep ] && ncu -i $O/r2n_$r.ncu-rep --page raw --csv > $O/r2n_${r}_raw.csv 2>/dev/null
done
# SASS view of the first captured launch of the two walk kernels (per-instruction executed counts / lanes), then drop the reports:
# gpurun_out/ travels back only below 64 MiB
ncu -i $O/r2n_c1.ncu-rep --page source --csv --kernel-name regex:k_wide_secondary --launch-count 1 > $O/r2n_c1_wsec_source.csv 2>/dev/null
ncu -i $O/r2n_rb.ncu-rep --page source --csv --kernel-name regex:k_trace_batch --launch-count 1 > $O/r2n_rb_source.csv 2>/dev/null
rm -f $O/*.ncu-rep
ls -la $O | head -40
bash tools/gpu_sanitize.sh memcheck
