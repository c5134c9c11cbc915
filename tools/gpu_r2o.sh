#!/bin/bash
# r2o (1 GPU): queue capacity A/B on C1 / C5, launch list of C1, --set full of k_wide_secondary on C1 at full spp (second step: the
# host's kernel choice is settled from the start)
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1"
for cap in 4194304 8388608 16777216 33554432; do
  HRT_WAVE_CAPACITY=$cap $B --workload teapot_scene --steps 3 --warmup 2 > $O/r2o_c1_cap$cap.json 2>/dev/null
  HRT_WAVE_CAPACITY=$cap $B --workload big_mesh --bvh lbvh_gpu --indexed --steps 2 --warmup 2 > $O/r2o_c5_cap$cap.json 2>/dev/null
  HRT_WAVE_CAPACITY=$cap $B --workload cornell_smoke --spp 512 --steps 2 --warmup 2 > $O/r2o_c4_cap$cap.json 2>/dev/null
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2o_c*_cap*.json")):
    try:
        d = json.load(open(f)); print(f, round(d["value"]), "ms", round(d["ms_per_step"], 2), "launches", d["gpu_launches"], {k: round(v, 1) for k, v in d["kernel_ms"].items() if k != "note"})
    except Exception as e:
        print(f, "FAILED", e)
PY
C1="$B --workload teapot_scene --steps 1 --warmup 1"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2o_c1_launches.csv $C1 > $O/ncu_c1_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_wide_secondary' -s 170 -c 3 -o $O/r2o_c1 $C1 > $O/ncu_c1.log 2>&1
ncu -i $O/r2o_c1.ncu-rep --page raw --csv > $O/r2o_c1_raw.csv 2>/dev/null
ncu -i $O/r2o_c1.ncu-rep --page source --csv > $O/r2o_c1_source.csv 2>/dev/null
rm -f $O/*.ncu-rep
ls -la $O | grep r2o
