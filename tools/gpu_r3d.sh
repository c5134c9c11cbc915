#!/bin/bash
# r3d: phase / refill thresholds of the wide walk, A/B at HEAD (tools/build_variant.sh)
O=gpurun_out; mkdir -p $O
B="python bench.py --no-cpu --e2e-steps 1 --steps 3 --warmup 2"
V=hobbyraytracer_b200/csrc/variants
for n in base p6 p12 p16 r16 r28; do
  so=$PWD/$V/libhrt_b200_$n.so; [ $n = base ] && so=$PWD/hobbyraytracer_b200/csrc/libhrt_b200.so
  HRT_PRODUCT_SO=$so $B --workload teapot_scene > $O/r3d_${n}_c1.json 2>/dev/null
  HRT_PRODUCT_SO=$so $B --workload raybatch > $O/r3d_${n}_rb.json 2>/dev/null
  HRT_PRODUCT_SO=$so $B --workload big_mesh --bvh lbvh_gpu --indexed --spp 16 > $O/r3d_${n}_c5.json 2>/dev/null
  HRT_PRODUCT_SO=$so $B --workload cornell_smoke --spp 256 > $O/r3d_${n}_c4.json 2>/dev/null
done
python - <<'PY'
import json, glob
for n in ("base", "p6", "p12", "p16", "r16", "r28"):
    row = []
    for w in ("c1", "rb", "c5", "c4"):
        try: row.append(round(json.load(open(f"gpurun_out/r3d_{n}_{w}.json"))["value"]))
        except Exception as e: row.append("FAIL")
    print(n, row)
PY
