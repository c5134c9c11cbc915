#!/bin/bash
# r2b: first B200 run of the 8-wide batch kernel: gpu tests, ray batch bench, ncu --set full of k_trace_batch
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2b_pytest.log 2>&1; tail -3 $O/r2b_pytest.log
B="python bench.py --no-cpu --e2e-steps 1"
$B --workload raybatch --rays 8388608 --steps 2 --warmup 1 > $O/r2b_raybatch.json 2> $O/r2b_raybatch.err; cat $O/r2b_raybatch.json | cut -c1-400
RB="$B --workload raybatch --rays 2097152 --steps 1 --warmup 1"
$RB > $O/plain_rb.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_trace_batch -s 1 -c 1 -o $O/r2b_rb $RB > $O/ncu_rb.log 2>&1
tail -2 $O/ncu_rb.log
