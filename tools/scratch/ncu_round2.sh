#!/bin/bash
# round-2 ncu evidence (one GPU). Each capture only after the plain command exited 0.
set -x
O=gpurun_out
R=/tmp/ncu; mkdir -p $R   # reports stay on the box (64 MiB limit): only summaries come back
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-c4"
timeout 600 $B > $O/n2_plain.json 2> $O/n2_plain.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file $O/r2c_launches_c2_nprobe24.csv $B > $O/n2_l.out 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:ivf_scan_tcr -c 2 -f -o $R/r2c_scan_one_c2 $B > $O/n2_f.out 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:"ivf_scan_tc_kernel|select_warp" -c 3 -f -o $R/r2c_coarse_select_c2 $B > $O/n2_g.out 2>&1
# C3 (d = 960): coarse + scan are both rows-major launches
B3="python bench.py --workload C3 --steps 2 --warmup 3 --no-cpu-baseline --no-c4"
timeout 600 $B3 > $O/n2_c3_plain.json 2> $O/n2_c3_plain.err && \
timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:ivf_scan_tcr -c 2 -f -o $R/r2c_scan_c3 $B3 > $O/n2_c3.out 2>&1
# C4 shard: one-MMA kernel at nprobe 16 and the exact warp kernel at nprobe 1 and 2
RG="python tools/bench_regimes.py --regime C4s --steps 1"
timeout 900 $RG --nprobes 16 > $O/n2_c4_plain.jsonl 2> $O/n2_c4_plain.err && \
timeout 1500 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:ivf_scan_tcr -c 1 -f -o $R/r2c_scan_one_c4s_np16 $RG --nprobes 16 > $O/n2_c4a.out 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:ivf_scan_kernel -c 1 -f -o $R/r2c_scan_warp_c4s_np1 $RG --nprobes 1 > $O/n2_c4b.out 2>&1
ls -la $R
for f in $R/*.ncu-rep; do b=$(basename $f .ncu-rep); python tools/ncu_summary.py $f $O/${b}_ncu_full.json; done
du -sh $O
