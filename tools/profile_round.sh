#!/bin/bash
# Round profile on the GPU box: launch lists of `bench.py --profile` (C2, C3) and one `ncu --set full` capture of the
# k_trace_queue launches of one step each. The .ncu-rep files (55-65 MB each) are summarised on the box and deleted: only
# text/CSV summaries travel back in gpurun_out/ (64 MiB limit); they are copied to profiles/ by hand afterwards.
set -x
cd "$(dirname "$0")/.."
O=gpurun_out
python bench.py --profile --steps 2 > $O/prof_c2_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_c2.csv python bench.py --profile --steps 2 > $O/ncu_c2.log 2>&1
python tools/ncu_launches.py $O/launches_c2.csv --sum > $O/launches_c2_summary.txt 2>&1
python bench.py --profile --steps 2 --workload synthetic > $O/prof_c3_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_c3.csv python bench.py --profile --steps 2 --workload synthetic > $O/ncu_c3.log 2>&1
python tools/ncu_launches.py $O/launches_c3.csv --sum > $O/launches_c3_summary.txt 2>&1
if [ "$1" == "full" ]; then
ncu --set full --clock-control none -k regex:k_trace_queue --launch-skip 102 -c 34 -f -o /tmp/trace_full_c2 python bench.py --profile --steps 2 > $O/ncu_full_c2.log 2>&1
python tools/ncu_traffic.py /tmp/trace_full_c2.ncu-rep cornell_1920x1080 1 > $O/traffic_c2.log 2>&1
python tools/ncu_summary.py /tmp/trace_full_c2.ncu-rep > $O/trace_full_c2_summary.txt 2>&1
ncu --set full --clock-control none -k regex:k_trace_queue --launch-skip 102 -c 34 -f -o /tmp/trace_full_c3 python bench.py --profile --steps 2 --workload synthetic > $O/ncu_full_c3.log 2>&1
python tools/ncu_traffic.py /tmp/trace_full_c3.ncu-rep synthetic_1920x1080 1 > $O/traffic_c3.log 2>&1
python tools/ncu_summary.py /tmp/trace_full_c3.ncu-rep > $O/trace_full_c3_summary.txt 2>&1
cp profiles/traffic.json $O/traffic.json
fi
du -sh $O
