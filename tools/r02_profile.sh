#!/bin/bash
# Round-2 profile on the GPU box (one GPU): launch list of `bench.py --profile` (C3 default: 8 passes per frame in flight), `ncu --set full`
# of every k_trace_queue launch of one frame (-> DRAM traffic per step, profiles/traffic.json) and of the shade / surface kernels of
# bounces 0-1. Reports are summarised on the box (the .ncu-rep files stay there); text summaries travel back in gpurun_out/.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
W=${1:-synthetic}
python bench.py --profile --steps 1 --workload $W > $O/p_${W}_plain.log 2>&1 || { tail -5 $O/p_${W}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/p_${W}_launches.csv python bench.py --profile --steps 1 --workload $W > $O/p_${W}_ncu_launches.log 2>&1
python tools/ncu_launches.py $O/p_${W}_launches.csv --sum > $O/p_${W}_launches_summary.txt 2>&1
tail -25 $O/p_${W}_launches_summary.txt
ncu --set full --clock-control none --import-source on -k regex:k_trace_queue --launch-skip 108 -c 36 -f -o /tmp/p_trace python bench.py --profile --steps 1 --workload $W > $O/p_${W}_ncu_trace.log 2>&1
python tools/ncu_traffic.py /tmp/p_trace.ncu-rep ${W}_1920x1080_spf16 1 > $O/p_${W}_traffic.log 2>&1
python tools/ncu_summary.py /tmp/p_trace.ncu-rep > $O/p_${W}_trace_full.txt 2>&1
cp profiles/traffic.json $O/traffic.json
ncu --set full --clock-control none --import-source on -k regex:"k_wf_(bsdf_resolved|env_bsdf_sample|env_light_eval|nee_light_eval|nee_bsdf_sample|surface|camera_surface|camera_|accumulate)" --launch-skip 222 -c 12 -f -o /tmp/p_shade python bench.py --profile --steps 1 --workload $W > $O/p_${W}_ncu_shade.log 2>&1
python tools/ncu_summary.py /tmp/p_shade.ncu-rep > $O/p_${W}_shade_full.txt 2>&1
grep -E "^kernel|gpu__time_duration|registers_per_thread|issue_active|thread_inst_executed_per_inst_executed.ratio|long_scoreboard_per|dram__bytes|lts__t_sector_hit|l1tex__t_sector_hit|local_op" $O/p_${W}_trace_full.txt | head -60
du -sh $O
