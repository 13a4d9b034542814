#!/bin/bash
# Round 2, GPU call 25: bake timings (all five tables, Fresnel included) and ncu --set full of the bake kernels on reduced tables.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
timeout 600 python tools/bake_tables.py --tables conductor,glossy,glass,thin_glass,fresnel > $O/r25_bake_times.json 2> $O/r25_bake_times.err; cat $O/r25_bake_times.json; tail -3 $O/r25_bake_times.err
timeout 300 python tools/bake_tables.py --tables conductor,glossy,glass,thin_glass --scale 8 > $O/r25_bake_small.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_bake -f -o /tmp/p_bake python tools/bake_tables.py --tables conductor,glossy,glass,thin_glass --scale 8 > $O/r25_ncu_bake.log 2>&1
python tools/ncu_summary.py /tmp/p_bake.ncu-rep > $O/r25_bake_ncu.txt 2>&1
grep -E "^kernel|gpu__time_duration|registers_per_thread|issue_active|thread_inst_executed_per_inst_executed.ratio|warps_active|pipe_(xu|fma|alu)|local_op|dram__bytes_read" $O/r25_bake_ncu.txt | head -70
