#!/bin/bash
# Round 2, GPU call 12: the whole GPU suite + smoke at HEAD, the default bench lines (16 sample passes per frame in flight) of C3 / C2 / C4 / C5,
# the launch list of C4 (ReSTIR DI kernels) and the refreshed ncu capture of C3's traversal kernel (profiles/traffic.json key of the default command).
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
( time timeout 2400 python -m pytest tests -x -q -m gpu ) > $O/r12_tests.log 2>&1
tail -8 $O/r12_tests.log
( time python -c "import __graft_entry__ as g; g.smoke()" ) > $O/r12_smoke.log 2>&1; tail -3 $O/r12_smoke.log
timeout 900 python bench.py > $O/r12_c3.json 2> $O/r12_c3.err
timeout 900 python bench.py --workload cornell --no-cpu-baseline > $O/r12_c2.json 2> $O/r12_c2.err
timeout 900 python bench.py --workload many_lights --no-cpu-baseline > $O/r12_c4.json 2> $O/r12_c4.err
timeout 900 python bench.py --workload adaptive_full --width 3840 --height 2160 --no-cpu-baseline > $O/r12_c5.json 2> $O/r12_c5.err
for f in $O/r12_c?.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step e2e", round(d["e2e"]["value"],1), "launches", d["gpu_launches"], {k:round(v,2) for k,v in r["stage_ms"].items()}, "frac", round(r["frac"],3), d["clocks"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
PY
done
bash tools/r02_profile.sh synthetic > $O/r12_profile_c3.log 2>&1; tail -12 $O/r12_profile_c3.log
W=many_lights
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/p_${W}_launches.csv python bench.py --profile --steps 1 --samples-per-frame 1 --workload $W > $O/p_${W}_ncu_launches.log 2>&1
python tools/ncu_launches.py $O/p_${W}_launches.csv --sum > $O/p_${W}_launches_summary.txt 2>&1
tail -30 $O/p_${W}_launches_summary.txt
ncu --set full --clock-control none --import-source on -k regex:"k_ReSTIR_DI" --launch-skip 9 -c 3 -f -o /tmp/p_restir python bench.py --profile --steps 1 --samples-per-frame 1 --workload $W > $O/p_${W}_ncu_restir.log 2>&1
python tools/ncu_summary.py /tmp/p_restir.ncu-rep > $O/p_${W}_restir_full.txt 2>&1
grep -E "^kernel|gpu__time_duration|registers_per_thread|issue_active|thread_inst_executed_per_inst_executed.ratio|long_scoreboard_per|dram__bytes|lts__t_sector_hit|l1tex__t_sector_hit|local_op|warps_active" $O/p_${W}_restir_full.txt | head -60
du -sh $O
