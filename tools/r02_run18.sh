#!/bin/bash
# Round 2, GPU call 18: knob sweeps at the final defaults (path lanes, queue grid size) on C3 and C4; ncu --set full of the phased ReSTIR DI kernels.
cd "$GRAFT_REPO_ROOT"
O=gpurun_out
mkdir -p $O
run() { name=$1; wl=$2; shift 2; env "$@" timeout 600 python bench.py --workload $wl --steps 2 --no-cpu-baseline > $O/r18_$name.json 2> $O/r18_$name.err; python - $O/r18_$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); r=d["roofline"]
    print(sys.argv[1], round(d["value"],1), "Mrays/s", round(d["ms_per_step"],2), "ms/step", {k:round(v,2) for k,v in r["stage_ms"].items()}, "lanes", d["config"]["path_lanes"])
except Exception as e:
    print("failed", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
}
run c3_lanes1 synthetic B200PT_PATH_LANES=1
run c3_lanes2 synthetic B200PT_PATH_LANES=2
run c3_lanes3 synthetic B200PT_PATH_LANES=3
run c3_lanes1_q24 synthetic B200PT_PATH_LANES=1 B200PT_QUEUE_BLOCKS_PER_SM=24
run c3_lanes2_q24 synthetic B200PT_PATH_LANES=2 B200PT_QUEUE_BLOCKS_PER_SM=24
run c3_lanes2_q32 synthetic B200PT_PATH_LANES=2 B200PT_QUEUE_BLOCKS_PER_SM=32
run c4_lanes1 many_lights B200PT_PATH_LANES=1
run c4_lanes2 many_lights B200PT_PATH_LANES=2
run c4_lanes3 many_lights B200PT_PATH_LANES=3
run c4_lanes2_q24 many_lights B200PT_PATH_LANES=2 B200PT_QUEUE_BLOCKS_PER_SM=24
W=many_lights
ncu --set full --clock-control none --import-source on -k regex:"k_ReSTIR_DI" --launch-skip 27 -c 9 -f -o /tmp/p_restir python bench.py --profile --steps 1 --samples-per-frame 1 --workload $W > $O/p18_${W}_ncu_restir.log 2>&1
python tools/ncu_summary.py /tmp/p_restir.ncu-rep > $O/p18_${W}_restir_phased_full.txt 2>&1
grep -E "^kernel|gpu__time_duration|registers_per_thread|smsp__issue_active|thread_inst_executed_per_inst_executed.ratio|long_scoreboard_per|dram__bytes_read|sm__warps_active" $O/p18_${W}_restir_phased_full.txt | head -60
