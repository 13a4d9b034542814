# node/triangle counts and trace time of the 10 M-triangle scene for different primitive costs of the collapse DP
for c in 0.2 0.4 0.7 1.0; do
  echo "c_prim $c"
  B200PT_BVH_PRIMITIVE_COST=$c timeout 600 python bench.py --workload synthetic --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['value'], d['ms_per_step'], r['stage_ms'], 'nodes/ray', r['nodes_per_ray'], 'tris/ray', r['triangles_per_ray'], d['config']['bvh'])"
done
