for s in 0 9000 13000 19000 28000; do
RG_BLEND_SMEM_F=$s RG_BLEND_SMEM_B=$s timeout 100 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-ref-cuda --no-graph 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('smem',$s,d['value'],d['kernels']['blend_forward']['ms_per_step'],d['kernels']['blend_backward']['ms_per_step'])"
done
