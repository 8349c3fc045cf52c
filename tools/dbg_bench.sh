for d in 0 1 2 4 7; do
  B200M_DBG=$d timeout 60 python bench.py --steps 3 --warmup 3 --no-also --images 60 2>gpurun_out/dbg_$d.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('dbg=$d', 'kernel_ms', round(d['roofline']['kernel_ms_per_step'],2), 'TOPS', round(d['roofline']['achieved'] or 0,1))
except Exception as e: print('dbg=$d failed', e)" >> gpurun_out/dbg.log 2>&1
done
cat gpurun_out/dbg.log
