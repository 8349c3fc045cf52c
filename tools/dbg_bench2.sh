for d in 64 66 79 91; do
  echo "=== dbg=$d" >> gpurun_out/dbg2.log
  B200M_DBG=$d timeout 60 python bench.py --steps 1 --warmup 3 --no-also --images 60 2>gpurun_out/dbg_$d.err | grep -E "role|metric" | cut -c1-200 | tail -5 >> gpurun_out/dbg2.log
done
cat gpurun_out/dbg2.log
