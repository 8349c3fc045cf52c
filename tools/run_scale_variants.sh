#!/bin/bash
# usage: tools/run_scale_variants.sh TAG N "extra args" ...   (one bench.py run per extra-args string on N ranks;
# JSON line -> gpurun_out/TAG_nN_vK.json)
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
i=0
for extra in "$@"; do
  i=$((i+1))
  out=gpurun_out/${TAG}_n${N}_v${i}
  if [ "$N" = "1" ]; then
    timeout ${BENCH_TIMEOUT:-300} python bench.py --gpus 1 $extra > $out.json 2> $out.err
  else
    timeout ${BENCH_TIMEOUT:-300} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+i)) bench.py --gpus $N $extra > $out.json 2> $out.err
  fi
  echo "$TAG N=$N v$i [$extra] rc=$?"
  grep -v "^\*\*\*\|OMP_NUM_THREADS\|^$\|NCCL version" $out.err | tail -3 | cut -c1-400
done
