#!/bin/bash
# weak-scaling solve at N ranks with deeper hierarchies (the coarsest chain grows with the rank count)
N=${1:-8}
O=gpurun_out/r2k_n$N
mkdir -p $O
export TORCH_NCCL_HEARTBEAT_TIMEOUT_SEC=120
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for L in 4 5; do
( time timeout 200 $TR --master-port $((29710 + L)) bench.py --gpus $N --steps 20 --warmup 5 --no-extra --levels $L ) > $O/bench_l$L.out 2> $O/bench_l$L.err; echo "rc=$?" >> $O/bench_l$L.err
done
grep "rc=\|^real" $O/*.err | tr '\n' ' '; echo
for f in $O/bench*.out; do echo $f; python - "$f" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('value','ms_per_step','n_gpus')}, d.get('solve'))
except Exception as e: print('ERR',e)
P
done
for f in $O/*.err; do tail -c 100000 $f > $f.t; mv $f.t $f; done
