#!/bin/bash
# N-GPU bench exactly as the driver launches it (+ the reference arm); short timeouts
N=${1:-8}
O=gpurun_out/r2j_n$N
mkdir -p $O
export TORCH_NCCL_HEARTBEAT_TIMEOUT_SEC=120
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
( time timeout 400 $TR --master-port 29701 bench.py --gpus $N --steps 20 --warmup 5 ) > $O/bench_full.out 2> $O/bench_full.err; echo "rc=$?" >> $O/bench_full.err
if grep -q "rc=0" $O/bench_full.err; then
( time timeout 200 $TR --master-port 29702 bench.py --impl reference --gpus $N --steps 5 --warmup 3 ) > $O/bench_ref.out 2> $O/bench_ref.err; echo "rc=$?" >> $O/bench_ref.err
[ -z "$SKIP_HALO_GRAPH" ] && ( time VFEM_HALO_GRAPH=1 timeout 200 $TR --master-port 29703 bench.py --gpus $N --steps 20 --warmup 5 --no-extra --no-solve ) > $O/bench_apply_halograph.out 2> $O/bench_apply_halograph.err; echo "rc=$?" >> $O/bench_apply_halograph.err
fi
grep "rc=\|^real" $O/*.err | tr '\n' ' '; echo
for f in $O/bench*.out; do echo $f; python - "$f" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('value','ms_per_step','n_gpus')}, d.get('solve'), {k:(v.get('k_apply_gdof_per_s'),v.get('solve',{}).get('ms_per_solve'),v.get('solve',{}).get('iterations')) if 'error' not in v else v for k,v in d.get('configs',{}).items()}, d.get('e2e',{}).get('value'))
except Exception as e: print('ERR',e)
P
done
for f in $O/*.err; do tail -c 100000 $f > $f.t; mv $f.t $f; done
