#!/bin/bash
# multi-GPU call (N ranks, default 2).  Every command runs under a short timeout and the script stops at the first one that
# times out: a stalled collective must not hold the box.
N=${1:-2}
O=gpurun_out/r2c_n$N
mkdir -p $O
export TORCH_NCCL_HEARTBEAT_TIMEOUT_SEC=120
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
PORT=29510
step() { # name, time limit, command...
  local name=$1 lim=$2; shift 2
  [ -e $O/STOP ] && return
  PORT=$((PORT + 1))
  ( time timeout $lim "$@" ) > $O/$name.out 2> $O/$name.err
  local rc=$?
  echo "rc=$rc" >> $O/$name.err
  if [ $rc -eq 124 ] || [ $rc -eq 137 ]; then echo "$name timed out" > $O/STOP; fi
}
[ -z "$SKIP_SLAB_CHECK" ] && step slab_check 200 $TR --master-port $PORT tools/slab_check.py
step bench_apply 200 $TR --master-port $((PORT + 1)) bench.py --gpus $N --steps 20 --warmup 5 --no-extra --no-solve
if [ -e $O/STOP ]; then   # diagnose: the same without the captured halo-shift + apply graph
  rm -f $O/STOP
  VFEM_NO_HALO_GRAPH=1 step bench_apply_nohalograph 200 $TR --master-port $((PORT + 1)) bench.py --gpus $N --steps 20 --warmup 5 --no-extra --no-solve
  echo "halo graph hang" >> $O/STOP
fi
step bench_solve 300 $TR --master-port $((PORT + 1)) bench.py --gpus $N --steps 20 --warmup 5 --no-extra
VFEM_SLAB_NO_GRAPH=1 step bench_solve_nograph 300 $TR --master-port $((PORT + 1)) bench.py --gpus $N --steps 20 --warmup 5 --no-extra
step bench_l4 300 $TR --master-port $((PORT + 1)) bench.py --gpus $N --steps 20 --warmup 5 --no-extra --levels 4
step bench_full 500 $TR --master-port $((PORT + 1)) bench.py --gpus $N --steps 20 --warmup 5
step bench_ref 200 $TR --master-port $((PORT + 1)) bench.py --impl reference --gpus $N --steps 5 --warmup 3
cat $O/STOP 2>/dev/null; grep "rc=\|^real" $O/*.err | tr '\n' ' '; echo
for f in $O/bench*.out; do echo $f; python - "$f" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('value','ms_per_step','n_gpus')}, d.get('solve'), {k:(v.get('k_apply_gdof_per_s'),v.get('solve',{}).get('ms_per_solve'),v.get('solve',{}).get('iterations')) if 'error' not in v else v for k,v in d.get('configs',{}).items()}, d.get('e2e',{}).get('value'))
except Exception as e: print('ERR',e)
P
done
tail -n 5 $O/slab_check.out
