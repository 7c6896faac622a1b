#!/bin/bash
# multi-GPU call (N ranks, default 2); every command under a short timeout: a stalled collective must not hold the box
N=${1:-2}
O=gpurun_out/r2c_n$N
mkdir -p $O
export TORCH_NCCL_HEARTBEAT_TIMEOUT_SEC=120
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29511 tools/slab_check.py > $O/slab_check.log 2>&1; echo "rc=$?" >> $O/slab_check.log
timeout 200 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --no-extra --no-solve > $O/bench_apply.json 2> $O/bench_apply.err; echo "rc=$?" >> $O/bench_apply.err
VFEM_NO_HALO_GRAPH=1 timeout 200 $TR --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 --no-extra --no-solve > $O/bench_apply_nohalograph.json 2>> $O/bench_apply.err; echo "rc=$?" >> $O/bench_apply.err
timeout 300 $TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 --no-extra > $O/bench_solve.json 2> $O/bench_solve.err; echo "rc=$?" >> $O/bench_solve.err
VFEM_SLAB_NO_GRAPH=1 timeout 300 $TR --master-port 29514 bench.py --gpus $N --steps 20 --warmup 5 --no-extra > $O/bench_solve_nograph.json 2>> $O/bench_solve.err; echo "rc=$?" >> $O/bench_solve.err
timeout 300 $TR --master-port 29515 bench.py --gpus $N --steps 20 --warmup 5 --no-extra --levels 4 > $O/bench_l4.json 2>> $O/bench_solve.err; echo "rc=$?" >> $O/bench_solve.err
timeout 500 $TR --master-port 29516 bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_full.json 2> $O/bench_full.err; echo "rc=$?" >> $O/bench_full.err
timeout 200 $TR --master-port 29518 bench.py --impl reference --gpus $N --steps 5 --warmup 3 > $O/bench_ref.json 2>> $O/bench_full.err
grep "rc=" $O/*.err $O/slab_check.log; for f in $O/bench*.json; do echo $f; python - "$f" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('value','ms_per_step','n_gpus')}, d.get('solve'), {k:(v.get('k_apply_gdof_per_s'),v.get('solve',{}).get('ms_per_solve'),v.get('solve',{}).get('iterations')) if 'error' not in v else v for k,v in d.get('configs',{}).items()}, d.get('e2e',{}).get('value'))
except Exception as e: print('ERR',e)
P
done
