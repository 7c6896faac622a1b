#!/bin/bash
# round 2, multi-GPU call: N ranks (default 2): slab correctness vs 1 GPU, bench with / without graph replay, levels 3 / 4
N=${1:-2}
O=gpurun_out/r2c_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tools/slab_check.py > $O/slab_check.log 2>&1; echo "rc=$?" >> $O/slab_check.log
timeout 900 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "rc=$?" >> $O/bench.err
VFEM_SLAB_NO_GRAPH=1 timeout 900 $TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 --no-extra > $O/bench_nograph.json 2>> $O/bench.err
timeout 900 $TR --master-port 29514 bench.py --gpus $N --steps 20 --warmup 5 --no-extra --levels 4 > $O/bench_l4.json 2>> $O/bench.err
timeout 900 $TR --master-port 29515 bench.py --gpus $N --steps 20 --warmup 5 --no-extra --levels 5 > $O/bench_l5.json 2>> $O/bench.err
timeout 600 $TR --master-port 29516 bench.py --impl reference --gpus $N --steps 5 --warmup 3 > $O/bench_ref.json 2>> $O/bench.err
tail -n 5 $O/slab_check.log $O/bench.err; for f in $O/bench*.json; do echo $f; python - "$f" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('value','ms_per_step','n_gpus')}, d.get('solve'), {k:(v.get('k_apply_gdof_per_s'),v.get('solve',{}).get('ms_per_solve'),v.get('solve',{}).get('iterations')) if 'error' not in v else v for k,v in d.get('configs',{}).items()}, d.get('e2e',{}).get('value'))
except Exception as e: print('ERR',e)
P
done
