#!/bin/bash
# halo'd K-apply on N ranks: boundary views side by side against one after the other; host issue time per step
N=${1:-4}
O=gpurun_out/r2o_n$N
mkdir -p $O
export TORCH_NCCL_HEARTBEAT_TIMEOUT_SEC=120 VFEM_BENCH_TRACE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
B="bench.py --gpus $N --steps 20 --warmup 5 --no-extra --no-solve --no-cpu"
( timeout 150 $TR --master-port 29801 $B ) > $O/split.out 2> $O/split.err; echo "rc=$?" >> $O/split.err
( VFEM_HALO_SERIAL_VIEWS=1 timeout 150 $TR --master-port 29802 $B ) > $O/serial.out 2> $O/serial.err; echo "rc=$?" >> $O/serial.err
( VFEM_NO_HALO_OVERLAP=1 timeout 150 $TR --master-port 29803 $B ) > $O/nooverlap.out 2> $O/nooverlap.err; echo "rc=$?" >> $O/nooverlap.err
for f in split serial nooverlap; do echo "== $f $(grep rc= $O/$f.err)"; grep "host issue" $O/$f.err | head -8; python -c "
import json,sys
d=json.loads(open('$O/$f.out').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'])"; done
for f in $O/*.err; do tail -c 50000 $f > $f.t; mv $f.t $f; done
