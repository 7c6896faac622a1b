#!/bin/bash
# NCCL protocol / channel settings for the one-plane halo shifts (100 KB messages): K-apply without overlap + solve
N=${1:-4}
O=gpurun_out/r2p_n$N
mkdir -p $O
export TORCH_NCCL_HEARTBEAT_TIMEOUT_SEC=120 VFEM_NO_HALO_OVERLAP=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
B="bench.py --gpus $N --steps 20 --warmup 5 --no-extra --no-cpu"
P=29810
for cfg in "default" "NCCL_PROTO=LL" "NCCL_PROTO=LL128" "NCCL_MAX_P2P_NCHANNELS=1" "NCCL_MIN_P2P_NCHANNELS=8"; do
  P=$((P + 1)); name=$(echo $cfg | tr '=' '_')
  ( [ "$cfg" != default ] && export $cfg; timeout 150 $TR --master-port $P $B ) > $O/$name.out 2> $O/$name.err; echo "rc=$?" >> $O/$name.err
  echo "== $cfg $(grep rc= $O/$name.err)"; python -c "
import json
d=json.loads(open('$O/$name.out').read().strip().splitlines()[-1]); print('  k-apply %.1f GDOF/s %.2f us/step; solve %.2f ms (%d it)' % (d['value'], d['ms_per_step']*1e3, d['solve']['ms_per_solve'], d['solve']['iterations']))"
done
for f in $O/*.err; do tail -c 30000 $f > $f.t; mv $f.t $f; done
