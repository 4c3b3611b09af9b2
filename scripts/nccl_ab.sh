#!/bin/bash
# usage (on an N-GPU box): nccl_ab.sh <N> <tag>  -> gpurun_out/<tag>_nccl_ab.txt
# Does the per-step all_gather's own kernel (NCCL channels = resident CTAs that spin until every rank's solve is done)
# take SM room from the solve kernels?  Same bench command under different NCCL channel limits.
n=$1; t=$2; o=gpurun_out; mkdir -p $o
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29520"
A="bench.py --gpus $n --steps 50 --warmup 3 --no-cpu-baseline --latency-samples 0"
: > $o/${t}_nccl_ab.txt
run() { echo "== $1" >> $o/${t}_nccl_ab.txt; env $1 timeout 200 $T $A 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], round(d['e2e']['value']))" >> $o/${t}_nccl_ab.txt; }
run RMPC_AB=default
run NCCL_MAX_NCHANNELS=1
run NCCL_MAX_CTAS=1
run "NCCL_MAX_NCHANNELS=1 NCCL_NTHREADS=64"
run RMPC_AB=default
cat $o/${t}_nccl_ab.txt
