#!/bin/bash
# BASELINE.json configs on N GPUs: FSDP2 bf16, tp=2 x dp_shard, Ulysses sp=N (8192 tokens per rank), cp=2, FSDP2 fp8 (MXFP8).
# Usage: run_configs.sh N [steps]
N=${1:-2}; STEPS=${2:-5}; OUT=gpurun_out/configs_n$N; mkdir -p $OUT
run() { name=$1; shift; timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps $STEPS --warmup 3 "$@" > $OUT/$name.json 2> $OUT/$name.err; echo "$name: $(tail -1 $OUT/$name.json | cut -c1-230)"; grep -E "Error|error:|Traceback" $OUT/$name.err | head -3; }
run fsdp_bf16 --parallelism fsdp
run tp2 --parallelism tp2
run sp$N --parallelism sp$N --seq-len $((8192 * N)) --micro-batch 1
[ $N -ge 4 ] && run cp2 --parallelism cp2 --seq-len 16384 --micro-batch 1
run fsdp_fp8 --parallelism fsdp --precision fp8
