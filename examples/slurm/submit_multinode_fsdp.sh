#!/bin/bash
# Several nodes, sharded (FSDP2 semantics), configuration taken from fsdp_config.yaml next to this script.  Hybrid sharding is the
# natural layout: shard inside a node (NVSwitch), replicate across nodes — set it with
#   --parallelism_config_dp_shard_size 8 --parallelism_config_dp_replicate_size $SLURM_NNODES
#SBATCH --job-name=b200-multinode-fsdp
#SBATCH -D .
#SBATCH --output=O-%x.%j
#SBATCH --error=E-%x.%j
#SBATCH --nodes=4
#SBATCH --ntasks-per-node=1
#SBATCH --gres=gpu:8
#SBATCH --cpus-per-task=96
#SBATCH --time=01:59:00

set -euo pipefail
export ACCELERATE_DIR="${ACCELERATE_DIR:-$PWD}"
export PYTHONPATH="$ACCELERATE_DIR:${PYTHONPATH:-}"
GPUS_PER_NODE=8
head_node_ip=$(scontrol show hostnames "$SLURM_JOB_NODELIST" | head -n 1)
export LAUNCHER="python -m accelerate_hpc_test_b200.commands.accelerate_cli launch \
    --config_file $ACCELERATE_DIR/examples/slurm/fsdp_config.yaml \
    --num_processes $((SLURM_NNODES * GPUS_PER_NODE)) \
    --num_machines $SLURM_NNODES \
    --rdzv_backend c10d \
    --main_process_ip $head_node_ip \
    --main_process_port 29500"
export SCRIPT="$ACCELERATE_DIR/examples/torch_native_parallelism/nd_parallel.py"
export SCRIPT_ARGS="--dp-shard-size 8 --dp-replicate-size $SLURM_NNODES --model llama3-8b --sequence-length 2048"

srun bash -c "$LAUNCHER --machine_rank \$SLURM_NODEID $SCRIPT $SCRIPT_ARGS"
