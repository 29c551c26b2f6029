#!/bin/bash
# Several nodes, data parallel.  srun starts ONE launcher per node; every launcher gets the same rendezvous endpoint (the first
# node of the allocation) and its own machine rank.  Inside a node the native engines use NVLink peer memory; across nodes
# `Accelerator.prepare` falls back to torch DDP over NCCL (peer mappings do not cross hosts).
#SBATCH --job-name=b200-multinode
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
    --num_processes $((SLURM_NNODES * GPUS_PER_NODE)) \
    --num_machines $SLURM_NNODES \
    --rdzv_backend c10d \
    --main_process_ip $head_node_ip \
    --main_process_port 29500 \
    --mixed_precision bf16"
export SCRIPT="$ACCELERATE_DIR/examples/complete_nlp_example.py"
export SCRIPT_ARGS="--output_dir $ACCELERATE_DIR/examples/output --checkpointing_steps epoch"

# --machine_rank must be evaluated on each node, hence the quoted \$SLURM_NODEID
srun bash -c "$LAUNCHER --machine_rank \$SLURM_NODEID $SCRIPT $SCRIPT_ARGS"
