#!/bin/bash
# CPU-only nodes (gloo): useful for testing a script's distributed logic before it gets GPU time.  One rank per node here;
# raise --num_processes / --num_cpu_threads_per_process to use more of each node.
#SBATCH --job-name=b200-multicpu
#SBATCH -D .
#SBATCH --output=O-%x.%j
#SBATCH --error=E-%x.%j
#SBATCH --nodes=2
#SBATCH --ntasks-per-node=1
#SBATCH --cpus-per-task=32
#SBATCH --time=00:59:00

set -euo pipefail
export ACCELERATE_DIR="${ACCELERATE_DIR:-$PWD}"
export PYTHONPATH="$ACCELERATE_DIR:${PYTHONPATH:-}"
head_node_ip=$(scontrol show hostnames "$SLURM_JOB_NODELIST" | head -n 1)
export LAUNCHER="python -m accelerate_hpc_test_b200.commands.accelerate_cli launch \
    --cpu --multi_gpu \
    --num_processes $SLURM_NNODES \
    --num_machines $SLURM_NNODES \
    --num_cpu_threads_per_process $SLURM_CPUS_PER_TASK \
    --rdzv_backend c10d \
    --main_process_ip $head_node_ip \
    --main_process_port 29500"
export SCRIPT="$ACCELERATE_DIR/examples/complete_nlp_example.py"
export SCRIPT_ARGS="--cpu --model tiny --output_dir $ACCELERATE_DIR/examples/output"

srun bash -c "$LAUNCHER --machine_rank \$SLURM_NODEID $SCRIPT $SCRIPT_ARGS"
