#!/bin/bash
# One node, every B200 of it: `accelerate launch` forks one process per GPU.
#SBATCH --job-name=b200-multigpu
#SBATCH -D .
#SBATCH --output=O-%x.%j
#SBATCH --error=E-%x.%j
#SBATCH --nodes=1
#SBATCH --ntasks-per-node=1          # one launcher per node; the launcher starts the ranks
#SBATCH --gres=gpu:8
#SBATCH --cpus-per-task=96
#SBATCH --time=01:59:00

set -euo pipefail
# activate the environment that has torch and this package, e.g.
#   source /path/to/venv/bin/activate
export ACCELERATE_DIR="${ACCELERATE_DIR:-$PWD}"
export PYTHONPATH="$ACCELERATE_DIR:${PYTHONPATH:-}"
GPUS_PER_NODE=${SLURM_GPUS_ON_NODE:-8}

python -m accelerate_hpc_test_b200.commands.accelerate_cli launch \
    --num_processes "$GPUS_PER_NODE" \
    --mixed_precision bf16 \
    "$ACCELERATE_DIR/examples/complete_nlp_example.py" \
    --output_dir "$ACCELERATE_DIR/examples/output" --checkpointing_steps epoch
