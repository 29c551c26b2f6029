#!/bin/bash
# Llama-3-8B on 65,536-token sequences, one sequence split over 8 GPUs (Ulysses sequence parallelism).
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
python -m accelerate_hpc_test_b200.commands.accelerate_cli launch \
    --config_file "$here/sp_alst.accelerate_config.yml" \
    "$here/sp_alst.py" --sp-size 8 --sequence-length 65536 --num-steps 10 "$@"
