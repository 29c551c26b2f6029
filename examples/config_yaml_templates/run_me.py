"""Prints what the current configuration resolves to — run it under each template to see what the template does:

    accelerate launch --config_file examples/config_yaml_templates/fsdp2.yaml examples/config_yaml_templates/run_me.py
"""

import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

from accelerate_hpc_test_b200 import Accelerator

accelerator = Accelerator()
accelerator.print(f"Accelerator state from the current environment:\n{accelerator.state}")
if accelerator.parallelism_config is not None:
    accelerator.print(f"Parallelism:\n{accelerator.parallelism_config}")
for name in ("fp8_recipe_handler", "te_recipe_handler", "ao_recipe_handler"):
    if getattr(accelerator, name, None) is not None:
        accelerator.print(f"FP8 config ({name}):\n{getattr(accelerator, name)}")
        break
accelerator.end_training()
