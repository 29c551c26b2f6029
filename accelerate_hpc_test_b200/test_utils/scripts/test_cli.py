"""Smallest possible launched program: every rank reports the device it was given (what ``accelerate launch`` smoke tests run)."""

import torch

from accelerate_hpc_test_b200 import PartialState


def main():
    state = PartialState()
    where = f"{torch.cuda.device_count()} GPUs visible" if torch.cuda.is_available() else "no GPU, running on CPU"
    print(f"rank {state.process_index} of {state.num_processes}: {where}; device {state.device}", flush=True)
    state.wait_for_everyone()
    if state.is_main_process:
        print("Successfully ran on", state.num_processes, "processes" if state.num_processes > 1 else "process")


if __name__ == "__main__":
    main()
