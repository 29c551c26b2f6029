"""Sharded checkpoints can be merged back into one file.  A small model is trained for a step on the sharded engine, its weights are
saved with ``state_dict_type=SHARDED_STATE_DICT`` (a ``torch.distributed.checkpoint`` directory, the layout the reference writes), and
the directory is merged twice — through ``merge_fsdp_weights`` and through the ``accelerate merge-weights`` command line — into
safetensors and into a torch pickle.  The merged weights must be the trained ones, not the initial ones.

    accelerate launch --num_processes 2 --cpu -m accelerate_hpc_test_b200.test_utils.scripts.test_merge_weights
"""

import os
import shutil
import subprocess
import sys
import tempfile

import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin, merge_fsdp_weights, set_seed
from accelerate_hpc_test_b200.utils.operations import broadcast_object_list


class Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.body = torch.nn.Sequential(torch.nn.Linear(16, 64), torch.nn.GELU(), torch.nn.Linear(64, 64), torch.nn.GELU())
        self.head = torch.nn.Linear(64, 8)

    def forward(self, x):
        return self.head(self.body(x))


def merged_tensors(path, safe):
    if safe:
        from safetensors.torch import load_file

        return load_file(os.path.join(path, "model.safetensors"))
    return torch.load(os.path.join(path, "pytorch_model.bin"), weights_only=True)


def main():
    plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="no_wrap", state_dict_type="SHARDED_STATE_DICT")
    accelerator = Accelerator(fsdp_plugin=plugin)
    if accelerator.num_processes < 2:
        raise SystemExit("a sharded checkpoint needs shards: launch with at least two processes")
    set_seed(7)
    model = Net().to(accelerator.device)
    initial = {name: tensor.detach().clone().cpu() for name, tensor in model.state_dict().items()}
    optimizer = torch.optim.AdamW(model.parameters(), lr=0.05)
    model, optimizer = accelerator.prepare(model, optimizer)
    for step in range(2):
        x = torch.randn(8, 16, generator=torch.Generator().manual_seed(step * 10 + accelerator.process_index)).to(accelerator.device)
        accelerator.backward(model(x).pow(2).mean())
        optimizer.step()
        optimizer.zero_grad()
    trained = {name: tensor.float().cpu() for name, tensor in accelerator.get_state_dict(model).items()}
    assert any(not torch.allclose(trained[name], initial[name]) for name in initial), "training did not change the weights"

    # every rank writes its shards into the same directory: rank 0 picks it
    holder = [tempfile.mkdtemp() if accelerator.is_main_process else None]
    broadcast_object_list(holder)
    root = holder[0]
    accelerator.save_state(os.path.join(root, "checkpoint"))
    shard_dir = os.path.join(root, "checkpoint", "pytorch_model_fsdp_0")
    assert os.path.isdir(shard_dir) and any(name.endswith(".distcp") for name in os.listdir(shard_dir)), os.listdir(os.path.join(root, "checkpoint"))
    accelerator.wait_for_everyone()

    for safe in (True, False):
        # in-process API: collective in spirit (every rank may call it), only the main process writes
        out = os.path.join(root, f"merged_api_{'safe' if safe else 'bin'}")
        merge_fsdp_weights(shard_dir, out, safe_serialization=safe)
        accelerator.wait_for_everyone()
        if accelerator.is_main_process:
            got = merged_tensors(out, safe)
            assert set(got) == set(trained), (sorted(got), sorted(trained))
            assert all(torch.allclose(got[name].float(), trained[name], atol=1e-6) for name in trained), "merged weights differ from the trained ones"
        # command line, on the main process only (it needs no process group)
        if accelerator.is_main_process:
            out = os.path.join(root, f"merged_cli_{'safe' if safe else 'bin'}")
            command = [sys.executable, "-m", "accelerate_hpc_test_b200.commands.accelerate_cli", "merge-weights", shard_dir, out]
            if not safe:
                command.append("--unsafe_serialization")
            env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
            done = subprocess.run(command, capture_output=True, text=True, env=env)
            assert done.returncode == 0, done.stderr[-2000:]
            got = merged_tensors(out, safe)
            assert all(torch.allclose(got[name].float(), trained[name], atol=1e-6) for name in trained)
        accelerator.wait_for_everyone()
    if accelerator.is_main_process:
        shutil.rmtree(root, ignore_errors=True)
    accelerator.print("test_merge_weights: all passed")
    accelerator.end_training()


if __name__ == "__main__":
    main()
