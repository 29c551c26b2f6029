"""Sharded training with memory accounting: a context manager records the peak device and host memory of the training and of the
evaluation phase of every epoch, and the numbers go to the tracker next to the metrics.  Sharding is what makes them interesting: with
``--use_fsdp`` (or an FSDP config from ``accelerate config``) parameters, gradients and optimizer state are divided by the number of
ranks (reference examples/by_feature/fsdp_with_peak_mem_tracking.py).

    accelerate launch --num_processes 2 --cpu examples/by_feature/fsdp_with_peak_mem_tracking.py --model tiny --use_fsdp --with_tracking
"""

import gc
import os
import sys
import threading

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import psutil
import torch

from _common import base_parser, build, default_config, evaluate

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.utils import FullyShardedDataParallelPlugin


# New Code #
def b2mb(x):
    return int(x / 2**20)


# New Code #
class TorchTracemalloc:
    """Peak memory of the ``with`` block: device memory from the caching allocator's statistics, host memory by polling the
    resident set size from a thread (there is no high-water mark to ask the OS for)."""

    def __enter__(self):
        gc.collect()
        self.on_gpu = torch.cuda.is_available()
        if self.on_gpu:
            torch.cuda.empty_cache()
            torch.cuda.reset_peak_memory_stats()
            self.begin = torch.cuda.memory_allocated()
        self.process = psutil.Process()
        self.cpu_begin = self.process.memory_info().rss
        self.cpu_peak = self.cpu_begin
        self._stop = threading.Event()
        self._poller = threading.Thread(target=self._poll, daemon=True)
        self._poller.start()
        return self

    def _poll(self):
        while not self._stop.wait(0.01):
            self.cpu_peak = max(self.cpu_peak, self.process.memory_info().rss)

    def __exit__(self, *exc):
        self._stop.set()
        self._poller.join()
        gc.collect()
        if self.on_gpu:
            torch.cuda.empty_cache()
            self.used = b2mb(torch.cuda.memory_allocated() - self.begin)
            self.peaked = b2mb(torch.cuda.max_memory_allocated() - self.begin)
        else:
            self.used = self.peaked = 0
        self.cpu_used = b2mb(self.process.memory_info().rss - self.cpu_begin)
        self.cpu_peaked = b2mb(max(self.cpu_peak, self.process.memory_info().rss) - self.cpu_begin)


def training_function(config, args):
    # New Code #
    plugin = None
    if args.use_fsdp:
        plugin = FullyShardedDataParallelPlugin(fsdp_version=2, auto_wrap_policy="transformer_based_wrap",
                                                transformer_cls_names_to_wrap=["BertLayer"])
    accelerator = Accelerator(cpu=args.cpu, mixed_precision=args.mixed_precision, fsdp_plugin=plugin,
                              log_with="jsonl" if args.with_tracking else None, project_dir=args.project_dir)
    if args.with_tracking:
        accelerator.init_trackers("fsdp_with_peak_mem_tracking", config)
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = build(accelerator, config, args)
    # with a sharded engine the model and the optimizer go through `prepare` together: the optimizer has to end up holding the shards
    model, optimizer, train_dataloader, eval_dataloader, lr_scheduler = accelerator.prepare(
        model, optimizer, train_dataloader, eval_dataloader, lr_scheduler
    )
    for epoch in range(int(config["num_epochs"])):
        # New Code #
        with TorchTracemalloc() as train_mem:
            model.train()
            total_loss = 0.0
            for batch in train_dataloader:
                loss = model(**batch).loss
                accelerator.backward(loss)
                optimizer.step()
                lr_scheduler.step()
                optimizer.zero_grad()
                total_loss += float(loss.detach())
        accelerator.print(f"epoch {epoch} train: device memory used {train_mem.used} MB, peak {train_mem.peaked} MB; "
                          f"host used {train_mem.cpu_used} MB, peak {train_mem.cpu_peaked} MB")
        # New Code #
        with TorchTracemalloc() as eval_mem:
            metrics = evaluate(accelerator, model, eval_dataloader)
        accelerator.print(f"epoch {epoch}:", metrics)
        if args.with_tracking:
            accelerator.log({"accuracy": metrics["accuracy"], "train_loss": total_loss / max(1, len(train_dataloader)),
                             "train_peak_device_mb": train_mem.peaked, "train_peak_host_mb": train_mem.cpu_peaked,
                             "eval_peak_device_mb": eval_mem.peaked, "eval_peak_host_mb": eval_mem.cpu_peaked}, step=epoch)
    accelerator.end_training()


def main():
    parser = base_parser()
    parser.add_argument("--use_fsdp", action="store_true", help="Shard parameters, gradients and optimizer state over the ranks.")
    parser.add_argument("--with_tracking", action="store_true", help="Whether to load in all available experiment trackers from the environment and use them for logging.")
    parser.add_argument("--project_dir", type=str, default="logs", help="Location on where to store experiment tracking logs`")
    args = parser.parse_args()
    training_function(default_config(args), args)


if __name__ == "__main__":
    main()
