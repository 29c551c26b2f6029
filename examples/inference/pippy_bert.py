"""Pipeline-parallel inference of the native BERT classifier: embeddings on the first stage, encoder layers spread over the stages,
pooler + classifier on the last; the padding mask travels with the activations (reference examples/inference/pippy/bert.py).

    accelerate launch --num_processes 2 examples/inference/pippy_bert.py
    accelerate launch --num_processes 2 --cpu examples/inference/pippy_bert.py --cpu --tiny
"""

import argparse
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

import torch

from accelerate_hpc_test_b200 import PartialState, prepare_pippy
from accelerate_hpc_test_b200.models.bert import BertConfig, BertForSequenceClassification, pad_collate, synthetic_mrpc


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--tiny", action="store_true")
    p.add_argument("--batch", type=int, default=16)
    args = p.parse_args()
    state = PartialState(cpu=args.cpu)
    cfg = BertConfig(vocab_size=1000, hidden_size=64, num_hidden_layers=4, num_attention_heads=4, intermediate_size=128) if args.tiny else BertConfig.base_cased()
    torch.manual_seed(0)
    model = BertForSequenceClassification(cfg).eval()
    batch = pad_collate(synthetic_mrpc(args.batch, vocab_size=cfg.vocab_size, max_len=64, seed=1))
    batch.pop("labels")
    # every process keeps only its stage's layers; 4 micro-batches flow through the stages so that all of them work at once
    model = prepare_pippy(model, split_points="auto", example_kwargs=batch, num_chunks=4, gather_output=True)
    with torch.no_grad():
        model(**batch) if state.is_main_process else model()          # warm-up: stage buffers, headers
        state.wait_for_everyone()
        start = time.perf_counter()
        logits = model(**batch) if state.is_main_process else model()
        if state.device.type == "cuda":
            torch.cuda.synchronize()
    state.print(f"{args.batch} sentence pairs through {state.num_processes} stages in {(time.perf_counter() - start) * 1e3:.1f} ms; "
                f"predictions {logits.argmax(-1).tolist()}")
    state.wait_for_everyone()


if __name__ == "__main__":
    main()
