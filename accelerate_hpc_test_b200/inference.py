"""Pipeline-parallel inference: ``prepare_pippy`` (reference src/accelerate/inference.py:31-186).

The model is traced and cut at split points (by default the boundaries ``infer_auto_device_map`` finds
for an even spread over the ranks), rank *r* keeps stage *r*, micro-batches flow through a GPipe
schedule whose activations travel over NCCL point-to-point sends (this path
stays on NCCL p2p by specification).  Rank 0 feeds the inputs, the last rank holds the output; with
``gather_output=True`` it is copied to every rank.
"""

from __future__ import annotations

import math
from types import MethodType
from typing import Any, Optional, Union

import torch

from .state import PartialState
from .utils.modeling import calculate_maximum_sizes, infer_auto_device_map
from .utils.operations import copy_tensor_to_devices, ignorant_find_batch_size, pad_input_tensors, send_to_device
from .utils.other import convert_bytes


def generate_device_map(model, num_processes: int = 1, no_split_module_classes=None, max_memory: dict = None):
    """A device map that spreads ``model`` evenly over ``num_processes`` virtual devices."""
    if num_processes == 1:
        return infer_auto_device_map(model, no_split_module_classes=no_split_module_classes, clean_result=False)
    if max_memory is None:
        model_size, shared = calculate_maximum_sizes(model)
        # split the model into `num_processes` parts, keeping room for the largest un-splittable layer
        memory = (model_size + shared[0]) / num_processes
        memory = convert_bytes(memory)
        value, ending = memory.split(" ")
        memory = math.ceil(float(value)) * 1.1
        memory = f"{memory} {ending}"
        max_memory = {i: memory for i in range(num_processes)}
    return infer_auto_device_map(model, max_memory=max_memory, no_split_module_classes=no_split_module_classes, clean_result=False)


def find_pippy_batch_size(args, kwargs):
    found_batch_size = None
    if args is not None:
        for arg in args:
            found_batch_size = ignorant_find_batch_size(arg)
            if found_batch_size is not None:
                break
    if kwargs is not None and found_batch_size is None:
        for kwarg in kwargs.values():
            found_batch_size = ignorant_find_batch_size(kwarg)
            if found_batch_size is not None:
                break
    return found_batch_size


def build_pipeline(model, split_points, args, kwargs, num_chunks):
    """Trace ``model`` with torch.distributed.pipelining, cut it at ``split_points``, keep this rank's stage."""
    from torch.distributed.pipelining import ScheduleGPipe, SplitPoint, pipeline

    state = PartialState()
    split_spec = {split_point: SplitPoint.BEGINNING for split_point in split_points}
    pipe = pipeline(model, mb_args=args, mb_kwargs=kwargs, split_spec=split_spec)
    stage = pipe.build_stage(state.local_process_index, device=state.device)
    return ScheduleGPipe(stage, num_chunks)


def pippy_forward(forward, num_chunks, gather_output, *args, **kwargs):
    state = PartialState()
    output = None
    if state.num_processes == 1:
        output = forward(*args, **kwargs)
    elif state.is_local_main_process:
        found_batch_size = find_pippy_batch_size(args, kwargs)
        if found_batch_size is None:
            raise ValueError("Could not find batch size from args or kwargs")
        if found_batch_size != num_chunks:
            args = pad_input_tensors(args, found_batch_size, num_chunks)
            kwargs = pad_input_tensors(kwargs, found_batch_size, num_chunks)
        forward(*args, **kwargs)
    elif state.is_last_process:
        output = forward()
    else:
        forward()
    if gather_output:
        # each rank needs the final output: broadcast it from the last stage
        output = copy_tensor_to_devices(output)
    return output


def prepare_pippy(model, split_points: Optional[Union[str, list]] = "auto", no_split_module_classes: Optional[list] = None,
                  example_args: Optional[tuple] = (), example_kwargs: Optional[dict] = None, num_chunks: Optional[int] = None,
                  gather_output: Optional[bool] = False):
    """Wrap ``model`` for pipeline-parallel inference over all processes; ``model(...)`` then runs the schedule."""
    state = PartialState()
    example_args = send_to_device(example_args, "cpu")
    example_kwargs = send_to_device(example_kwargs, "cpu")
    if num_chunks is None:
        num_chunks = state.num_processes
    if split_points == "auto":
        device_map = generate_device_map(model, num_chunks, no_split_module_classes=no_split_module_classes)
        split_points = []
        for i in range(1, num_chunks):
            split_points.append(next(k for k, v in device_map.items() if v == i))
    model.hf_split_points = split_points
    stage = build_pipeline(model, split_points, example_args, example_kwargs, num_chunks)
    model._original_forward = model.forward
    model._original_call = model.__call__
    model.pippy_stage = stage
    model.hf_split_points = split_points

    def forward(*args, **kwargs):
        return pippy_forward(stage.step, num_chunks, gather_output, *args, **kwargs)

    # a closure keeps the stage alive and the signature generic
    model_forward = MethodType(forward, model)
    forward.__wrapped__ = model_forward
    model.forward = forward
    return model
