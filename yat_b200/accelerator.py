"""Minimal stand-in for the slice of ``accelerate.Accelerator`` the reference trainer uses
(common/trainer.py:31-37,252-253,317,344-347,359,377): one process per GPU, torch DDP over NCCL.

Differences from the reference on purpose: NCCL P2P/IB are NOT disabled, so the
gradient all-reduce runs over NVLink 5 / NVSwitch; DDP uses ``gradient_as_bucket_view`` and a static
graph (the step has no unused parameters) instead of ``find_unused_parameters=True``.
"""
import contextlib
import datetime
import os

import torch
import torch.distributed as dist


class Accelerator:
    def __init__(self, gradient_accumulation_steps: int = 1, backend: str = None, find_unused_parameters=False):
        self.gradient_accumulation_steps = int(gradient_accumulation_steps)
        self.find_unused_parameters = find_unused_parameters
        world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_process_index = int(os.environ.get("LOCAL_RANK", "0"))
        use_cuda = torch.cuda.is_available()
        if use_cuda:
            torch.cuda.set_device(self.local_process_index)
            self.device = torch.device("cuda", self.local_process_index)
        else:
            self.device = torch.device("cpu")
        if world > 1 and not dist.is_initialized():
            dist.init_process_group(backend=backend or ("nccl" if use_cuda else "gloo"),
                                    timeout=datetime.timedelta(seconds=3600))
        self.process_index = dist.get_rank() if dist.is_initialized() else 0
        self.num_processes = dist.get_world_size() if dist.is_initialized() else 1
        self.is_main_process = self.process_index == 0
        self.sync_gradients = True
        self._micro_step = 0

    # -- model / optimizer wrapping
    def prepare(self, *objs):
        out = []
        for o in objs:
            if isinstance(o, torch.nn.Module) and self.num_processes > 1:
                o = torch.nn.parallel.DistributedDataParallel(
                    o, device_ids=[self.local_process_index] if self.device.type == "cuda" else None,
                    gradient_as_bucket_view=True, find_unused_parameters=self.find_unused_parameters,
                    static_graph=(not self.find_unused_parameters) and self.gradient_accumulation_steps == 1,
                    bucket_cap_mb=256)
            out.append(o)
        return out[0] if len(out) == 1 else tuple(out)

    def unwrap_model(self, model):
        return model.module if isinstance(model, torch.nn.parallel.DistributedDataParallel) else model

    @contextlib.contextmanager
    def accumulate(self, model):
        self._micro_step += 1
        self.sync_gradients = (self._micro_step % self.gradient_accumulation_steps) == 0
        if not self.sync_gradients and isinstance(model, torch.nn.parallel.DistributedDataParallel):
            with model.no_sync():
                yield
        else:
            yield

    def backward(self, loss):
        if self.gradient_accumulation_steps > 1:
            loss = loss / self.gradient_accumulation_steps
        loss.backward()

    def clip_grad_norm_(self, parameters, max_norm):
        return torch.nn.utils.clip_grad_norm_(parameters, max_norm, foreach=True)

    # -- collectives
    def wait_for_everyone(self):
        if self.num_processes > 1:
            dist.barrier()

    def gather(self, tensor):
        if self.num_processes == 1:
            return tensor.reshape(1, *tensor.shape) if tensor.dim() == 0 else tensor
        t = tensor.reshape(1) if tensor.dim() == 0 else tensor
        out = [torch.empty_like(t) for _ in range(self.num_processes)]
        dist.all_gather(out, t.contiguous())
        return torch.cat(out, dim=0)

    def reduce(self, tensor, reduction="mean"):
        if self.num_processes == 1:
            return tensor
        t = tensor.clone()
        dist.all_reduce(t)
        return t / self.num_processes if reduction == "mean" else t
