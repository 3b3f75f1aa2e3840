"""Import shim for ``accelerate`` (absent from this image, no network): the minimal ``Accelerator`` the
reference's training engine touches — constructor kwargs ``mixed_precision`` / ``gradient_accumulation_steps``
and the members ``is_main_process, num_processes, prepare, backward, clip_grad_norm_, save_model``
(reference llmctl/runtime/engine.py:107-117,265-267,289,297,378).  It reproduces what accelerate does on that
path: NCCL process group from the torchrun environment, ``DistributedDataParallel`` when world > 1, bf16
autocast around the model forward, batches moved to the device.  Nothing from llmctl_b200 is used here.
"""
import os

import torch
import torch.distributed as dist

__version__ = "0.0.0"


class _DeviceLoader:
    def __init__(self, loader, device):
        self.loader, self.device = loader, device

    def __len__(self):
        return len(self.loader)

    def __iter__(self):
        for batch in self.loader:
            if isinstance(batch, dict):
                yield {k: (v.to(self.device, non_blocking=True) if torch.is_tensor(v) else v) for k, v in batch.items()}
            else:
                yield batch


class Accelerator:
    def __init__(self, mixed_precision=None, gradient_accumulation_steps=1, **kwargs):
        self.mixed_precision = mixed_precision
        self.gradient_accumulation_steps = gradient_accumulation_steps
        world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if torch.cuda.is_available():
            torch.cuda.set_device(self.local_rank)
            self.device = torch.device("cuda", self.local_rank)
        else:
            self.device = torch.device("cpu")
        if world > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            dist.init_process_group("nccl" if self.device.type == "cuda" else "gloo")
        self.num_processes = dist.get_world_size() if dist.is_initialized() else 1
        self.process_index = dist.get_rank() if dist.is_initialized() else 0
        self.is_main_process = self.process_index == 0

    def _wrap_model(self, model):
        model = model.to(self.device)
        dtype = {"bf16": torch.bfloat16, "fp16": torch.float16}.get(self.mixed_precision)
        if dtype is not None and self.device.type == "cuda":
            fwd = model.forward

            def autocast_forward(*a, **k):
                with torch.autocast("cuda", dtype=dtype):
                    return fwd(*a, **k)
            model.forward = autocast_forward
        if self.num_processes > 1:
            model = torch.nn.parallel.DistributedDataParallel(
                model, device_ids=[self.local_rank] if self.device.type == "cuda" else None)
        return model

    def prepare(self, *objs):
        out = []
        for o in objs:
            if isinstance(o, torch.nn.Module):
                out.append(self._wrap_model(o))
            elif isinstance(o, torch.optim.Optimizer) or o is None:
                out.append(o)
            elif hasattr(o, "__iter__"):
                out.append(_DeviceLoader(o, self.device))
            else:
                out.append(o)
        return tuple(out) if len(out) > 1 else out[0]

    def backward(self, loss, **kwargs):
        loss.backward(**kwargs)

    def clip_grad_norm_(self, parameters, max_norm, norm_type=2):
        return torch.nn.utils.clip_grad_norm_(parameters, max_norm, norm_type)

    def unwrap_model(self, model):
        return model.module if hasattr(model, "module") else model

    def save_model(self, model, save_directory, **kwargs):
        if os.environ.get("LLMCTL_REF_SKIP_SAVE") == "1":
            return
        from safetensors.torch import save_file
        os.makedirs(save_directory, exist_ok=True)
        sd = {k: v.detach().cpu().contiguous() for k, v in self.unwrap_model(model).state_dict().items()}
        save_file(sd, os.path.join(save_directory, "model.safetensors"))

    def wait_for_everyone(self):
        if dist.is_initialized():
            dist.barrier()
