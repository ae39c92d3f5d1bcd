"""`DistributedDataParallel` for the drop-in classes (`pretrain.py:166`, `finetune.py:137`).

torch's DDP reduces gradients from per-parameter autograd hooks; the K3M classes run the whole step as ONE autograd
node over a flat gradient arena, so those hooks never fire.  This shim keeps the driver edit to one import:

    from k3m_b200.parallel import DistributedDataParallel        # instead of torch.nn.parallel
    model = DistributedDataParallel(model, device_ids=[local_rank])

It registers the process group with the model (backward then all-reduces the gradient arena and divides by the world
size: DDP's mean), exposes `.module`, forwards calls / attribute access, and prefixes `state_dict()` keys with
`module.` like DDP does (`pretrain.py:1238-1251` strips it again)."""
import torch
from torch import nn


class DistributedDataParallel(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, process_group=None, **unused):
        super(DistributedDataParallel, self).__init__()
        if not (torch.distributed.is_available() and torch.distributed.is_initialized()):
            raise RuntimeError("DistributedDataParallel needs torch.distributed.init_process_group() first")
        self.module = module
        group = process_group if process_group is not None else torch.distributed.group.WORLD
        module.set_process_group(group)
        # same starting point on every rank, as DDP's constructor broadcast guarantees
        with torch.no_grad():
            torch.distributed.broadcast(module._arena, src=torch.distributed.get_global_rank(group, 0), group=group)

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def __getattr__(self, name):
        try:
            return super(DistributedDataParallel, self).__getattr__(name)
        except AttributeError:
            return getattr(self.module, name)
