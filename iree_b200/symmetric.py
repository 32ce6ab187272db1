"""Host-side plumbing for symmetric regions: allocation and peer / multicast mappings come
from torch.distributed._symmetric_memory (CUDA VMM + NVSwitch multicast objects exchanged
between the ranks of a torch process group); the pointers are handed to the C-ABI
(iree_b200_channel_attach_symmetric_region, include/iree_b200_hal.h). Nothing here touches
the data path: the all-reduce itself is kernels/collective_kernels.cu.
"""

from __future__ import annotations

from . import hal


class SymmetricRegion:
    """|data_bytes| of symmetric device memory followed by one zeroed flag block per channel
    that will be attached (|flag_blocks|)."""

    def __init__(self, data_bytes: int, device, group=None, flag_blocks: int = 1):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        group = group or dist.group.WORLD
        self.data_bytes = (data_bytes + 1023) // 1024 * 1024
        self.flag_blocks = flag_blocks
        total = self.data_bytes + flag_blocks * hal.SYMMETRIC_FLAG_BYTES
        try:
            symm_mem.enable_symm_mem_for_group(group.group_name)
        except Exception:
            pass  # newer torch: not needed (deprecated no-op)
        self.tensor = symm_mem.empty(total, dtype=torch.uint8, device=device)
        self.tensor.zero_()
        torch.cuda.synchronize(device)
        self.handle = symm_mem.rendezvous(self.tensor, group)
        self.handle.barrier()
        self.rank = self.handle.rank
        self.world = self.handle.world_size
        self.peer_bases = [int(p) for p in self.handle.buffer_ptrs]
        mc = 0
        try:
            if self.handle.has_multicast_support:
                mc = int(self.handle.multicast_ptr or 0)
        except Exception:
            mc = 0
        self.multicast_base = mc
        assert self.peer_bases[self.rank] == self.tensor.data_ptr()

    @property
    def local_base(self) -> int:
        return self.peer_bases[self.rank]

    def view(self, offset: int, shape, dtype):
        """A torch view of the local copy: |shape| elements of |dtype| at byte |offset|."""
        import math
        n = math.prod(shape) * dtype.itemsize
        assert offset % 16 == 0 and offset + n <= self.data_bytes
        return self.tensor[offset:offset + n].view(dtype).view(*shape)

    def attach(self, channel: "hal.Channel", executable, flag_block: int = 0,
               use_multicast: bool = True, max_blocks: int = 0):
        assert 0 <= flag_block < self.flag_blocks
        off = self.data_bytes + flag_block * hal.SYMMETRIC_FLAG_BYTES
        channel.attach_symmetric_region(
            executable, self.data_bytes, self.peer_bases, [b + off for b in self.peer_bases],
            multicast_base=self.multicast_base if use_multicast else 0, max_blocks=max_blocks)
