#!/usr/bin/env python
"""Probe: does torch symmetric memory (peer pointers + NVSwitch multicast) come up on this box?  (torchrun, N >= 2)"""
import os
import sys
import time

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import torch.distributed._symmetric_memory as symm
    t0 = time.time()
    buf = symm.empty(64 << 20, dtype=torch.uint8, device=dev)
    h = symm.rendezvous(buf, dist.group.WORLD)
    torch.cuda.synchronize()
    info = dict(rank=rank, world=h.world_size, multicast=bool(h.has_multicast_support), mc_ptr=hex(h.multicast_ptr or 0),
                ptrs=[hex(p) for p in h.buffer_ptrs], secs=round(time.time() - t0, 2))
    print("SYMM", info, flush=True)
    # peer visibility through torch only: rank 0 fills its buffer, everyone reads rank 0's copy
    if rank == 0:
        buf.fill_(7)
    h.barrier()
    peer0 = h.get_buffer(0, (1024,), torch.uint8)
    ok = bool((peer0 == 7).all().item())
    h.barrier()
    print("SYMM_READ", rank, ok, flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
