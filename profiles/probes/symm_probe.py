"""Probe: does torch symmetric memory give a multicast (NVLS) pointer on this box?
torchrun --nproc-per-node 2 profiles/probes/symm_probe.py"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
t = symm_mem.empty(1 << 20, dtype=torch.float32, device=dev)
t.zero_()
hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
print(f"rank {rank}: has_multicast_support={hdl.has_multicast_support} multicast_ptr={hdl.multicast_ptr:#x} "
      f"buffer_ptrs={[hex(p) for p in hdl.buffer_ptrs]} signal_pad_size={hdl.signal_pad_size}", flush=True)
hdl.barrier(channel=0)
# peer write through get_buffer
peer = hdl.get_buffer((rank + 1) % world, (16,), torch.float32)
peer.fill_(float(rank + 1))
hdl.barrier(channel=0)
torch.cuda.synchronize()
print(f"rank {rank}: first elems after peer write {t[:4].tolist()}", flush=True)
dist.destroy_process_group()
