import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
rank = int(os.environ["RANK"]); torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", torch.cuda.current_device())
dist.init_process_group("nccl", device_id=dev)
try:
    t = symm_mem.empty(1 << 20, dtype=torch.float32, device=dev)
    t.fill_(rank + 1)
    h = symm_mem.rendezvous(t, dist.group.WORLD)
    print(rank, "rendezvous ok", "buffers", [hex(p) for p in h.buffer_ptrs], "signal", [hex(p) for p in h.signal_pad_ptrs][:2],
          "multicast", hex(h.multicast_ptr) if h.multicast_ptr else None,           "signal_pad_size", h.signal_pad_size, "buffer_size", h.buffer_size, flush=True)
    h.barrier()
    peer = h.get_buffer((rank + 1) % dist.get_world_size(), (4,), torch.float32)
    print(rank, "peer first values", peer.tolist(), flush=True)
except Exception as e:
    print(rank, "FAILED", type(e).__name__, str(e)[:500], flush=True)
dist.barrier(); dist.destroy_process_group()
