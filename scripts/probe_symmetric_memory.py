"""Does torch's symmetric memory rendezvous work on this box, and does it expose a multicast (NVLS) mapping?  Run under
torchrun, one rank per GPU; the answer decides whether extra={"residual_allreduce": "nvls"} is available
(solver.py::_ensure_plan).  Diagnostic only -- nothing imports this."""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as sm
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
try:
    t = sm.empty(4096, dtype=torch.float32, device=torch.device("cuda", rank))
    h = sm.rendezvous(t, dist.group.WORLD.group_name)
    print(rank, "rendezvous ok; multicast_ptr", hex(h.multicast_ptr) if h.multicast_ptr else 0, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs][:2],
          "signal_pad", len(h.signal_pad_ptrs), flush=True)
except Exception as e:
    print(rank, "symm_mem failed:", repr(e)[:300], flush=True)
dist.destroy_process_group()
