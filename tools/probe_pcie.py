"""Developer probe (run under torchrun, one rank per GPU): the copy-engine route without the kernel
(cm_pipe_probe_copies) for several input sizes, all ranks at once -- is the host/PCIe ceiling of the
pipelined path set by bytes or by the number of copies?"""
import ctypes, os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, ".")
from cassmantle_b200 import _lib
from cassmantle_b200.engine import ScoringEngine, ScoringPipe

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
lib = _lib.load()
eng = ScoringEngine.synthetic(10_000, 300, 0, local)
B = 4096
for n_slots, in_flight in [(16, 8), (12, 8), (8, 8), (4, 8), (16, 16), (16, 4)]:
    pipe = ScoringPipe(eng, depth=32, max_guesses=B, max_slots=n_slots)
    for s in range(32):
        v = pipe.buffers(s, B, n_slots)
        v["guess_idx"][:] = 1
    secs = ctypes.c_double(0)
    for rep in range(2):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        _lib.check(lib.cm_pipe_probe_copies(pipe._h, 4000, in_flight, ctypes.byref(secs)))
    t = torch.tensor([secs.value], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    in_bytes = B * n_slots * 4 + B * 9
    if rank == 0:
        us = t.item() / 4000 * 1e6
        print(f"world {world}: {B}x{n_slots} ({in_bytes/1e3:.0f} KB in + 16 KB out), {in_flight} in flight: {us:.2f} us/step = "
              f"{(in_bytes + B*4)/us/1e3:.1f} GB/s per GPU", flush=True)
    pipe.close()
eng.close()
if world > 1:
    dist.destroy_process_group()
