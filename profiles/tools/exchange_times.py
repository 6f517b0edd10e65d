"""torchrun diagnostic: device time of each piece of the multi-GPU exchange at C2 size (P = 1M, one view per rank).
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 profiles/tools/exchange_times.py"""
import ctypes, os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from langsplat_b200 import dist as D, _lib as L

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
P = 1_000_000
ex = D.make_peer_exchange(P, 1, dev, grad_floats_per_gaussian=14)
assert ex is not None
lib = L.load()
means = torch.randn(P, 3, device=dev)
flat = torch.randn(14 * P, device=dev)
cam = torch.tensor([0.0, 0.0, 5.0], device=dev)

def timed(fn, n=20):
    for _ in range(3): fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())

def step_sh_only():
    s = ex.begin_step(); f = s.new_factor(P, dev, cam, 3); f.normal_()
    s.finish(means, 16)

def step_full_copy():
    s = ex.begin_step(); f = s.new_factor(P, dev, cam, 3); f.normal_()
    s.exchange(means, 16, [flat])

def step_full_inplace():
    s = ex.begin_step(); lead = s.new_flat(14 * P, dev); f = s.new_factor(P, dev, cam, 3); f.normal_(); lead.normal_()
    s.exchange(means, 16, [lead])

def fill_only():
    s = ex.begin_step(); lead = s.new_flat(14 * P, dev); f = s.new_factor(P, dev, cam, 3); f.normal_(); lead.normal_()

def nccl():
    dist.all_reduce(flat)

stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
def ar_only():  # barrier + all-reduce kernel + signal + wait on the last slot's staging area
    off = L.PEER_HEADER_BYTES + (ex.epoch & 1) * ex.slot_bytes
    ins = (ctypes.c_void_p * world)(*[b + off + ex.in_off * 4 for b in ex.bases])
    outs = (ctypes.c_void_p * world)(*[b + off + ex.out_off * 4 for b in ex.bases])
    ex._barriers += 1; lib.lsr_peer_barrier(rank, world, ex._headers, ex._barriers, stream)
    lib.lsr_peer_allreduce(rank, world, ins, outs, 14 * P, stream)
    ex._barriers += 1; lib.lsr_peer_signal(rank, world, ex._headers, ex._barriers, stream)
    lib.lsr_peer_wait(rank, world, ex._headers, ex._barriers, stream)

def bar_only():
    ex._barriers += 1; lib.lsr_peer_barrier(rank, world, ex._headers, ex._barriers, stream)

res = {}
for name, fn in (("barrier", bar_only), ("barrier+allreduce+signal+wait", ar_only), ("fill (normal_ x2, baseline)", fill_only),
                 ("fill + barrier + SH kernel", step_sh_only), ("fill + exchange (in place)", step_full_inplace),
                 ("randn + exchange (staging copy)", step_full_copy), ("NCCL all_reduce 56 MB", nccl)):
    res[name] = timed(fn)
if rank == 0:
    print("world", world)
    for k, v in res.items(): print("  %-36s %8.1f us" % (k, v))
ex.check(); ex.close()
dist.destroy_process_group()
