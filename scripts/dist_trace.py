"""Phase timing of the multi-GPU sort (torchrun). Synchronises after every phase, so overlap is lost:
the numbers show what each phase costs on its own."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from ray_project_ray_b200 import distributed as D, ops
from ray_project_ray_b200.block import DeviceBlock, DeviceBlockAccessor
from ray_project_ray_b200.sort_task_spec import SortKey

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
PAY = int(sys.argv[2]) if len(sys.argv) > 2 else 64   # payload bytes per row (90: gensort)
ITERS = int(sys.argv[3]) if len(sys.argv) > 3 else 3
g = torch.Generator(device=dev).manual_seed(rank)
k = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
p = torch.empty((n, PAY), dtype=torch.uint8, device=dev)
local = DeviceBlock({"id": k, "payload": p})
sk = SortKey("id")
def T(label, t0):
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t = time.perf_counter()
    if rank == 0: print(f"{label:34s} {(t - t0) * 1e3:8.2f} ms", flush=True)
    return time.perf_counter()
for it in range(ITERS):
    torch.cuda.synchronize(); dist.barrier(); t = time.perf_counter()
    if rank == 0: print("---- iteration", it)
    assert D.p2p_available(dev)
    t = T("p2p_available", t)
    m = 4096
    pos = ((np.arange(m, dtype=np.float64) + 0.5) * (n / m)).astype(np.int64).astype(np.int32)
    sc_ = ops.encode_keys(ops.gather(k, torch.from_numpy(pos).to(dev)), False)
    allc = D._gather_samples(sc_, 4096, world, dev, None); spl = D.choose_splitters(allc, world)
    sp = torch.from_numpy(spl.view(np.int64)).to(dev); t = T("samples + splitters", t)
    target, counts = ops.partition_targets(k, False, sp); t = T("partition_targets", t)
    sc = [int(c) for c in counts.cpu().tolist()[:world]]
    cols = {"id": (k, None), "payload": (p, None)}
    ex = D.P2PExchange(cols, n, sc, dev, target=target); t = T("P2PExchange setup (counts, buffers)", t)
    ex.fence(); t = T("fence", t)
    ex.send(["id"]); t = T("send keys", t)
    ex.fence(); t = T("fence", t)
    kb = DeviceBlock(ex.view(["id"]), num_rows=ex.n_recv, device=dev)
    perm2, _, sk2 = DeviceBlockAccessor(kb)._argsort(sk, False); t = T("argsort received keys", t)
    ex.send(["payload"]); t = T("send payload", t)
    ex.fence(); t = T("fence", t)
    out = ops.gather(ex.view(["payload"])["payload"], perm2); t = T("final payload gather", t)
# overlap experiment: received-key sort on a side stream while the payload scatter runs on the main stream
side = D._side_stream(dev)
from ray_project_ray_b200 import _lib
for bps in ((0, 0, 1, 2, 4) if len(sys.argv) <= 4 else ()):
    _lib.call("b200x_set_tuning", 1, bps)
    torch.cuda.synchronize(); dist.barrier(); t = time.perf_counter()
    ex.fence(); ex.send(["payload"]); ex.fence()
    t = T(f"payload send alone, blocks/SM={bps}", t)
    torch.cuda.synchronize(); dist.barrier(); t = time.perf_counter()
    ex.fence()
    ev = torch.cuda.Event(); ev.record()
    with torch.cuda.stream(side):
        side.wait_event(ev)
        perm2, _, sk2 = DeviceBlockAccessor(kb)._argsort(sk, False)
    ex.send(["payload"]); ex.fence()
    side.synchronize()
    t = T(f"CONCURRENT sort || payload send, blocks/SM={bps}", t)
dist.destroy_process_group()
