import os, sys, time, torch
if "RANK" not in os.environ:
    print("peer access 0->1:", torch.cuda.can_device_access_peer(0, 1))
    a = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:0"); b = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:1")
    for _ in range(2): b.copy_(a); torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    t = time.perf_counter()
    for _ in range(5): b.copy_(a)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    print("cudaMemcpyPeer 1 GiB: %.1f GB/s" % (5 * 1.0737 / (time.perf_counter() - t)))
    sys.exit(0)
import torch.distributed as dist
r = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(r)
dist.init_process_group("nccl", device_id=torch.device("cuda", r))
for mb in (64, 256, 1024):
    n = mb << 20
    s = torch.empty(2 * n, dtype=torch.uint8, device="cuda"); o = torch.empty(2 * n, dtype=torch.uint8, device="cuda")
    for _ in range(2): dist.all_to_all_single(o, s)
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(5): dist.all_to_all_single(o, s)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 5
    if r == 0: print("all_to_all_single %d MiB per peer: %.2f ms, %.1f GB/s out per rank" % (mb, dt * 1e3, n / dt / 1e9))
    sl = [s[:n].view(torch.int64), s[n:].view(torch.int64)]; ol = [o[:n].view(torch.int64), o[n:].view(torch.int64)]
    for _ in range(2): dist.all_to_all(ol, sl)
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(5): dist.all_to_all(ol, sl)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 5
    if r == 0: print("all_to_all (lists)  %d MiB per peer: %.2f ms, %.1f GB/s out per rank" % (mb, dt * 1e3, n / dt / 1e9))
dist.destroy_process_group()
