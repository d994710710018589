"""torchrun check of DistributedJoin.probe_chunks on real GPUs (one rank per GPU): every pair of every chunk must join equal
keys (checked on the rank that produced it, against the rows it received), and the pair count over all ranks must be the
1:1 workload's probe row count.  torchrun --nproc-per-node N scripts/check_dist.py [--mode p2p|nccl]"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import datafusion_parallelism_b200 as d
from datafusion_parallelism_b200.exchange import DistributedJoin, _RawCudaArray

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="p2p")
ap.add_argument("--build", type=int, default=5_000_000)
ap.add_argument("--probe", type=int, default=40_000_000)
ap.add_argument("--chunk", type=int, default=6_000_000)
ap.add_argument("--hot-share", type=float, default=0.0, help="this fraction of every rank's probe rows carries ONE key (extreme heavy hitter)")
ap.add_argument("--zipf", action="store_true", help="Zipf(1) probe keys: heavy hitters, routed with enable_hot_routing on the second pass")
a = ap.parse_args()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
g = torch.Generator(device=dev)
g.manual_seed(11 + rank)
bk = torch.randperm(a.build, device=dev, generator=g, dtype=torch.int64) * world + rank       # unique over all ranks
pk = torch.randint(0, a.build * world, (a.probe,), device=dev, generator=g, dtype=torch.int64)
if a.zipf:
    import math
    u = torch.rand(a.probe, device=dev, generator=g, dtype=torch.float64)
    ks = a.build * world
    pk = (torch.exp(u * math.log(ks + 0.5)) - 0.5).clamp_(0, ks - 1).to(torch.int64)
    pk = (pk * 2654435761) % ks                       # hot keys spread over the key space (and over the owners)
if a.hot_share > 0:
    m = torch.rand(a.probe, device=dev, generator=g) < a.hot_share
    pk = torch.where(m, torch.full_like(pk, 12345 * world + 1), pk)
    a.zipf = True                                     # second pass routes the heavy hitter
stream = torch.cuda.Stream(device=dev)
j = DistributedJoin("inner", key_width=8, stream=stream, mode=a.mode, layout="radix", radix_bits=5)
for rep in range(2):                                    # second pass re-uses the receive buffers
    j.reset()
    j.build_push(0, [bk])
    j.build_finish(0)
    n_b = j.build_rows_local
    bkeys = torch.as_tensor(_RawCudaArray(j.build.recv_keys_ptr, n_b, "<i8"), device=dev).clone()
    lo_of = [0]
    n_hot_keys = j.enable_hot_routing(pk) if (a.zipf and rep == 1) else 0
    tot, recv_rows, hot_rows = 0, 0, 0
    torch.cuda.synchronize(); dist.barrier()
    import time
    t_start = time.perf_counter()
    for r in j.probe_chunks(0, [pk[lo:lo + a.chunk] for lo in range(0, a.probe, a.chunk)], want=d.OUT_PAIRS):
        side = j.probe
        n_recv = int(side.plan.recv_rows[rank])
        pkeys = torch.as_tensor(_RawCudaArray(side.recv_keys_ptr, n_recv, "<i8"), device=dev)
        b_, p_ = r.device_arrays()
        bi = torch.as_tensor(b_, device=dev).to(torch.int64)
        pi = torch.as_tensor(p_, device=dev).to(torch.int64) & 0xffffffff
        assert r.n == n_recv, (r.n, n_recv)
        assert bool((bkeys[bi] == pkeys[pi]).all()), "a pair joins different keys"
        assert int(torch.unique(pi).numel()) == n_recv, "a probe row is missing or doubled"
        tot += r.n
        recv_rows += n_recv
        if r.hot is not None:
            hb_, hp_ = r.hot.device_arrays()
            if r.hot.n:
                hb = torch.as_tensor(hb_, device=dev).to(torch.int64)
                hp = torch.as_tensor(hp_, device=dev).to(torch.int64) & 0xffffffff
                assert bool((j.hot_build_keys[hb] == side.hot_keys[: side.n_hot][hp]).all()), "a hot pair joins different keys"
                assert int(torch.unique(hp).numel()) == side.n_hot == r.hot.n, "a hot probe row is missing or doubled"
                assert bool((pk[lo_of[0] + r.hot_rid.to(torch.int64)[hp]] == side.hot_keys[: side.n_hot][hp]).all()), "hot row id does not point at its row"
            tot += r.hot.n
            hot_rows += r.hot.n
            r.hot.release()
        lo_of[0] += a.chunk
        r.release()
    torch.cuda.synchronize(); dist.barrier()
    ms = (time.perf_counter() - t_start) * 1e3
    t = torch.tensor([tot], device=dev, dtype=torch.int64)
    dist.all_reduce(t)
    assert int(t.item()) == a.probe * world, (int(t.item()), a.probe * world)
    rr = torch.tensor([recv_rows + hot_rows], device=dev, dtype=torch.int64)
    allr = [torch.zeros_like(rr) for _ in range(world)]
    dist.all_gather(allr, rr)
    if rank == 0:
        loads = [int(x.item()) for x in allr]
        print(f"pass {rep}: hot keys {n_hot_keys}, rows joined per rank {loads} (max/mean {max(loads) / (sum(loads) / world):.3f}), probe side {ms:.1f} ms wall")
# host batches: pinned keys in, pinned pairs out, copies pipelined around the exchange + join (probe_host_chunks)
j.reset()
hbk = torch.empty(a.build, dtype=torch.int64).pin_memory(); hbk.copy_(bk)
hpk = torch.empty(a.probe, dtype=torch.int64).pin_memory(); hpk.copy_(pk)
with torch.cuda.stream(stream):
    j.build_push_host(0, hbk.numpy())
    j.build_finish(0)
    bkeys_h = torch.as_tensor(_RawCudaArray(j.build.recv_keys_ptr, j.build_rows_local, "<i8"), device=dev).cpu().numpy()
    for want_h in (d.OUT_PAIRS, d.OUT_PAIRS | d.OUT_BUILD_U32):          # (u64, u32) pairs, then 4-byte build indices
        tot = 0
        for r in j.probe_host_chunks(0, hpk.numpy(), chunk_rows=a.chunk, want=want_h):
            assert r.host_build_idx.dtype == (np.uint32 if want_h & d.OUT_BUILD_U32 else np.uint64)
            side = r.side
            n_recv = int(side.plan.recv_rows[rank])
            pkeys_h = torch.as_tensor(_RawCudaArray(side.recv_keys_ptr, n_recv, "<i8"), device=dev).cpu().numpy()
            assert r.n == n_recv == r.host_build_idx.shape[0] == r.host_probe_idx.shape[0]
            assert (bkeys_h[r.host_build_idx.astype("int64")] == pkeys_h[r.host_probe_idx.astype("int64")]).all(), "host pair joins different keys"
            src_rank, src_row = j.origin_probe(r.host_probe_idx[:1000], side=side)
            for sr in range(world):                      # a pair's origin row on THIS rank really holds that key
                if sr == rank:
                    m = src_rank == sr
                    rows = r.chunk_lo + src_row[m].astype("int64")
                    assert (hpk.numpy()[rows] == pkeys_h[r.host_probe_idx[:1000].astype("int64")][m]).all(), "origin_probe points at the wrong row"
            tot += r.n
t = torch.tensor([tot], device=dev, dtype=torch.int64)
dist.all_reduce(t)
assert int(t.item()) == a.probe * world, (int(t.item()), a.probe * world)
if rank == 0:
    print(f"host chunks ok: {int(t.item())} pairs")
if rank == 0:
    print(f"check_dist ok: {world} ranks, mode {a.mode}, {a.probe * world} pairs, every pair key-equal")
j.close()
dist.barrier()
dist.destroy_process_group()
