"""torchrun check of heavy-hitter routing for the LEFT-family joins (left, full, left_semi, left_anti) on real GPUs.

The build rows of hot keys are probed on every rank against their broadcast copy; DistributedJoin.probe_finish OR-reduces
the "matched" flags and only the owner emits (or withholds) those rows. Checked here against counts computed from the
all-gathered keys: number of pairs, of NULL-build rows (full), and the finaliser batch row by row (distinct build rows, every
one satisfying the join type's condition, and as many as the rank owns).
  torchrun --nproc-per-node N scripts/check_dist_left.py"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import datafusion_parallelism_b200 as d
from datafusion_parallelism_b200.exchange import DistributedJoin, _RawCudaArray

ap = argparse.ArgumentParser()
ap.add_argument("--build", type=int, default=200_000)
ap.add_argument("--probe", type=int, default=3_000_000)
ap.add_argument("--chunk", type=int, default=1_100_000)
ap.add_argument("--joins", default="left,full,left_semi,left_anti,inner")
a = ap.parse_args()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
g = torch.Generator(device=dev)
g.manual_seed(101 + rank)
KS = 100_000                      # id space; the uniform probes only cover its lower half, so half of the build rows stay unmatched
HOT, HOT1, HOT_NOBUILD = 4242, 777, KS + 77
# HOT: probe rows on rank 0 only, HOT1: on the last rank only (their build rows live on every rank, so a rank that saw no probe
# row of the key reports them unmatched and only the OR over the ranks gets it right); HOT_NOBUILD: no build row anywhere
bk = torch.randint(0, KS, (a.build,), device=dev, generator=g, dtype=torch.int64)
bk[:5] = HOT                      # every rank owns five build rows of each hot key
bk[5:8] = HOT1
pk = torch.randint(0, KS // 2, (a.probe,), device=dev, generator=g, dtype=torch.int64)
u = torch.rand(a.probe, device=dev, generator=g)
pk = torch.where(pk == HOT, pk + 1, pk)
pk = torch.where(pk == HOT1, pk + 1, pk)
if rank == 0:
    pk = torch.where(u < 0.3, torch.full_like(pk, HOT), pk)
if rank == world - 1:
    pk = torch.where((u >= 0.5) & (u < 0.7), torch.full_like(pk, HOT1), pk)
pk = torch.where((u >= 0.3) & (u < 0.4), torch.full_like(pk, HOT_NOBUILD), pk)

# global key counts (the oracle of this check)
allb = [torch.empty_like(bk) for _ in range(world)]
allp = [torch.empty_like(pk) for _ in range(world)]
dist.all_gather(allb, bk)
dist.all_gather(allp, pk)
cb = torch.bincount(torch.cat(allb), minlength=KS + 100)
cp = torch.bincount(torch.cat(allp), minlength=KS + 100)
want_pairs = int((cb * cp).sum().item())
want_null_build = int(cp[cb == 0].sum().item())
stream = torch.cuda.Stream(device=dev)
for jt in a.joins.split(","):
    j = DistributedJoin(jt, key_width=8, stream=stream, mode="p2p", layout="radix", radix_bits=4)
    j.build_push(0, [bk])
    j.build_finish(0)
    n_b = j.build_rows_local
    bkeys = torch.as_tensor(_RawCudaArray(j.build.recv_keys_ptr, n_b, "<i8"), device=dev).clone()
    n_hot = j.enable_hot_routing(pk, hot_fraction=0.05)
    assert n_hot == 3, f"{jt}: expected the three heavy hitters to be found, got {n_hot}"
    pairs = null_build = 0
    for lo in range(0, a.probe, a.chunk):
        r = j.probe_batch(0, [pk[lo:lo + a.chunk]], want=d.OUT_PAIRS)
        side = j.probe
        for res, bsrc, psrc in ((r, bkeys, torch.as_tensor(_RawCudaArray(side.recv_keys_ptr, int(side.plan.recv_rows[rank]), "<i8"), device=dev)),
                                (r.hot, j.hot_build_keys, side.hot_keys[: side.n_hot])):
            if res is None or res.n == 0:
                continue
            b_, p_ = res.device_arrays()
            bi = torch.as_tensor(b_, device=dev)
            pi = torch.as_tensor(p_, device=dev).to(torch.int64) & 0xffffffff
            isnull = bi == -1
            null_build += int(isnull.sum().item())
            ok = ~isnull
            assert bool((bsrc[bi[ok]] == psrc[pi[ok]]).all()), f"{jt}: a pair joins different keys"
            assert bool((cb[psrc[pi[isnull]]] == 0).all()), f"{jt}: a NULL-build row whose key has build rows"
            pairs += int(ok.sum().item())
        if r.hot is not None:
            r.hot.release()
        r.release()
    fin = j.probe_finish(0, want=d.OUT_PAIRS | d.OUT_HOST)
    fb, fp = fin.numpy()
    fin.release()
    t = torch.tensor([pairs, null_build], device=dev, dtype=torch.int64)
    dist.all_reduce(t)
    tot_pairs, tot_null = int(t[0].item()), int(t[1].item())
    if jt in ("inner", "left", "full"):
        assert tot_pairs == want_pairs, (jt, tot_pairs, want_pairs)
    else:
        assert tot_pairs == 0, (jt, tot_pairs)
    assert tot_null == (want_null_build if jt == "full" else 0), (jt, tot_null, want_null_build)
    # the finaliser: distinct local build rows, each satisfying the condition, and all of them
    fb = fb.astype(np.int64)
    assert len(np.unique(fb)) == len(fb), f"{jt}: a build row twice in the finaliser"
    assert (fp == d.NULL_PROBE).all() if len(fp) else True
    matched_local = (cp[bkeys] > 0).cpu().numpy()
    if jt == "inner":
        assert len(fb) == 0
    elif jt == "left_semi":
        assert matched_local[fb].all(), f"{jt}: an unmatched build row in the finaliser"
        assert len(fb) == int(matched_local.sum()), (jt, len(fb), int(matched_local.sum()))
    else:
        assert not matched_local[fb].any(), f"{jt}: a matched build row in the finaliser"
        assert len(fb) == int((~matched_local).sum()), (jt, len(fb), int((~matched_local).sum()))
    if rank == 0:
        print(f"{jt}: ok — {tot_pairs} pairs, {tot_null} NULL-build rows, finaliser rows on rank 0: {len(fb)} (hot keys routed: {n_hot})", flush=True)
    j.close()
    dist.barrier()
if rank == 0:
    print(f"check_dist_left ok: {world} ranks")
dist.destroy_process_group()
