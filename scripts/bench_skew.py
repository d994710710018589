"""C3 (BASELINE.json configs[2]): skewed / duplicate-heavy keys on one GPU, device-resident, timed with the library's own
CUDA events. Not the contract bench (bench.py measures C2); the lines this prints go to profiles/ and DESIGN.md.

Workloads:
  exp     build = make_exponential_int_array (src/api_utils.rs:15-23) per 1024-row block, exactly as
          benches/exponential_distribution.rs:178-188 builds its join tables, scaled to --build rows (Int32 keys, up to ~6
          copies of a key at the low end of every block, gaps at the high end); probe = uniform over the same id range.
  zipf-b  build = Zipf(theta = 1) over the id range (real heavy hitters: duplicate runs of millions of rows), probe uniform.
  zipf-p  build = unique keys, probe = Zipf(theta = 1) (hot probe keys).
Every run checks the output cardinality against sum_k count_build(k) * count_probe(k) computed with torch.
"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import datafusion_parallelism_b200 as d
import oracle

ap = argparse.ArgumentParser()
ap.add_argument("--build", type=int, default=100_000_000)
ap.add_argument("--probe", type=int, default=1_000_000_000)
ap.add_argument("--chunk", type=int, default=0, help="probe rows per call; 0 = whole probe side (zipf-b: 2^28, its pair buffer is sized for 5.5 pairs per probe row)")
ap.add_argument("--workloads", default="exp,zipf-b,zipf-p")
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--join", default="inner")
a = ap.parse_args()
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
HBM = 6531.3e9
try:
    HBM = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] * 1e9
except Exception:
    pass


def zipf_keys(n, key_space, seed):
    """Zipf(theta = 1) over [0, key_space) by inverse CDF: P(k) ~ 1 / (k + 1). cdf(k) = H(k + 1) / H(N) ~ ln(k + 1.5) / ln(N + 0.5)."""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    u = torch.rand(n, device=dev, generator=g, dtype=torch.float64)
    k = torch.exp(u * np.log(key_space + 0.5)) - 0.5
    k = k.clamp_(0, key_space - 1).to(torch.int32)
    # scatter the ranks over the id range so that hot keys are not the small integers (multiplicative permutation)
    mult = 2654435761 % key_space
    while np.gcd(mult, key_space) != 1:
        mult += 1
    return ((k.to(torch.int64) * mult) % key_space).to(torch.int32)


def make(workload, nb, npr):
    g = torch.Generator(device=dev)
    g.manual_seed(5)
    if workload == "exp":
        pat = torch.from_numpy(np.asarray(oracle.make_exponential_int_array(0, 1024), dtype=np.int32)).to(dev)
        i = torch.arange(nb, device=dev, dtype=torch.int64)
        bk = ((i // 1024) * 1024 + pat[i % 1024].to(torch.int64)).to(torch.int32)
        del i
        pk = torch.randint(0, nb, (npr,), device=dev, generator=g, dtype=torch.int32)
    elif workload == "zipf-b":
        bk = zipf_keys(nb, nb, 5)
        pk = torch.randint(0, nb, (npr,), device=dev, generator=g, dtype=torch.int32)
    elif workload == "zipf-p":
        bk = torch.randperm(nb, device=dev, generator=g, dtype=torch.int32)
        pk = zipf_keys(npr, nb, 6)
    else:
        raise SystemExit(f"unknown workload {workload}")
    return bk, pk


def expected_rows(bk, pk, key_space):
    cnt = torch.bincount(bk.to(torch.int64), minlength=key_space)
    tot = 0
    for lo in range(0, pk.numel(), 1 << 27):
        tot += int(cnt[pk[lo:lo + (1 << 27)].to(torch.int64)].sum().item())
    top = int(cnt.max().item())
    del cnt
    return tot, top


for wl in a.workloads.split(","):
    bk, pk = make(wl, a.build, a.probe)
    want, top = expected_rows(bk, pk, a.build)
    chunk = a.chunk if a.chunk > 0 else ((1 << 28) if wl == "zipf-b" else a.probe)
    torch.cuda.synchronize()
    best = None
    for rep in range(a.reps + 1):
        with d.HashJoin(a.join, key_widths=(4,), layout="auto", timing=True, expected_build_rows=a.build) as j:
            t0 = time.perf_counter()
            j.build_push(0, [bk])
            j.build_finish(0)
            tot = 0
            for lo in range(0, a.probe, chunk):
                r = j.probe_batch(0, [pk[lo:lo + chunk]], want=d.OUT_PAIRS)
                tot += r.n
                r.release()
            wall = (time.perf_counter() - t0) * 1e3
            st = j.stats()
        assert tot == want, f"{wl}: {tot} result rows, expected {want}"
        bms = st["ms_build_copy"] + st["ms_build_partition"] + st["ms_build_insert"]
        pms = st["ms_probe_partition"] + st["ms_probe"]
        line = {"workload": wl, "build_rows": a.build, "probe_rows": a.probe, "probe_chunk_rows": chunk, "key": "int32", "join": a.join, "result_rows": tot,
                "longest_duplicate_run": top, "distinct_keys": st["distinct_keys"], "duplicate_rows": st["duplicate_rows"],
                "layout": st["layout"], "radix_bits": st["radix_bits"], "build_ms": bms, "probe_ms": pms, "wall_ms": wall,
                "build_Gtps": a.build / bms / 1e6, "probe_Gtps": a.probe / pms / 1e6,
                "probe_frac_hbm_36B": a.probe * 36 / (pms * 1e-3) / HBM, "ms": {k: st[k] for k in st if k.startswith("ms_")},
                "dense_probe_launches": st.get("dense_probe_launches"), "probe_retries": st["probe_retries"]}
        if rep > 0 and (best is None or line["build_ms"] + line["probe_ms"] < best["build_ms"] + best["probe_ms"]):
            best = line
    print(json.dumps(best), flush=True)
    del bk, pk
    torch.cuda.empty_cache()
