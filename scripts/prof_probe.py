"""Small fixed workload for ncu: one build + a few probe launches (device-resident)."""
import argparse, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import datafusion_parallelism_b200 as d

ap = argparse.ArgumentParser()
ap.add_argument("--build", type=int, default=100_000_000)
ap.add_argument("--probe", type=int, default=1 << 27)
ap.add_argument("--layout", default="direct")
ap.add_argument("--radix-bits", type=int, default=0)
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
torch.cuda.set_device(0)
g = torch.Generator(device="cuda"); g.manual_seed(3)
bk = torch.randperm(a.build, device="cuda", generator=g, dtype=torch.int64)
pk = torch.randint(0, a.build, (a.probe,), device="cuda", generator=g, dtype=torch.int64)
torch.cuda.synchronize()
with d.HashJoin("inner", layout=a.layout, radix_bits=a.radix_bits, timing=True, expected_build_rows=a.build) as j:
    j.build_push(0, [bk]); j.build_finish(0)
    for _ in range(a.reps):
        r = j.probe_batch(0, [pk], want=d.OUT_PAIRS)
        assert r.n == a.probe
        r.release()
    st = j.stats()
    print("probe ms/launch", st["ms_probe"] / st["probe_launches"], "Gtps", a.probe / (st["ms_probe"] / st["probe_launches"]) / 1e6)
