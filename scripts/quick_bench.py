"""Ad-hoc device-resident timing of build + probe (not the contract bench: see bench.py)."""
import argparse, json, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import datafusion_parallelism_b200 as d

ap = argparse.ArgumentParser()
ap.add_argument("--build", type=int, default=100_000_000)
ap.add_argument("--probe", type=int, default=1_000_000_000)
ap.add_argument("--chunk", type=int, default=1 << 28)
ap.add_argument("--layout", default="direct")
ap.add_argument("--radix-bits", type=int, default=0)
ap.add_argument("--lf", type=float, default=0.5)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--miss-mult", type=int, default=1, help="probe keys uniform over build x this many ids: hit rate = 1 / miss_mult")
a = ap.parse_args()

torch.cuda.set_device(0)
g = torch.Generator(device="cuda"); g.manual_seed(3)
bk = torch.randperm(a.build, device="cuda", generator=g, dtype=torch.int64)
pk = torch.randint(0, a.build * a.miss_mult, (a.probe,), device="cuda", generator=g, dtype=torch.int64)
torch.cuda.synchronize()
for rep in range(a.reps):
    with d.HashJoin("inner", layout=a.layout, radix_bits=a.radix_bits, load_factor=a.lf, timing=True, expected_build_rows=a.build) as j:
        t0 = time.time()
        j.build_push(0, [bk]); j.build_finish(0)
        st = j.stats()
        tb = time.time() - t0
        tot = 0
        t0 = time.time()
        for lo in range(0, a.probe, a.chunk):
            r = j.probe_batch(0, [pk[lo:lo + a.chunk]], want=d.OUT_PAIRS)
            tot += r.n
            r.release()
        tp = time.time() - t0
        st2 = j.stats()
    bms = st["ms_build_copy"] + st["ms_build_partition"] + st["ms_build_insert"]
    pms = st2["ms_probe"] + st2["ms_probe_partition"]
    print(json.dumps({"rep": rep, "layout": st["layout"], "radix_bits": st["radix_bits"], "table_MB": st["table_bytes"] / 1e6,
                      "build_ms": bms, "build_copy_ms": st["ms_build_copy"], "build_part_ms": st["ms_build_partition"], "build_insert_ms": st["ms_build_insert"],
                      "build_Gtps": a.build / bms / 1e6, "build_wall_ms": tb * 1e3,
                      "probe_ms": pms, "probe_part_ms": st2["ms_probe_partition"], "probe_kernel_ms": st2["ms_probe"],
                      "probe_Gtps": a.probe / pms / 1e6, "probe_wall_ms": tp * 1e3, "probe_roofline_frac_36B": a.probe * 36 / (pms * 1e-3) / 6537.3e9,
                      "matches": tot, "launches": st2["kernel_launches"]}))
