"""Times b200j_shuffle_local_scatter (single-pass scatter by destination into local regions) on one GPU."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import datafusion_parallelism_b200 as d
from datafusion_parallelism_b200.exchange import shuffle_local_scatter
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000_000
g = torch.Generator(device="cuda"); g.manual_seed(1)
keys = torch.randint(0, 1 << 40, (n,), device="cuda", generator=g, dtype=torch.int64)
for W in (2, 4, 8, 16):
    cap = (int(n / W * 1.15) + 65536 + 63) // 64 * 64
    sk = torch.empty((W + 2) * cap, dtype=torch.int64, device="cuda")
    sr = torch.empty((W + 2) * cap, dtype=torch.int32, device="cuda")
    with d.HashJoin("inner", key_widths=(8,), timing=True) as j:
        for rep in range(3):
            j.reset_stats()
            c = shuffle_local_scatter(j, keys, W, sk.data_ptr(), sr.data_ptr(), cap)
            ms = j.stats()["ms_exchange_partition"]
        print(f"n_dest {W}: {ms:.3f} ms per {n} rows  ({n * 20 / ms / 1e6:.0f} GB/s of 20 B/row)  counts max/min {int(c[:W].max())}/{int(c[:W].min())}")
    del sk, sr
