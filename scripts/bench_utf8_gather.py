"""Ad-hoc: device time of b200j_gather with one Utf8 build payload column (short strings), by lanes per row of the copy kernel.
usage: B200J_UTF8_LANES=<0|4|8|16|32> python scripts/bench_utf8_gather.py [avg_len]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pyarrow as pa, torch
import datafusion_parallelism_b200 as d

avg = int(sys.argv[1]) if len(sys.argv) > 1 else 12
nb, npr = 4_000_000, 100_000_000
rng = np.random.default_rng(1)
lens = rng.integers(max(1, avg // 2), avg * 3 // 2 + 1, nb)
offs = np.zeros(nb + 1, dtype=np.int32); np.cumsum(lens, out=offs[1:])
data = rng.integers(97, 123, int(offs[-1]), dtype=np.uint8)
strs = pa.Array.from_buffers(pa.string(), nb, [None, pa.py_buffer(offs.tobytes()), pa.py_buffer(data.tobytes())])
bk = rng.permutation(nb).astype(np.int64)
pk = torch.randint(0, nb, (npr,), device="cuda", dtype=torch.int64)
with d.HashJoin("inner", key_widths=(8,), build_payload_types=(d.T_UTF8,), timing=True) as j:
    j.build_push(0, [bk], [strs]); j.build_finish(0)
    res = j.probe_batch(0, [pk], want=d.OUT_PAIRS)
    ms = []
    for rep in range(3):
        s0 = j.stats()["ms_gather"]
        g = j.gather(res, [pk], want=0)
        ms.append(round(j.stats()["ms_gather"] - s0, 3))
        vb = g.column_buffers(1)["values_bytes"]
        g.release()
    res.release()
print(json.dumps({"lanes": os.environ.get("B200J_UTF8_LANES", "auto"), "avg_len": avg, "rows": npr, "utf8_bytes": vb, "ms_gather": ms}))
