"""Ad-hoc: device time of the probe phase (partition + probe kernel) with and without B200J_OUT_GATHER_HINT, C2 shape."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import datafusion_parallelism_b200 as d

nb, npr = 100_000_000, 1_000_000_000
g = torch.Generator(device="cuda"); g.manual_seed(3)
bk = torch.randperm(nb, device="cuda", generator=g, dtype=torch.int64)
pk = torch.randint(0, nb, (npr,), device="cuda", generator=g, dtype=torch.int64)
with d.HashJoin("inner", layout="auto", timing=True, build_payload_types=(8,)) as j:
    j.build_push(0, [bk], [torch.arange(nb, device="cuda", dtype=torch.int64)]); j.build_finish(0)
    out = {}
    for name, want in (("plain", d.OUT_PAIRS), ("hint", d.OUT_PAIRS | d.OUT_GATHER_HINT), ("hint_u32", d.OUT_PAIRS | d.OUT_GATHER_HINT | d.OUT_BUILD_U32)):
        ms = []
        for rep in range(3):
            s0 = j.stats()
            r = j.probe_batch(0, [pk], want=want)
            s1 = j.stats()
            r.release()
            ms.append(round(s1["ms_probe"] - s0["ms_probe"], 3))
        out[name] = ms
    print(json.dumps(out))
