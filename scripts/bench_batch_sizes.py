"""Probe throughput through the C ABI against the probe-batch size: one b200j_probe_batch call per batch
of 8 192 / 65 536 / 1 M / 16 M rows, device-resident and host-resident keys, against the coalescing feeder
(b200j_probe_push per batch + one b200j_probe_flush per 4 M staged rows). 10 M-row unique build side, 64 M probe rows."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import datafusion_parallelism_b200 as d

nb, npr = 10_000_000, 1 << 26
g = torch.Generator(device="cuda"); g.manual_seed(5)
bk = torch.randperm(nb, device="cuda", generator=g, dtype=torch.int64)
pk = torch.randint(0, nb, (npr,), device="cuda", generator=g, dtype=torch.int64)
hp = d.PinnedArray(npr, np.int64)
torch.from_numpy(hp.array).copy_(pk)
torch.cuda.synchronize()
with d.HashJoin("inner", key_widths=(8,)) as j:
    j.build_push(0, [bk]); j.build_finish(0)
    for mem, keys in (("device", pk), ("host", hp.array)):
        for bs in (8192, 65536, 1 << 20, 1 << 24):
            n_b = min(npr // bs, 2048 if bs == 8192 else 1 << 30)          # bounded: 8192-row calls are slow
            for feeder in (False, True):
                if feeder and bs > (1 << 20):
                    continue
                want = d.OUT_PAIRS | (d.OUT_HOST if mem == "host" else 0)
                if not feeder:                      # warm-up: the handle's block cache gets its buffers of this size
                    j.probe_batch(0, [keys[:bs]], want=want).release()
                    j.probe_batch(0, [keys[:bs]], want=want).release()
                torch.cuda.synchronize(); t0 = time.perf_counter(); tot = 0
                for i in range(n_b):
                    kb = keys[i * bs:(i + 1) * bs]
                    if feeder:
                        j.probe_push(0, [kb])
                        if j.staged_rows(0) >= (1 << 22) or i == n_b - 1:
                            r = j.probe_flush(0, want=want); tot += r.n
                            gth = j.gather_staged(0, r, want=0); gth.release(); r.release()
                    else:
                        r = j.probe_batch(0, [kb], want=want); tot += r.n; r.release()
                torch.cuda.synchronize(); dt = time.perf_counter() - t0
                assert tot == n_b * bs
                print(json.dumps({"keys": mem, "batch_rows": bs, "batches": n_b, "path": "push + flush per 4 M rows (+ gather of the key columns)" if feeder else "probe_batch per batch",
                                  "us_per_batch": round(dt / n_b * 1e6, 1), "probe_rows_per_s": round(n_b * bs / dt)}), flush=True)
