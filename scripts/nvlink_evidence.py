"""NVLink byte evidence for the two exchange kernels ("no ncu/nsys evidence of NVLink traffic").

ONE process, two GPUs (ncu is never pointed at a multi-rank command): a handle on cuda:1 probes a batch whose pieces sit on
cuda:0 and cuda:1 -- what mode "pull" does, the partition pass reads the peer's send region over NVLink -- and a handle on
cuda:0 scatters a batch into receive buffers on cuda:0 and cuda:1 -- what mode "p2p" does, the scatter's stores cross NVLink.
Run plain first (prints CUDA-event times and the bytes that HAVE to cross), then under
    ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum,gpu__time_duration.sum --clock-control none \
        -k regex:'part_block_kernel|part_scatter_kernel' --csv --log-file gpurun_out/nvlink.csv python scripts/nvlink_evidence.py
Peer access between the two devices is switched on by torch (its first cross-device copy enables it for cudaMalloc memory)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import datafusion_parallelism_b200 as d
from datafusion_parallelism_b200.exchange import shuffle_count, shuffle_scatter, _dev_column

N_PIECE = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 26          # rows per piece (512 MB of int64 keys)
N_BUILD = 10_000_000


def main():
    assert torch.cuda.device_count() >= 2, "needs two GPUs"
    d0, d1 = torch.device("cuda", 0), torch.device("cuda", 1)
    # torch enables peer access lazily, per direction, on the first copy
    a = torch.ones(8, device=d0)
    b = a.to(d1)
    a.copy_(b)
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    g0 = torch.Generator(device=d0).manual_seed(1)
    g1 = torch.Generator(device=d1).manual_seed(2)
    out = {"rows_per_piece": N_PIECE}

    # ---- pull: GPU 1 probes pieces that live on GPU 0 (remote) and GPU 1 (local) ---------------------------------------
    torch.cuda.set_device(1)
    build = torch.randperm(N_BUILD * 4, device=d1, generator=g1)[:N_BUILD].to(torch.int64)
    remote = torch.randint(0, N_BUILD * 4, (N_PIECE,), dtype=torch.int64, device=d0, generator=g0)
    local = torch.randint(0, N_BUILD * 4, (N_PIECE,), dtype=torch.int64, device=d1, generator=g1)
    s1 = torch.cuda.Stream(device=d1)
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    with d.HashJoin("right", key_widths=(8,), layout="radix", radix_bits=5, device=1, stream=s1.cuda_stream) as j1:
        j1.build_push(0, [build])
        j1.build_finish(0)
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s1)
            res = j1.probe_pieces(0, [(remote.data_ptr(), N_PIECE), (local.data_ptr(), N_PIECE)], want=d.OUT_PAIRS)
            e1.record(s1)
            e1.synchronize()
            assert res.n == 2 * N_PIECE
            res.release()
        out["pull"] = {"ms_probe_pieces": round(e0.elapsed_time(e1), 3), "bytes_that_must_cross": N_PIECE * 8,
                       "expect": "part_block_kernel of the REMOTE piece: nvlrx__bytes ~ rows x 8 B (+ request packets on nvltx)"}
        st = j1.stats()
        out["pull"]["seg_probe_launches"] = st.get("seg_probe_launches")

    # ---- push: GPU 0 scatters a batch into receive buffers on GPU 0 and GPU 1 ------------------------------------------
    torch.cuda.set_device(0)
    s0 = torch.cuda.Stream(device=d0)
    keys = torch.randint(0, 1 << 40, (N_PIECE,), dtype=torch.int64, device=d0, generator=g0)
    torch.cuda.synchronize(0)
    with d.HashJoin("inner", key_widths=(8,), device=0, stream=s0.cuda_stream) as j0:
        counts = shuffle_count(j0, keys, 2)
        rk = [torch.empty(int(counts[r]) + 64, dtype=torch.int64, device=dv) for r, dv in ((0, d0), (1, d1))]
        rr = [torch.empty(int(counts[r]) + 64, dtype=torch.int32, device=dv) for r, dv in ((0, d0), (1, d1))]
        torch.cuda.synchronize(0)
        torch.cuda.synchronize(1)
        for rep in range(3):
            c = shuffle_count(j0, keys, 2)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s0)
            shuffle_scatter(j0, 2, [t.data_ptr() for t in rk] + [0], [t.data_ptr() for t in rr] + [0], [0, 0, 0])
            e1.record(s0)
            e1.synchronize()
        out["push"] = {"ms_scatter": round(e0.elapsed_time(e1), 3), "rows_to_peer": int(counts[1]), "bytes_that_must_cross": int(counts[1]) * 12,
                       "GBps_over_nvlink": round(int(counts[1]) * 12 / e0.elapsed_time(e1) / 1e6, 1),
                       "expect": "part_scatter_kernel<8,1>: nvltx__bytes ~ rows_to_peer x 12 B"}
        # the peer's buffer holds what the local one would: same multiset of keys as a host-side split
        got = torch.sort(rk[1][: int(counts[1])].to(d0))[0]
        out["push"]["peer_rows_checked"] = int(got.shape[0])
    print(json.dumps(out))


if __name__ == "__main__":
    main()
