#!/usr/bin/env python3
"""Multi-GPU correctness check of the sharded all-vs-all path (run under torchrun, one rank per GPU):
orbx_allpairs_sharded (NCCL all-gather inside liborbx.so) and the torch.distributed variant of orb_slam_cuda_b200/sharding.py must
both give, on every rank, exactly the rows a single GPU computes over the full database — for block sizes that do not
divide evenly, several chunk counts and a rank with an empty block.  Prints one JSON line on rank 0; exit code 1 on mismatch.
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/multi_gpu_check.py"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import orb_slam_cuda_b200 as X  # noqa: E402
from orb_slam_cuda_b200 import sharding  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    m = X.ORBmatcher(0.6, True)
    comm = sharding.NcclComm(torch.device("cuda", local))
    # a real (non-default) stream for the torch-side work: a stream argument of 0 means "the library's own stream" in the C ABI, which
    # is not ordered with torch's default stream
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)
    results = []
    ok_all = True
    for k_total, n, chunks in [(37, 300, 4), (world * 6, 500, 3), (world - 1 if world > 1 else 1, 200, 2), (64, 2000, 1), (5 * world + 3, 257, 8)]:
        rng = np.random.default_rng(1000 + k_total)
        base = rng.integers(0, 256, size=(8, n, 32), dtype=np.uint8)
        full_np = np.stack([np.roll(base[i % 8], i // 8, axis=0) ^ (np.random.default_rng(i).random((n, 32)) < 0.03).astype(np.uint8)
                            for i in range(k_total)])
        b, e = sharding.block_range(k_total, rank, world)
        full = torch.from_numpy(full_np).cuda()
        local_desc = full[b:e].contiguous()
        want = torch.zeros((e - b, k_total), dtype=torch.int32, device="cuda")
        if e > b:
            m.match_allpairs_device(local_desc.data_ptr(), e - b, full.data_ptr(), k_total, n, want.data_ptr())
        torch.cuda.synchronize()
        got, moved = sharding.allpairs_sharded_native(comm, local_desc, k_total, 50, 0.6, chunks=chunks)
        got2, _ = sharding.allpairs_sharded_native(comm, local_desc, k_total, 50, 0.6, chunks=chunks)  # workspace reuse
        torch.cuda.synchronize()

        def score_fn(qq, db, o):
            m.match_allpairs_device(qq.data_ptr(), qq.shape[0], db.data_ptr(), db.shape[0], n, o.data_ptr(),
                                    stream=torch.cuda.current_stream().cuda_stream)

        got3, _ = sharding.allpairs_sharded(local_desc, k_total, score_fn, chunks=chunks)
        torch.cuda.synchronize()
        ok = bool(torch.equal(got, want) and torch.equal(got2, want) and torch.equal(got3, want))
        nz = int((want > 0).sum())
        t = torch.tensor([1 if ok else 0, nz], dtype=torch.int64, device="cuda")
        dist.all_reduce(t[:1], op=dist.ReduceOp.MIN)
        dist.all_reduce(t[1:], op=dist.ReduceOp.SUM)
        ok_all = ok_all and int(t[0]) == 1
        results.append({"k_total": k_total, "n": n, "chunks": chunks, "ok": int(t[0]) == 1, "nonzero_scores": int(t[1]), "gathered_bytes": moved})
    comm.destroy()
    if rank == 0:
        print(json.dumps({"world": world, "ok": ok_all, "cases": results}))
    dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()
