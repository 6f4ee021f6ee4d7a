"""Host-side logic of the row-sharded search, exercised with world_size 2 and 3 on CPU (gloo).

Each rank plays one GPU: it owns the striped rows ``plan_rows`` assigns to it, produces its local
top-k as packed 64-bit keys (here with the numpy oracle standing in for the kernel — this test is
about routing, packing and merging, not arithmetic), all-gathers the keys with gloo and merges them
with ``merge_keys_host`` (the host restatement of kernel K4).  The result must equal the oracle's
search on the whole database for any world size.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _ord(s):
    b = np.asarray(s, dtype=np.float32).view(np.uint32)
    return np.where(b & np.uint32(0x80000000), ~b, b | np.uint32(0x80000000)).astype(np.uint64)


def _worker(rank, world, port, n, d, B, k, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import synth
        from oracle import nns_oracle
        from cslam_b200.sharded import plan_rows, merge_keys_host

        db = synth.unit_rows(77, n, d, "rand")
        qs = synth.queries_for(78, db, B, "rand")
        # rows arrive in two blocks, like incremental adds: plan_rows routes each block
        mine = []
        for first, count in ((0, n // 3), (n // 3, n - n // 3)):
            pos, local = plan_rows(first, count, rank, world)
            assert (local == (first + pos) // world).all()
            mine.append(first + pos)
        mine = np.concatenate(mine)
        assert (mine % world == rank).all()
        shard = db[mine]
        keys = np.zeros((B, k), dtype=np.uint64)
        if len(shard):
            # similarities are computed on the whole matrix and sliced, so that BLAS blocking cannot
            # make a shard's fp32 rounding differ from the whole-DB oracle's: routing is under test
            kk = min(k, len(shard))
            for b in range(B):
                s = nns_oracle.cosine_similarities(db, qs[b])[mine]
                order = np.argsort(-s, kind="stable")[:kk]
                grow = mine[order].astype(np.uint64)
                keys[b, :kk] = (_ord(s[order]) << np.uint64(32)) | (~grow & np.uint64(0xFFFFFFFF))
        t = torch.from_numpy(keys.view(np.int64))
        gathered = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)
        allk = np.stack([g.numpy().view(np.uint64) for g in gathered])
        rows, sims = merge_keys_host(allk, k)
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), rows=rows, sims=sims)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,k", [(2, 501, 10), (3, 100, 7), (2, 5, 10)])
def test_sharded_merge_equals_whole_db(tmp_path, world, n, k):
    import synth
    from oracle import nns_oracle
    d, B = 24, 6
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, d, B, k, str(tmp_path)), nprocs=world, join=True)
    db = synth.unit_rows(77, n, d, "rand")
    qs = synth.queries_for(78, db, B, "rand")
    ref_idx, ref_sims = nns_oracle.search_batch(db, qs, k)
    kk = ref_idx.shape[1]
    for r in range(world):
        z = np.load(os.path.join(str(tmp_path), "rank%d.npz" % r))
        assert (z["rows"][:, :kk] == ref_idx).all(), "rank %d" % r
        np.testing.assert_array_equal(z["sims"][:, :kk], ref_sims)
        assert (z["rows"][:, kk:] == -1).all()


def test_plan_rows_partitions_every_row_once():
    from cslam_b200.sharded import plan_rows
    for world in (1, 2, 4, 8):
        seen = []
        for rank in range(world):
            pos, local = plan_rows(13, 100, rank, world)
            seen.extend((13 + pos).tolist())
            assert ((13 + pos) // world == local).all()
        assert sorted(seen) == list(range(13, 113))
