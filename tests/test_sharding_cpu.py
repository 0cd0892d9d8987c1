"""CPU tests (no GPU) of the multi-GPU host logic with world_size=2 over gloo: keyframe-range shards of the
descriptor database + all-gather + merge == one scan; frame sharding covers every frame once."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ar_orbslam2_b200.sharding import frames_of_rank, shard_range, sharded_top2
from ar_orbslam2_b200.synth import synth_descriptors, synth_queries


def test_shard_ranges_partition():
    for n in (0, 1, 7, 256, 50000):
        for world in (1, 2, 3, 8):
            r = [shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n and all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(e - b for b, e in r) - min(e - b for b, e in r) <= 1
    assert sorted(i for k in range(4) for i in frames_of_rank(10, k, 4)) == list(range(10))


def _merge_numpy(bd, bi, sd, si):
    """Reference merge in shard order (what orb_top2_merge does on the device)."""
    world, nq = bd.shape
    out = [np.full(nq, 256, np.int32), np.full(nq, -1, np.int64), np.full(nq, 256, np.int32), np.full(nq, -1, np.int64)]
    for q in range(nq):
        d1, i1, d2, i2 = 256, -1, 256, -1
        for p in range(world):
            for d, i in ((int(bd[p, q]), int(bi[p, q])), (int(sd[p, q]), int(si[p, q]))):
                if i < 0:
                    continue
                if d < d1:
                    d2, i2, d1, i1 = d1, i1, d, i
                elif d < d2:
                    d2, i2 = d, i
        out[0][q], out[1][q], out[2][q], out[3][q] = d1, i1, d2, i2
    return out


def _worker(rank, world, port, n, nq, seed, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import orbref
    b, e = shard_range(n, rank, world)
    rows = synth_descriptors(e - b, seed, first_row=b)           # each rank generates only its shard
    local = orbref.hamming_top2(q, rows, idx_base=b, fast=True)
    local_t = tuple(torch.from_numpy(np.ascontiguousarray(x)) for x in local)
    merged = sharded_top2(local_t, lambda bd, bi, sd, si: _merge_numpy(bd.numpy(), bi.numpy(), sd.numpy(), si.numpy()))
    want = orbref.hamming_top2(q, synth_descriptors(n, seed), fast=True)
    ok = all(np.array_equal(m, w) for m, w in zip(merged, want))
    t = torch.tensor([1 if ok else 0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    assert t.item() == 1


def test_sharded_db_world2_gloo():
    n, nq, seed = 4001, 64, 12
    full = synth_descriptors(n, seed)
    q = synth_queries(lambda idx: full[idx], n, nq, 4)
    q[5] = full[10]; full_dup = full.copy()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(2, port, n, nq, seed, q), nprocs=2, join=True)


def _kf_case(n_kf, kf_rows, nq, seed):
    full = synth_descriptors(n_kf * kf_rows, seed)
    q = synth_queries(lambda idx: full[idx], n_kf * kf_rows, nq, seed + 1)
    return full, q


def test_keyframe_counts_oracle_against_numpy():
    """orbref_db_keyframe_counts (ORBmatcher.cc:216-231's acceptance per (query, keyframe)) against a numpy restatement."""
    from oracle import orbref
    n_kf, kf_rows, nq = 23, 37, 90
    full, q = _kf_case(n_kf, kf_rows, nq, 31)
    full[5 * kf_rows + 3] = full[5 * kf_rows + 2]          # duplicate rows inside a keyframe: best == second -> rejected
    q[1] = full[5 * kf_rows + 2]
    got = orbref.db_keyframe_counts(q, full, kf_rows, 50, 0.75, threads=3)
    bits = np.unpackbits(full, axis=1).astype(np.int32)
    qb = np.unpackbits(q, axis=1).astype(np.int32)
    dist = (qb[:, None, :] != bits[None, :, :]).sum(-1).reshape(nq, n_kf, kf_rows)
    srt = np.sort(dist, axis=2)
    ok = (srt[:, :, 0] <= 50) & (srt[:, :, 0].astype(np.float32) < np.float32(0.75) * srt[:, :, 1].astype(np.float32))
    assert np.array_equal(got, ok.sum(0).astype(np.int32))
    assert got.sum() >= nq // 2 - 2, "planted queries are counted"
    assert orbref.db_keyframe_counts(q, full[:kf_rows], kf_rows)[0] == got[0]


def _kf_worker(rank, world, port, n_kf, kf_rows, nq, seed, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ar_orbslam2_b200.sharding import allgather_keyframe_counts, shard_keyframes
    from oracle import orbref
    b, e = shard_keyframes(n_kf, kf_rows, rank, world)
    rows = synth_descriptors(e - b, seed, first_row=b)
    local = torch.from_numpy(orbref.db_keyframe_counts(q, rows, kf_rows))
    allc = allgather_keyframe_counts(local, n_kf).numpy()
    want = orbref.db_keyframe_counts(q, synth_descriptors(n_kf * kf_rows, seed), kf_rows)
    t = torch.tensor([1 if np.array_equal(allc, want) else 0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    assert t.item() == 1


def test_keyframe_counts_world2_gloo():
    """Keyframe-range shards of unequal size (7 keyframes over 2 ranks): concatenated counts == one pass over the database."""
    n_kf, kf_rows, nq, seed = 7, 50, 40, 77
    full, q = _kf_case(n_kf, kf_rows, nq, seed)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_kf_worker, args=(2, port, n_kf, kf_rows, nq, seed, q), nprocs=2, join=True)
