"""Host logic of the multi-GPU layer (b200recon/parallel.py) on the gloo backend, world_size 2 and 3, CPU
tensors: slab partition, halo exchange, variable-length gather, vertex bases, ICP pair assignment.  The same
functions run over NCCL on the GPUs (tests/test_multi_gpu.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b200recon import _lib, parallel


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, R, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = _lib.PAD_LO, _lib.PAD_HI
        g = torch.Generator().manual_seed(7)
        full = torch.rand((R, 5, 5, 2), generator=g)                 # "global volume", plane-major
        fullc = torch.randint(0, 1 << 30, (R, 5, 5), generator=g, dtype=torch.int32)
        zb, ze = parallel.slab_range(R, rank, world)
        n = ze - zb
        slab = torch.zeros((lo + n + hi, 5, 5, 2))
        slabc = torch.zeros((lo + n + hi, 5, 5), dtype=torch.int32)
        slab[lo:lo + n] = full[zb:ze]
        slabc[lo:lo + n] = fullc[zb:ze]
        parallel.exchange_halos([slab, slabc, None], n)
        ok = True
        for p in range(lo + n + hi):
            z = zb - lo + p
            want = full[z] if 0 <= z < R else torch.zeros_like(full[0])
            wantc = fullc[z] if 0 <= z < R else torch.zeros_like(fullc[0])
            ok &= bool(torch.equal(slab[p], want)) and bool(torch.equal(slabc[p], wantc))
        # variable-length gather + bases
        rows = torch.full((rank * 3 + 1, 3), float(rank), dtype=torch.float64)
        counts = parallel.all_gather_counts([rows.shape[0], 2 * rank], torch.device("cpu"))
        ok &= counts.shape == (world, 2) and all(int(counts[r, 0]) == r * 3 + 1 for r in range(world))
        bases = parallel.exclusive_bases(counts[:, 0])
        ok &= int(bases[0]) == 0 and all(int(bases[r]) == sum(3 * i + 1 for i in range(r)) for r in range(world))
        got = parallel.gather_rows(rows, counts[:, 0], dst=0)
        if rank == 0:
            want = torch.cat([torch.full((r * 3 + 1, 3), float(r), dtype=torch.float64) for r in range(world)])
            ok &= bool(torch.equal(got, want))
        else:
            ok &= got is None
        # empty contribution
        e = torch.zeros((0 if rank == 1 else 2, 3), dtype=torch.int32)
        ce = parallel.all_gather_counts([e.shape[0]], torch.device("cpu"))[:, 0]
        ge = parallel.gather_rows(e, ce, dst=0)
        if rank == 0:
            ok &= ge.shape[0] == int(ce.sum())
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,R", [(2, 8), (3, 11)])
def test_halo_and_gather_gloo(world, R):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, R, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(r, True) for r in range(world)]


def test_slab_partition_and_round_robin():
    for R, W in [(1024, 8), (512, 3), (10, 4), (2048, 8)]:
        rs = parallel.slab_ranges(R, W)
        assert rs[0][0] == 0 and rs[-1][1] == R
        assert all(rs[i][1] == rs[i + 1][0] for i in range(W - 1))
        sizes = [b - a for a, b in rs]
        assert max(sizes) - min(sizes) <= 1
    owned = [parallel.round_robin(64, r, 8) for r in range(8)]
    assert sorted(sum(owned, [])) == list(range(64)) and all(len(o) == 8 for o in owned)
    assert owned[3][:3] == [3, 11, 19]


def test_halo_plan_messages_pair_up():
    """Every send has the matching recv on the peer with the same plane count."""
    W, n = 4, 6
    plans = [parallel.halo_plan(r, W, n) for r in range(W)]
    for r in range(W):
        for kind, peer, first, cnt in plans[r]:
            other = "recv" if kind == "send" else "send"
            assert any(k == other and p == r and c == cnt for k, p, _, c in plans[peer])
    with pytest.raises(RuntimeError):
        parallel.halo_plan(1, 2, 1)


def test_world_info_without_process_group():
    assert parallel.world_info() == (0, 1)
    assert np.array_equal(parallel.exclusive_bases([3, 0, 5]), [0, 3, 3])


def test_balanced_slab_ranges():
    from b200recon import parallel
    R = 1024
    w = np.zeros(R)
    w[100:330] = np.linspace(3.0, 1.0, 230)          # all the work in the lower third (table-top orbit)
    for world in (1, 2, 3, 4, 8):
        rg = parallel.balanced_slab_ranges(w, world)
        assert len(rg) == world and rg[0][0] == 0 and rg[-1][1] == R
        assert all(rg[i][1] == rg[i + 1][0] for i in range(world - 1))
        assert all(b - a >= 4 for a, b in rg)
        share = np.array([w[a:b].sum() for a, b in rg]) / w.sum()
        assert share.max() <= 1.0 / world + 0.02, (world, share)
    # no work at all: still a valid partition, roughly equal
    rg = parallel.balanced_slab_ranges(np.zeros(64), 4)
    assert [b - a for a, b in rg] == [16, 16, 16, 16]
    # minimum thickness wins over balance when the work sits in very few planes
    w2 = np.zeros(64)
    w2[10] = 1.0
    rg = parallel.balanced_slab_ranges(w2, 8)
    assert all(b - a >= 4 for a, b in rg) and rg[-1][1] == 64
    with pytest.raises(RuntimeError):
        parallel.balanced_slab_ranges(np.ones(16), 8)


def test_rebalance_slab_ranges_equalises_measured_time():
    from b200recon import parallel
    R = 512
    work = np.zeros(R)
    work[50:200] = 1.0
    true_cost = work.copy()
    true_cost[50:80] *= 3.0            # a region whose voxels cost three times what the update count suggests
    rg = parallel.balanced_slab_ranges(work, 4)
    for _ in range(3):
        times = [true_cost[a:b].sum() + 1e-9 for a, b in rg]
        rg = parallel.rebalance_slab_ranges(rg, work, times)
    times = np.array([true_cost[a:b].sum() for a, b in rg])
    assert times.max() / times.mean() < 1.12, times
    assert rg[0][0] == 0 and rg[-1][1] == R and all(rg[i][1] == rg[i + 1][0] for i in range(3))


def _agree_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        R = 64
        base = np.zeros(R)
        base[10:30] = np.linspace(1.0, 3.0, 20)
        noisy = base * (1.0 + 0.2 * (rank + 1))          # every rank "measures" a different profile
        local = parallel.balanced_slab_ranges(noisy + 1e-3 * rank * np.arange(R), world)
        agreed = parallel.agree_plane_work(noisy)
        ranges = parallel.balanced_slab_ranges(agreed, world)
        ok = True
        parallel.check_same_ranges(ranges, R)                            # passes: same numbers everywhere
        want = base * (1.0 + 0.2 * (sum(range(1, world + 1)) / world))
        ok &= bool(np.allclose(agreed, want))
        # a per-rank list that differs is caught on every rank
        bad = [(0, 20 + rank), (20 + rank, R)] if world == 2 else [(0, 20 + rank), (20 + rank, 40), (40, R)]
        try:
            parallel.check_same_ranges(bad, R)
            ok = False
        except RuntimeError:
            pass
        q.put((rank, bool(ok), local))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_ranks_agree_on_measured_profile_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_agree_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)


def test_check_same_ranges_tiling_without_group():
    parallel.check_same_ranges([(0, 10), (10, 64)], 64)
    for bad in ([(0, 10), (11, 64)], [(1, 10), (10, 64)], [(0, 10), (10, 63)]):
        with pytest.raises(RuntimeError):
            parallel.check_same_ranges(bad, 64)
    assert np.array_equal(parallel.agree_plane_work([1.0, 2.0]), [1.0, 2.0])


def test_mesh_checksum_is_order_and_numbering_independent():
    """parallel.mesh_checksum (the digest bench.py prints for the N-GPU mesh): invariant under vertex order, triangle
    order, vertex renumbering and the start corner of a triangle; sensitive to orientation, to a moved vertex and to
    a dropped triangle; degenerate triangles (repeated corners, which Open3D keeps) included."""
    rng = np.random.default_rng(0)
    V = rng.uniform(-1, 1, (1000, 3))
    T = rng.integers(0, 1000, (3000, 3))
    T[:50, 1] = T[:50, 0]
    a = parallel.mesh_checksum(V, T, 1e-7)
    perm = rng.permutation(1000)
    inv = np.empty(1000, np.int64)
    inv[perm] = np.arange(1000)
    T2 = np.roll(inv[T], 1, axis=1)[rng.permutation(3000)]
    assert parallel.mesh_checksum(V[perm], T2, 1e-7) == a
    T3 = T2.copy()
    T3[100] = T3[100][::-1]
    assert parallel.mesh_checksum(V[perm], T3, 1e-7) != a
    V4 = V.copy()
    V4[7, 1] += 1e-6
    assert parallel.mesh_checksum(V4, T, 1e-7)["vertex_sum"] != a["vertex_sum"]
    assert parallel.mesh_checksum(V, T[:-1], 1e-7)["triangle_sum"] != a["triangle_sum"]
    e = parallel.mesh_checksum(np.zeros((0, 3)), np.zeros((0, 3), np.int64), 1e-7)
    assert e["vertices"] == 0 and e["triangle_sum"] == "0" * 16


def test_local_slabs_rejects_bad_partitions():
    with pytest.raises(RuntimeError):
        parallel.LocalSlabs(1.0, 32, 0.1, x_ranges=[(0, 10), (12, 32)])
    with pytest.raises(RuntimeError):
        parallel.LocalSlabs(1.0, 32, 0.1, x_ranges=[(0, 1), (1, 32)])          # thinner than the halo
