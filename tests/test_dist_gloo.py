"""The N>1 plumbing (pathtracer_b200/dist.py) on CPU: world_size 2 and 3 over gloo.  The render step is
injected from the oracle (same Philox streams as the CUDA kernels), so what is tested is the partition
and the exchange: tiles -> gather -> bit-identical frame; sample ranges -> reduce -> same sums."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT

W, H, PP, SEED = 70, 50, 3, 4


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import pt_oracle as O
    from pathtracer_b200 import dist as D
    from pathtracer_b200 import host

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = host.HostWorld(3).scene(W, H)
    if mode == "tiles":
        frame = np.zeros((H, W), np.uint32)
        tx, _ = D.tile_grid(W, H)
        for t in D.my_tiles(W, H, rank, world):
            x0, y0 = (t % tx) * 32, (t // tx) * 32
            w, h = min(32, W - x0), min(32, H - y0)
            frame[y0:y0 + h, x0:x0 + w] = O.render(sc, W, H, PP, O.RNG_PHILOX, SEED, window=(x0, y0, w, h))["ldr"]
        packed = torch.from_numpy(D.pack_tiles_host(frame, rank, world).view(np.int32).reshape(-1).copy())
        bufs = D.gather_packed_tiles(packed, W, H)
        if rank == 0:
            out = np.zeros((H, W), np.uint32)
            for r in range(world):
                n = D.slots_per_rank(W, H, world)
                D.unpack_tiles_host(out, bufs[r].numpy().view(np.uint32)[:n * 1024].reshape(n, 32, 32), r, world)
            # the one-kernel scatter of the whole gathered buffer (k_unpack_gathered) must give the same frame
            assert np.array_equal(D.unpack_gathered_host(bufs.numpy().view(np.uint32), W, H, world), out)
            np.save(os.path.join(out_dir, "frame.npy"), out)
    else:
        lo, hi = D.sample_range(PP, rank, world)
        r = O.render(sc, W, H, PP, O.RNG_PHILOX, SEED, s_begin=lo, s_end=hi) if hi > lo else None
        acc = np.zeros((H, W, 4), np.float32)
        if r is not None:
            acc[..., :3] = r["sum"]
            acc[..., 3] = hi - lo
        t = torch.from_numpy(acc)
        share = D.pixel_share(W * H, rank, world)
        if share is not None:
            # reduce-scatter path: my share of the summed pixels, gathered in rank order = the full sum
            mine = torch.empty((share[1] - share[0]) * 4)
            D.reduce_scatter_accum(t.view(-1).clone(), mine, None)
            full = torch.empty(W * H * 4) if rank == 0 else None
            D.gather_shares(mine, full if rank == 0 else mine, None)
            if rank == 0:
                np.save(os.path.join(out_dir, "accum_rs.npy"), full.numpy().reshape(H, W, 4))
        D.reduce_accum(t)
        if rank == 0:
            np.save(os.path.join(out_dir, "accum.npy"), t.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("mode", ["tiles", "samples"])
def test_sharded_frame_over_gloo(tmp_path, world, mode):
    import torch.multiprocessing as mp
    from oracle import pt_oracle as O
    from pathtracer_b200 import host
    port = _free_port()
    mp.spawn(_worker, args=(world, port, mode, str(tmp_path)), nprocs=world, join=True)
    sc = host.HostWorld(3).scene(W, H)
    full = O.render(sc, W, H, PP, O.RNG_PHILOX, SEED)
    if mode == "tiles":
        got = np.load(tmp_path / "frame.npy")
        assert np.array_equal(got, full["ldr"])
    else:
        got = np.load(tmp_path / "accum.npy")
        assert (got[..., 3] == PP * PP).all()
        assert np.allclose(got[..., :3], full["sum"], rtol=1e-5, atol=1e-6)
        if (W * H) % world == 0:
            rs = np.load(tmp_path / "accum_rs.npy")
            assert (rs[..., 3] == PP * PP).all() and np.allclose(rs[..., :3], got[..., :3], rtol=1e-6, atol=1e-7)


def test_partitions_cover_exactly_once():
    from pathtracer_b200 import dist as D
    for (w, h) in [(1, 1), (33, 31), (1920, 1080), (3840, 2160), (70, 50)]:
        tx, ty = D.tile_grid(w, h)
        for world in [1, 2, 3, 4, 8]:
            ids = sorted(t for r in range(world) for t in D.my_tiles(w, h, r, world))
            assert ids == list(range(tx * ty))
            counts = [len(D.my_tiles(w, h, r, world)) for r in range(world)]
            assert max(counts) <= D.slots_per_rank(w, h, world) and all(len(D.tile_slots(w, h, r, world)) == D.slots_per_rank(w, h, world) for r in range(world))
            # the diagonal interleave deals every tile row evenly: shares differ by at most one tile per row
            assert max(counts) - min(counts) <= ty
            for r in range(world):
                for j, t in enumerate(D.tile_slots(w, h, r, world)):
                    if t >= 0:
                        assert D.tile_owner(w, h, world, t % tx, t // tx) == (r, j)
    for pp in [0, 1, 3, 4, 23, 32]:
        for world in [1, 2, 3, 4, 8]:
            rs = [D.sample_range(pp, r, world) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == pp * pp
            assert all(a[1] == b[0] for a, b in zip(rs[:-1], rs[1:]))
            sizes = [b - a for a, b in rs]
            assert max(sizes) - min(sizes) <= 1


def test_tile_map_matches_the_library():
    """The Python mirror of the tile -> rank map (dist.tile_owner / tile_slots) equals the C one (pt_tile_owner,
    pt_tile_slots): the kernels, the host accounting and the gloo tests agree on who owns which tile.  No vertical stripes:
    at 1280 and 3840 pixels and 8 ranks (tile rows of 40 and 120) a rank's tiles are spread over every column class."""
    import ctypes as C
    from pathtracer_b200 import dist as D
    from pathtracer_b200 import render
    L = render.lib()
    for (w, h) in [(1, 1), (33, 31), (300, 170), (1280, 720), (1920, 1080), (3840, 2160)]:
        tx, ty = D.tile_grid(w, h)
        for world in [1, 2, 3, 4, 5, 6, 7, 8]:
            assert L.pt_tile_slots(w, h, world) == D.slots_per_rank(w, h, world)
            for (x, y) in {(0, 0), (tx - 1, ty - 1), (tx // 2, ty // 3), (min(7, tx - 1), min(5, ty - 1))}:
                slot = C.c_uint32()
                r = L.pt_tile_owner(w, h, world, x, y, C.byref(slot))
                assert (r, slot.value) == D.tile_owner(w, h, world, x, y)
    for w, h in [(1280, 720), (3840, 2160)]:
        tx, _ = D.tile_grid(w, h)
        cols = {t % tx % 8 for t in D.my_tiles(w, h, 3, 8)}
        assert cols == set(range(8))


def test_pack_unpack_host_roundtrip():
    from pathtracer_b200 import dist as D
    rng = np.random.default_rng(0)
    img = rng.integers(0, 2 ** 32, size=(50, 70), dtype=np.uint64).astype(np.uint32)
    out = np.zeros_like(img)
    for r in range(3):
        D.unpack_tiles_host(out, D.pack_tiles_host(img, r, 3), r, 3)
    assert np.array_equal(out, img)
