"""Sharding one frame over the GPUs of a box: one process per GPU, torch.distributed for the plumbing.

The reference's only parallelism is CPU threads over 32x32 tiles handed out by MasterThread
(src/win32_main.cpp:888-960).  Two partitions of the same work, both with exactly one exchange:

  * "tiles"   rank r renders the tiles (tx, ty) with (tx + k*ty) % G == r (a diagonal interleave, k coprime
              with G, so sky tiles and Mario tiles spread evenly whatever the frame width) and resolves them straight into a tile-major pack (pt_resolve_pack_tiles);
              the packs are GATHERED into one contiguous buffer on rank 0 (4 B/pixel in total, no
              reduction) and scattered into the frame by ONE kernel (pt_unpack_gathered).  Counter-based
              RNG keyed by (pixel, sample) => the frame is bit-identical for any G.
  * "samples" rank r renders samples [r*spp/G, (r+1)*spp/G) of every pixel into its float4
              accumulation buffer.  REDUCE-SCATTER(sum, fp32) leaves rank r with the sums of pixels
              [r*n/G, (r+1)*n/G); it resolves them (pt_resolve_range) and the 4 B/pixel results are
              gathered straight into rank 0's frame.  (Where n % G != 0, or on a backend without
              reduce-scatter: one REDUCE to rank 0, which resolves.)
  * "tiles_p2p" the same partition as "tiles", exchanged without a collective on the data path: rank 0 exports
              its BGRA frame as a CUDA IPC handle once, the other ranks map it, and every rank's resolve kernel
              stores its tiles straight into that frame over NVLink (pt_resolve_tiles_to) -- resolve + pack +
              gather + unpack become one kernel per rank.  Two stream-ordered one-word all-reduces per frame
              order the ranks: (A) rank 0 has finished with the previous frame, (B) every rank's tiles have landed.

The collectives run on torch tensors, so the same code runs on NCCL (GPU) and on gloo (CPU tests,
where the render step is injected).  Nothing here does render arithmetic.
"""
import numpy as np

TILE = 32


def tile_grid(width, height):
    return (width + TILE - 1) // TILE, (height + TILE - 1) // TILE


def tile_skew(world):
    """pt_tile_skew (csrc/pt_device.cuh): the smallest k >= world/2 + 1 coprime with world, mod world; 0 for one rank."""
    from math import gcd
    if world <= 1:
        return 0
    k = world // 2 + 1
    while gcd(k, world) != 1:
        k += 1
    return k % world


def slots_per_rank(width, height, world):
    """Tile slots in every rank's pack (pt_tile_slots): tiles_y rows of ceil(tiles_x / world); some may be empty."""
    tx, ty = tile_grid(width, height)
    return ty * ((tx + world - 1) // world)


def tile_owner(width, height, world, tx_, ty_):
    """(rank, slot) of tile (tx_, ty_): rank = (tx + skew * ty) % world, slot = ty * slots_per_row + tx // world."""
    tx, _ = tile_grid(width, height)
    return (tx_ + tile_skew(world) * ty_) % world, ty_ * ((tx + world - 1) // world) + tx_ // world


def tile_slots(width, height, rank, world):
    """Row-major tile id held by each slot of `rank` (mirror of pt_slot_tile), -1 for an empty slot."""
    tx, ty = tile_grid(width, height)
    spr, k = (tx + world - 1) // world, tile_skew(world)
    out = []
    for y in range(ty):
        first = (rank - k * y) % world
        for m in range(spr):
            x = first + world * m
            out.append(y * tx + x if x < tx else -1)
    return out


def my_tiles(width, height, rank, world):
    """Row-major ids of the tiles owned by `rank`, in slot order."""
    return [t for t in tile_slots(width, height, rank, world) if t >= 0]


def max_tiles_per_rank(width, height, world):
    """Length of a rank's pack in tiles (= slots_per_rank: equal for every rank)."""
    return slots_per_rank(width, height, world)


def sample_range(pp, rank, world):
    """Contiguous slice of the pp*pp sample list owned by `rank` (sizes differ by at most one)."""
    spp = pp * pp
    base, extra = divmod(spp, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def pack_tiles_host(bgra, rank, world):
    """numpy mirror of k_pack_tiles (csrc/pt_kernels.cuh): [n_slots, 32, 32] uint32, zero padded (empty slots: zeros)."""
    h, w = bgra.shape
    tx, _ = tile_grid(w, h)
    ids = tile_slots(w, h, rank, world)
    out = np.zeros((len(ids), TILE, TILE), np.uint32)
    for j, t in enumerate(ids):
        if t < 0:
            continue
        x0, y0 = (t % tx) * TILE, (t // tx) * TILE
        blk = bgra[y0:y0 + TILE, x0:x0 + TILE]
        out[j, :blk.shape[0], :blk.shape[1]] = blk
    return out


def unpack_tiles_host(bgra, packed, rank, world):
    """numpy mirror of k_unpack_tiles."""
    h, w = bgra.shape
    tx, _ = tile_grid(w, h)
    for j, t in enumerate(tile_slots(w, h, rank, world)):
        if t < 0:
            continue
        x0, y0 = (t % tx) * TILE, (t // tx) * TILE
        hh, ww = min(TILE, h - y0), min(TILE, w - x0)
        bgra[y0:y0 + hh, x0:x0 + ww] = packed[j, :hh, :ww]
    return bgra


def unpack_gathered_host(gathered, width, height, world):
    """numpy mirror of k_unpack_gathered: `gathered` is [world, n_max*1024] uint32, rank r's pack in row r."""
    tx, _ = tile_grid(width, height)
    y, x = np.mgrid[0:height, 0:width]
    spr = (tx + world - 1) // world
    rank = (x // TILE + tile_skew(world) * (y // TILE)) % world
    slot = (y // TILE) * spr + (x // TILE) // world
    return gathered[rank, slot * (TILE * TILE) + (y % TILE) * TILE + (x % TILE)]


def gather_packed_tiles(packed, width, height, group=None, out=None):
    """Gather every rank's packed tiles ([n_r*1024] int32 tensor) into ONE contiguous [world, n_max*1024] tensor on
    rank 0 (`out` if given); returns it there, None elsewhere."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n_max = max_tiles_per_rank(width, height, world) * TILE * TILE
    if packed.numel() < n_max:
        packed = torch.cat([packed, packed.new_zeros(n_max - packed.numel())])
    if world == 1:
        return packed.view(1, n_max)
    if rank == 0 and out is None:
        out = packed.new_empty((world, n_max))
    dist.gather(packed, list(out.unbind(0)) if rank == 0 else None, dst=0, group=group)
    return out if rank == 0 else None


def reduce_accum(accum, group=None):
    """Sum the ranks' float accumulation buffers into rank 0's (in place)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM, group=group)
    return accum


def pixel_share(n_pixels, rank, world):
    """[lo, hi) of the pixel list owned by `rank` after the reduce-scatter; None when it does not divide evenly."""
    if world < 1 or n_pixels % world:
        return None
    c = n_pixels // world
    return rank * c, (rank + 1) * c


def reduce_scatter_accum(accum_flat, out, group=None):
    """out <- this rank's share of the element-wise sum of every rank's accum_flat (len(accum_flat) = world*len(out)).
    NCCL: one reduce-scatter.  gloo has none: all-reduce a copy and slice (CPU tests of the partition)."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    if dist.get_backend(group) == "nccl":
        dist.reduce_scatter_tensor(out, accum_flat, op=dist.ReduceOp.SUM, group=group)
    else:
        t = accum_flat.clone()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        out.copy_(t.view(world, -1)[rank])
    return out


def gather_shares(mine, frame_flat, group=None):
    """Gather every rank's resolved share (`mine`, equal lengths) into rank 0's flat frame, in rank order."""
    import torch.distributed as dist
    rank = dist.get_rank(group)
    dist.gather(mine, list(frame_flat.split(mine.numel())) if rank == 0 else None, dst=0, group=group)


class ShardedFrame:
    """Render one frame on this rank's GPU and exchange it; `core` is a render.RenderCore whose
    stream is torch's current stream (so the collectives order after the kernels)."""

    def __init__(self, core, width, height, mode="tiles", group=None):
        import torch
        import torch.distributed as dist
        self.core, self.w, self.h, self.mode, self.group = core, width, height, mode, group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.torch = torch
        self._packed = self._gathered = self._share = None
        self.launches = 0
        self._remote = None          # tiles_p2p: rank 0's frame mapped into this process
        self._flag = None
        self.phase_events = None     # a list: render() appends a CUDA event recorded right after the render launch (bench.py)

    def _mark_rendered(self):
        if self.phase_events is not None:
            ev = self.torch.cuda.Event(enable_timing=True)
            ev.record()
            self.phase_events.append(ev)

    def _p2p_setup(self):
        """Once per frame size: rank 0 exports its BGRA frame, everybody else maps it."""
        import torch.distributed as dist
        torch, core = self.torch, self.core
        self._flag = torch.zeros(1, dtype=torch.int32, device="cuda")
        h = torch.zeros(80, dtype=torch.uint8, device="cuda")      # PT_IPC_HANDLE_BYTES
        if self.rank == 0:
            raw = core.frame_ipc_handle(self.w, self.h)
            h.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
        if self.world > 1:
            dist.broadcast(h, src=0, group=self.group)
            if self.rank != 0:
                self._remote = core.open_peer_frame(bytes(h.cpu().numpy().tobytes()))
        self._p2p_ready = True

    def close(self):
        if self._remote:
            self.core.close_peer_frame(self._remote)
            self._remote = None

    def render(self, pp, seed):
        """Returns rank 0's resolved BGRA frame as a device int32 tensor [H, W] (None on other ranks)."""
        torch, core = self.torch, self.core
        self.launches = 0
        if self.mode == "tiles_p2p":
            import torch.distributed as dist
            if not getattr(self, "_p2p_ready", False):
                self._p2p_setup()
            core.render(self.w, self.h, pp, seed=seed, tile_rank=self.rank, tile_count=self.world)
            self._mark_rendered()
            if self.world > 1:
                dist.all_reduce(self._flag, group=self.group)      # (A) rank 0's stream is past everything that read the previous frame
            core.resolve_tiles_to(self._remote if self.rank != 0 else None, self.rank, self.world)
            self.launches += 2
            if self.world > 1:
                dist.all_reduce(self._flag, group=self.group)      # (B) ordered after every rank's stores on its stream
            if self.rank != 0:
                return None
            _, b = core.device_buffers()
            return torch.as_tensor(b, device="cuda")
        if self.mode == "tiles":
            core.render(self.w, self.h, pp, seed=seed, tile_rank=self.rank, tile_count=self.world)
            self._mark_rendered()
            if self.world == 1:
                core.resolve()
                self.launches += 2
                _, b = core.device_buffers()
                return torch.as_tensor(b, device="cuda")
            n_max = max_tiles_per_rank(self.w, self.h, self.world) * TILE * TILE
            if self._packed is None:
                self._packed = torch.zeros(n_max, dtype=torch.int32, device="cuda")
                if self.rank == 0:
                    self._gathered = torch.zeros((self.world, n_max), dtype=torch.int32, device="cuda")
            core.resolve_pack_tiles(self._packed.data_ptr(), self.rank, self.world)     # resolve of my tiles only, tile-major
            self.launches += 2
            gather_packed_tiles(self._packed, self.w, self.h, self.group, out=self._gathered)
            if self.rank != 0:
                return None
            core.unpack_gathered(self._gathered.data_ptr(), self.world, n_max)           # every rank's tiles in one kernel
            self.launches += 1
            _, b = core.device_buffers()
            return torch.as_tensor(b, device="cuda")
        lo, hi = sample_range(pp, self.rank, self.world)
        if hi > lo:
            core.render(self.w, self.h, pp, seed=seed, sample_begin=lo, sample_end=hi)
        else:
            core.render(self.w, self.h, 0, seed=seed)      # more ranks than samples: contribute zeros
        self._mark_rendered()
        self.launches += 1
        a, b = core.device_buffers()
        acc = torch.as_tensor(a, device="cuda")
        frame = torch.as_tensor(b, device="cuda")
        share = pixel_share(self.w * self.h, self.rank, self.world)
        if self.world > 1 and share is not None:
            n = share[1] - share[0]
            if self._share is None:
                self._share = torch.empty(n * 4, dtype=torch.float32, device="cuda")
            reduce_scatter_accum(acc.view(-1), self._share, self.group)
            mine = frame.view(-1)[share[0]:share[1]]
            core.resolve_range(self._share.data_ptr(), mine.data_ptr(), n)               # my share of the pixels
            self.launches += 1
            gather_shares(mine, frame.view(-1), self.group)
            if self.rank != 0:
                return None
            core.mark_resolved()
            return frame
        reduce_accum(acc, self.group)
        if self.rank != 0:
            return None
        core.resolve()
        self.launches += 1
        return frame
