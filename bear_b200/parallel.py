"""Multi-GPU host logic: one process per GPU, `torch.distributed` for the plumbing.

The render path is linear in its inputs (visr_bear/test/test_changes.py:341-361 pins this for the
reference at 1e-5), so a scene shards by OBJECT GROUP: rank g runs the full pipeline on its objects
with the same listener tables.  On GPUs the exchange lives in the library (bear_cuda_comm_connect,
include/bear_cuda.h): each rank mixes its objects onto the 48 + 24 bus rows, then convolves 24 / W of
the loudspeakers on the sum of all ranks' rows, read from the peers over NVLink, and the stereo
partial mixes [2, T*B] are summed with NCCL.  The CPU tests (gloo) sum the partial mixes of oracle
renderers here.  DirectSpeakers and HOA channels, which are few, stay on rank 0.  Batches of independent files shard BY FILE with no exchange.
The reference has no distributed code at all; this module is new design.
"""
import numpy as np

from .control import BlockTables


def object_group(rank, world, num_objects):
    """Contiguous slice of objects rendered by `rank` (sizes differ by at most one)."""
    base, extra = divmod(num_objects, world)
    start = rank * base + min(rank, extra)
    return slice(start, start + base + (1 if rank < extra else 0))


def file_shard(rank, world, num_files):
    """Indices of the files rendered by `rank` in a batch-of-files job (round robin)."""
    return list(range(rank, num_files, world))


def shard_tables(tb, rank, world):
    """Per-rank view of a scene's tables: the rank's object columns; DS / HOA only on rank 0."""
    N = tb["direct_gains"].shape[2]
    sl = object_group(rank, world, N)
    out = BlockTables(tb)
    out["direct_gains"] = np.ascontiguousarray(tb["direct_gains"][:, :, sl])
    out["diffuse_gains"] = np.ascontiguousarray(tb["diffuse_gains"][:, :, sl])
    out["direct_delays"] = np.ascontiguousarray(tb["direct_delays"][:, 2 * sl.start:2 * sl.stop])
    if rank != 0:
        T, L = tb["direct_speakers_gains"].shape[:2]
        out["direct_speakers_gains"] = np.zeros((T, L, 0), np.float32)
        out["direct_speakers_delays"] = np.zeros((T, 0), np.float32)
        out["hoa_gains"] = np.zeros((T, tb["hoa_gains"].shape[1], 0), np.float32)
    return out, sl


def shard_counts(rank, world, N, M, H):
    sl = object_group(rank, world, N)
    return sl.stop - sl.start, (M if rank == 0 else 0), (H if rank == 0 else 0)


def reduce_partial_mix(partial, group=None):
    """Sum the ranks' stereo partial mixes in place; `partial` is a torch tensor (CUDA for NCCL, CPU
    for gloo) or a numpy array (gloo).  Every rank ends up with the full mix."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return partial
    if isinstance(partial, np.ndarray):
        t = torch.from_numpy(partial)
        dist.all_reduce(t, group=group)
        return partial
    dist.all_reduce(partial, group=group)
    return partial


def connect_renderers(renderer, rank, world, root=0, group=None):
    """Make the ranks' CUDA renderers one (bear_cuda_comm_connect): the comm blobs travel through
    torch.distributed (any backend -- they are 256 bytes of host data); from then on the exchange is
    the library's own: bus rows read from the peers' memory over NVLink, stereo partial mixes summed
    on the root (root < 0: all-reduced with NCCL, every rank receives the mix)."""
    import torch.distributed as dist
    blobs = [None] * world
    dist.all_gather_object(blobs, renderer.comm_blob(rank), group=group)
    renderer.comm_connect(rank, world, blobs, root)


class ShardedScene:
    """Render one scene across the ranks of a process group.

    `make_renderer(n_obj, n_ds, n_hoa)` builds the local renderer (bear_b200.renderer.Renderer on a
    GPU); it must offer `render(obj, ds, hoa, tables)` -> [2, T*B]."""

    def __init__(self, make_renderer, N, M, H, rank, world, group=None, root=-1):
        self.rank, self.world, self.group, self.root = rank, world, group, root
        self.sl = object_group(rank, world, N)
        self.counts = shard_counts(rank, world, N, M, H)
        self.renderer = make_renderer(*self.counts)
        # a CUDA renderer exchanges inside the library; anything else (the CPU oracle in the gloo
        # tests) renders its partial mix and the sum is taken here
        self.connected = world > 1 and hasattr(self.renderer, "comm_connect")
        if self.connected:
            connect_renderers(self.renderer, rank, world, root, group)

    def render(self, obj, ds, hoa, tables):
        """The whole scene's render on the root (every rank for root < 0), None elsewhere."""
        tb, sl = shard_tables(tables, self.rank, self.world)
        n, m, h = self.counts
        args = (obj[sl] if n else None, ds if m else None, hoa if h else None, tb)
        if self.connected:
            receives = self.root < 0 or self.root == self.rank
            out = self.renderer.render(*args, output=None if receives else False)
            return out if receives else None
        partial = np.ascontiguousarray(self.renderer.render(*args), dtype=np.float32)
        return reduce_partial_mix(partial, self.group)


def render_file_batch(renderer, files, rank=0, world=1):
    """Batch-of-files render (BASELINE configs[4]): rank `rank` renders files rank, rank+world, ... one
    after the other on ONE renderer instance, resetting its state between files (`Renderer.reset`),
    so the BRIR spectra are transformed once per process instead of once per file.  `files`: sequence
    of (objects, direct_speakers, hoa, tables).  Returns {file index: [2, T*B] render}."""
    out = {}
    for idx in file_shard(rank, world, len(files)):
        obj, ds, hoa, tb = files[idx]
        renderer.reset()
        out[idx] = renderer.render(obj, ds, hoa, tb)
    return out
