"""Multi-GPU plumbing: one process per GPU (torch.distributed / NCCL), structured box cut into
x-slabs.  The reference is single-device (main.py:40); this is new.

Each element lives on exactly one rank; the vertex plane between two neighbouring slabs exists
on both.  After the tile kernel every rank packs the per-vertex sum of its local contributions on
the shared planes, exchanges it with its (at most two) neighbours, and finishes the shared vertices
redundantly -- no positions are ever exchanged, and a+b == b+a keeps both copies bit-identical.
"""
import numpy as np


def slab_range(n_layers, rank, world):
    """Cell layers [i0, i1) of `rank` when `n_layers` x-layers are dealt to `world` ranks."""
    base, rem = divmod(n_layers, world)
    i0 = rank * base + min(rank, rem)
    return i0, i0 + base + (1 if rank < rem else 0)


def slab_halo_vertices(rank, world, local_layers, ny, nz):
    """Local vertex ids (slab-local numbering of meshgen.kuhn_box(x_cells=...)) of the planes shared
    with rank-1 (left, local plane 0) and rank+1 (right, local plane `local_layers`).
    Returns (ids, n_left, n_right); the order inside a plane is the global (j, k) order on both sides."""
    plane = (ny + 1) * (nz + 1)
    left = np.arange(plane, dtype=np.int32) if rank > 0 else np.zeros(0, np.int32)
    right = (local_layers * plane + np.arange(plane, dtype=np.int32)) if rank < world - 1 else np.zeros(0, np.int32)
    return np.concatenate([left, right]), left.size, right.size


def owner_mask(rank, world, local_layers, ny, nz):
    """1.0 for vertices this rank counts in global reductions (a shared plane belongs to the LEFT rank)."""
    plane = (ny + 1) * (nz + 1)
    m = np.ones((local_layers + 1) * plane, dtype=np.float32)
    if rank > 0:
        m[:plane] = 0.0
    return m


class SlabHalo:
    """Halo exchange of one rank of an x-slab partition, over torch.distributed P2P (NCCL on GPUs)."""

    def __init__(self, rank, world, local_layers, ny, nz):
        self.rank, self.world = rank, world
        self.vertex_ids, self.n_left, self.n_right = slab_halo_vertices(rank, world, local_layers, ny, nz)
        self.n = int(self.vertex_ids.size)
        self.send = self.recv = None
        self._graphs = {}

    def attach(self, handle, device=None):
        """Allocate the send/recv buffers (torch, on the current CUDA device), bind them to the
        library object and complete the lumped masses of the shared vertices."""
        import torch

        self.torch = torch
        self.send = torch.zeros((self.n, 4), dtype=torch.float32, device=device or "cuda")
        self.recv = torch.zeros((self.n, 4), dtype=torch.float32, device=device or "cuda")
        handle.halo_bind(self.send.data_ptr(), self.recv.data_ptr())
        handle.halo_mass(1)
        self.exchange()
        handle.halo_mass(2)
        torch.cuda.current_stream().synchronize()

    def exchange(self):
        """send[left part] -> rank-1, send[right part] -> rank+1; recv likewise (one batched P2P group)."""
        import torch.distributed as dist

        ops = []
        nl = self.n_left
        if self.n_left:
            ops.append(dist.P2POp(dist.isend, self.send[:nl], self.rank - 1))
            ops.append(dist.P2POp(dist.irecv, self.recv[:nl], self.rank - 1))
        if self.n_right:
            ops.append(dist.P2POp(dist.isend, self.send[nl:], self.rank + 1))
            ops.append(dist.P2POp(dist.irecv, self.recv[nl:], self.rank + 1))
        if ops:
            for r in dist.batch_isend_irecv(ops):
                r.wait()

    def _steps_eager(self, handle, dt, g_dir, n):
        for _ in range(n):
            handle.explicit_step_phase(dt, g_dir, 1)
            self.exchange()
            handle.explicit_step_phase(dt, g_dir, 2)

    def explicit_steps(self, handle, dt, g_dir, n, use_graph=True):
        """n fused explicit steps with a halo exchange in the middle of each.  The whole sequence
        (kernels + NCCL P2P) is captured once into a CUDA graph and replayed."""
        torch = self.torch
        if not use_graph or n < 4:
            return self._steps_eager(handle, dt, g_dir, n)
        key = (n, float(dt), tuple(float(x) for x in g_dir))
        from . import _capi

        g = self._graphs.get(key)
        if g is None:
            # (NCCL channels were opened by attach(); capturing does not execute anything)
            torch.cuda.current_stream().synchronize()
            graph = torch.cuda.CUDAGraph()
            before = _capi.launch_count()
            with torch.cuda.graph(graph, stream=torch.cuda.current_stream(), capture_error_mode="thread_local"):
                self._steps_eager(handle, dt, g_dir, n)
            g = self._graphs[key] = (graph, _capi.launch_count() - before)
            self.captured_launches = getattr(self, "captured_launches", 0) + g[1]
        g[0].replay()
        self.replayed_launches = getattr(self, "replayed_launches", 0) + g[1]
