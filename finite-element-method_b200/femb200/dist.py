"""Multi-GPU plumbing: one process per GPU (torch.distributed / NCCL), structured box cut into
x-slabs.  The reference is single-device (main.py:40); this is new.

Each element lives on exactly one rank; the vertex plane between two neighbouring slabs exists
on both.  After the tile kernel every rank packs the per-vertex sum of its local contributions on
the shared planes, exchanges it with its (at most two) neighbours, and finishes the shared vertices
redundantly -- no positions are ever exchanged, and a+b == b+a keeps both copies bit-identical.
"""
import numpy as np


def rank_world():
    """(rank, world) of this process: torch.distributed if it is initialised, else RANK / WORLD_SIZE from the
    environment (torchrun), else (0, 1)."""
    import os
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except ImportError:
        pass
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def ensure_process_group(backend=None):
    """Under torchrun (WORLD_SIZE > 1) initialise the default process group once: NCCL with one GPU per rank
    (LOCAL_RANK), gloo when no GPU is present (CPU tests).  Returns (rank, world)."""
    import os
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            if backend == "nccl":
                torch.cuda.set_device(local_device())
                dist.init_process_group("nccl", device_id=torch.device("cuda", local_device()))
            else:
                dist.init_process_group(backend)
    return rank_world()


def local_device():
    import os
    return int(os.environ.get("LOCAL_RANK", "0"))


def make_stream(handle):
    """A non-default CUDA stream made current for torch and bound to the handle (the legacy default stream
    cannot be graph-captured)."""
    import torch
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    handle.set_stream(stream.cuda_stream)
    return stream


def all_gather_arrays(arr):
    """List of every rank's numpy array (different lengths allowed), on every rank."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size()
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    a = np.ascontiguousarray(arr)
    n = torch.tensor([a.shape[0]], dtype=torch.int64, device=dev)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    t = torch.zeros((max(sizes),) + a.shape[1:], dtype=torch.from_numpy(a).dtype, device=dev)
    t[:a.shape[0]] = torch.from_numpy(a).to(dev)
    outs = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(outs, t)
    if dev.type == "cuda":
        torch.cuda.current_stream().synchronize()
    return [o[:k].cpu().numpy() for o, k in zip(outs, sizes)]


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


def slab_halo_owned(rank, world, local_layers, ny, nz):
    """Per halo vertex (order of slab_halo_vertices): 1 if THIS rank counts it in global reductions.
    A shared plane belongs to the LEFT rank, so a rank owns its right plane and not its left one."""
    _, nl, nr = slab_halo_vertices(rank, world, local_layers, ny, nz)
    return np.concatenate([np.zeros(nl, np.uint8), np.ones(nr, np.uint8)])


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
        self.owned = slab_halo_owned(rank, world, local_layers, ny, nz)
        self.allred = None
        self.n = int(self.vertex_ids.size)
        self.send = self.recv = None
        self.peer = False
        self._graphs = {}

    def attach_peer(self, handle):
        """Peer-memory mode: one symmetric comm buffer per rank, mapped by every rank over NVLink
        (torch symmetric memory); the library's own kernels then do the halo sums and the dot-product
        sums with remote stores + flags -- no NCCL call on the data path.  Returns False (after telling
        why on stderr) when symmetric memory cannot be set up; the caller then uses the NCCL path."""
        import sys

        import torch
        import torch.distributed as dist

        from . import _capi

        self.torch = torch
        try:
            import torch.distributed._symmetric_memory as symm

            plane = max(self.n_left, self.n_right)
            nbytes, stride = _capi.comm_bytes(2 * plane)       # the same layout on every rank
            if hasattr(symm, "enable_symm_mem_for_group"):
                import warnings

                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")
                    symm.enable_symm_mem_for_group(dist.group.WORLD.group_name)
            buf = symm.empty(nbytes, dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
            buf.zero_()
            torch.cuda.synchronize()
            try:
                hdl = symm.rendezvous(buf, dist.group.WORLD)
            except TypeError:
                hdl = symm.rendezvous(buf, dist.group.WORLD.group_name)
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            dist.barrier()                                      # every rank's buffer is zeroed and mapped
            torch.cuda.synchronize()
        except Exception as ex:  # noqa: BLE001 -- any failure here just selects the NCCL path
            print("femb200.dist: peer-memory mode unavailable (%s: %s); using NCCL" % (type(ex).__name__, ex),
                  file=sys.stderr, flush=True)
            return False
        self._symm = (buf, hdl)                                 # keep the mapping alive
        left = self.rank - 1 if self.n_left else -1
        right = self.rank + 1 if self.n_right else -1
        # my left plane is the left neighbour's RIGHT plane: it follows that rank's own left plane (if any)
        left_dst_off = self.n_left if self.rank - 1 > 0 else 0
        handle.peer_bind(self.rank, self.world, left, right, self.n_left, left_dst_off, 0, stride, ptrs)
        handle.peer_mass()
        dist.barrier()
        self.peer = True
        return True

    def attach(self, handle, device=None, mode=None):
        """Set up the halo machinery and complete the lumped masses of the shared vertices.
        mode "peer": in-kernel communication over peer memory (default when available);
        mode "nccl": send/recv buffers exchanged by torch.distributed P2P, scalars by all_reduce."""
        import os

        import torch

        self.torch = torch
        mode = mode or os.environ.get("FEMB200_COMM", "peer")
        if mode == "peer" and self.world > 1 and self.attach_peer(handle):
            return
        self.send = torch.zeros((self.n, 4), dtype=torch.float32, device=device or "cuda")
        self.recv = torch.zeros((self.n, 4), dtype=torch.float32, device=device or "cuda")
        handle.halo_bind(self.send.data_ptr(), self.recv.data_ptr())
        self.allred = torch.zeros(4, dtype=torch.float64, device=device or "cuda")   # up to 4 sums per all-reduce
        handle.reduce_bind(self.allred.data_ptr())
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
        if self.peer:                  # the library runs (and graph-captures) the whole sequence itself
            return handle.explicit_step(dt, g_dir, n)
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

    # ---- implicit solve: the library enqueues device work up to a communication point, we communicate
    def allreduce(self):
        """Sum of the bound scalar over all ranks (identical bits on every rank, so the device-side
        convergence flags never diverge between ranks)."""
        import torch.distributed as dist

        dist.all_reduce(self.allred, op=dist.ReduceOp.SUM)

    def _until_iter(self, handle):
        """Advance to the end of the prologue / of one solver iteration."""
        from . import _capi

        n = 0
        while True:
            k = handle.solve_advance()
            n += 1
            if k == _capi.COMM_HALO:
                self.exchange()
            elif k == _capi.COMM_ALLREDUCE:
                self.allreduce()
            elif k == _capi.COMM_ITER:
                return n
            else:
                raise RuntimeError("unexpected communication request %r" % (k,))

    def implicit_solve(self, handle, dt, method, preconditioned, max_iter=0, tol=-1.0, use_graph=True, fetch=True):
        """implicit_solver_neo_hookean (implicit.py:201-223) on a slab-partitioned mesh: halo sums after every
        operator pass, one scalar allreduce per dot product.  One iteration (kernels + NCCL) is captured
        into a CUDA graph and replayed; convergence is polled every few iterations."""
        from . import _capi

        if self.peer:                  # halo sums and dot products happen inside the library's kernels
            return handle.implicit_solve(dt, method, preconditioned, max_iter, tol, fetch=fetch)
        fixed = (tol == 0.0)
        n_cap = max_iter if max_iter > 0 else (500 if method == 1 else 20000)
        # single-reduction CG (A^T A) sees r.r of update k in iteration k + 1; Jacobi needs 2 more passes (api_solve.cu)
        n_it = (n_cap + (1 if (preconditioned == 1 and not fixed) else 0)) if method == 1 else n_cap + 2   # (+1: harmless for the textbook loop)
        handle.solve_begin(dt, method, preconditioned, max_iter, tol)
        self._until_iter(handle)                       # prologue: assemble, A^T b, initial residual
        graph = None
        if use_graph and n_it >= 4 and self.send.is_cuda:
            torch = self.torch
            key = ("solve", float(dt), int(method), int(preconditioned), int(max_iter), float(tol))
            graph = self._graphs.get(key)
            if graph is None:
                self._until_iter(handle)               # iteration 1 runs eagerly (warms NCCL for these sizes)
                n_it -= 1
                torch.cuda.current_stream().synchronize()
                g = torch.cuda.CUDAGraph()
                before = _capi.launch_count()
                with torch.cuda.graph(g, stream=torch.cuda.current_stream(), capture_error_mode="thread_local"):
                    self._until_iter(handle)           # captured, not executed
                graph = self._graphs[key] = (g, _capi.launch_count() - before)
        it, chunk = 0, 4
        while it < n_it:
            n = n_it - it if fixed else min(chunk, n_it - it)
            for _ in range(n):
                if graph is not None:
                    graph[0].replay()
                    self.replayed_launches = getattr(self, "replayed_launches", 0) + graph[1]
                else:
                    self._until_iter(handle)
            it += n
            if not fixed:
                done, _, _ = handle.solve_poll()
                if done:
                    break
                chunk = min(chunk * 2, 32)
        return handle.solve_finish(fetch)

    def close(self):
        """Drop the captured CUDA graphs (they hold NCCL work: the process group cannot be destroyed while
        they are alive)."""
        self._graphs.clear()
        if self.peer or (self.send is not None and self.send.is_cuda):
            self.torch.cuda.synchronize()


def shutdown(halo=None, timeout=20.0):
    """Release the graphs, then destroy the process group; if NCCL teardown stalls (seen with graph-captured
    collectives), leave after `timeout` seconds -- all results have been produced by then."""
    import os
    import sys
    import threading

    import torch.distributed as dist

    if halo is not None:
        halo.close()
    sys.stdout.flush()
    sys.stderr.flush()
    t = threading.Timer(timeout, lambda: os._exit(0))
    t.daemon = True
    t.start()
    dist.destroy_process_group()
    t.cancel()
