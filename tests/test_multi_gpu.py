"""Multi-GPU slab path.  CPU part (gloo, world_size 2): partition helpers and the P2P exchange
pattern of SlabHalo.  GPU part: 2-rank slab run vs the single-GPU run (needs >= 2 GPUs)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_slab_partition_helpers():
    from femb200 import dist as fdist, meshgen

    nx, ny, nz = 7, 3, 2
    plane = (ny + 1) * (nz + 1)
    covered = []
    for world in (1, 2, 3, 4):
        owned = 0
        for r in range(world):
            i0, i1 = fdist.slab_range(nx, r, world)
            covered.append((i0, i1))
            ids, nl, nr = fdist.slab_halo_vertices(r, world, i1 - i0, ny, nz)
            assert nl == (plane if r > 0 else 0) and nr == (plane if r < world - 1 else 0)
            verts, tets = meshgen.kuhn_box(nx, ny, nz, 0.1, x_cells=(i0, i1))
            assert len(verts) == (i1 - i0 + 1) * plane and len(tets) == (i1 - i0) * ny * nz * 6
            gverts, _ = meshgen.kuhn_box(nx, ny, nz, 0.1)
            assert np.allclose(verts, gverts[i0 * plane:(i1 + 1) * plane])
            if nr:   # my right plane is the neighbour's left plane, same (j,k) order
                assert np.array_equal(ids[nl:] - (i1 - i0) * plane, np.arange(plane))
            owned += int(fdist.owner_mask(r, world, i1 - i0, ny, nz).sum())
        assert owned == (nx + 1) * plane     # every global vertex counted exactly once
        assert fdist.slab_range(nx, world - 1, world)[1] == nx


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist

    sys.path[:0] = [os.path.join(ROOT, "finite-element-method_b200"), ROOT]
    from femb200 import dist as fdist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    halo = fdist.SlabHalo(rank, world, 3, 2, 2)
    halo.send = torch.full((halo.n, 4), float(rank + 1))
    halo.recv = torch.zeros((halo.n, 4))
    halo.exchange()
    nl = halo.n_left
    ok = True
    if nl:
        ok &= bool((halo.recv[:nl] == float(rank)).all())        # from rank-1
    if halo.n_right:
        ok &= bool((halo.recv[nl:] == float(rank + 2)).all())     # from rank+1
    q.put((rank, ok))
    dist.destroy_process_group()


def test_halo_exchange_pattern_gloo_world2():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


class _FakeHandle:
    """Scripted stand-in for the library's solve state machine (femb_implicit_solve_begin/_advance/...):
    the same communication requests in the same order, with host tensors, so the driver logic of
    SlabHalo.implicit_solve runs under gloo without a GPU."""

    PROLOGUE = [1, 2, 3]            # HALO, ALLREDUCE, ITER
    ITERATION = [1, 1, 2, 2, 3]     # A^T A-CG: two operator passes, two dots

    def __init__(self, halo, rank, stop_after):
        self.halo, self.rank, self.stop_after = halo, rank, stop_after
        self.script, self.pc, self.iters, self.log = None, 0, 0, []
        self.pending = None

    def solve_begin(self, dt, method, pre, max_iter, tol):
        self.script, self.pc, self.iters = self.PROLOGUE, 0, 0

    def _consume(self):
        if self.pending == 1:
            self.log.append(("halo", self.halo.recv.clone()))
        elif self.pending == 2:
            self.log.append(("sum", float(self.halo.allred[0])))
        self.pending = None

    def solve_advance(self):
        self._consume()
        k = self.script[self.pc]
        self.pc += 1
        if k == 1:
            self.halo.send[:] = float(10 * self.iters + self.rank + 1)
        elif k == 2:
            self.halo.allred[0] = float(self.rank + 1 + self.iters)
        elif k == 3:
            if self.script is self.ITERATION:
                self.iters += 1
            self.script, self.pc = self.ITERATION, 0
        self.pending = k
        return k

    def solve_poll(self):
        return self.iters >= self.stop_after, self.iters, 0.0

    def solve_finish(self, fetch=True):
        self._consume()
        return self.iters, 0.0


def _gloo_solve_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist

    sys.path[:0] = [os.path.join(ROOT, "finite-element-method_b200"), ROOT]
    from femb200 import dist as fdist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    halo = fdist.SlabHalo(rank, world, 3, 2, 2)
    halo.send, halo.recv = torch.zeros((halo.n, 4)), torch.zeros((halo.n, 4))
    halo.allred = torch.zeros(1, dtype=torch.float64)
    fh = _FakeHandle(halo, rank, stop_after=10)
    iters, _ = halo.implicit_solve(fh, 1e-4, 1, 1, max_iter=50, tol=1e-5)
    other = 1 - rank
    ok = iters == 12                      # polled after 4 (not done) and after 4 + 8 (done): chunks double
    sums = [v for k, v in fh.log if k == "sum"]
    halos = [v for k, v in fh.log if k == "halo"]
    ok &= len(sums) == 1 + 2 * iters and len(halos) == 1 + 2 * iters
    ok &= sums[0] == 3.0 and all(s == 3.0 + 2 * (i // 2) for i, s in enumerate(sums[1:]))
    ok &= all(bool((hv == hv[0, 0]).all()) for hv in halos)
    ok &= float(halos[0][0, 0]) == other + 1 and float(halos[-1][0, 0]) == 10 * (iters - 1) + other + 1
    q.put((rank, bool(ok), iters))
    dist.destroy_process_group()


def test_implicit_solve_driver_gloo_world2():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_solve_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True, 12), (1, True, 12)], res


@pytest.mark.gpu
@pytest.mark.parametrize("comm,extra", [
    ("peer", {}),                                   # default: halo tiles | sender CTAs + other tiles | boundary kernel
    ("peer", {"FEMB_PEER_MODE": "one"}),            # ONE tile kernel per pass, sender CTAs inside it (halo-tile counter)
    ("peer", {"FEMB_PEER_FUSED_WAVES": "0"}),       # thick-slab variant: persistent operator kernels + send kernel
    ("peer", {"FEMB_PEER_FUSED": "0"}),             # separate send / receive / reduce kernels
    ("peer", {"FEMB_PEER_MODE": "late"}),           # one tile kernel; sender CTAs at the front of the boundary kernel
    ("nccl", {}),
])
def test_two_rank_slabs_match_single_gpu(comm, extra):
    """comm=peer: halo sums / dot products inside the library's kernels over NVLink peer memory (every
    kernel arrangement of csrc/object.h: peer_pass); comm=nccl: the same phases driven from Python with
    torch.distributed P2P + all_reduce."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, FEMB200_COMM=comm, **extra)
    if extra.get("FEMB_PEER_MODE") == "one":        # lives in the diagnostics build (csrc/dev.cuh: FEMB_DIAG_BUILD)
        diag = os.path.join(ROOT, "finite-element-method_b200", "femb200", "libfemb200_diag.so")
        if not os.path.exists(diag):
            pytest.skip("libfemb200_diag.so not built (make -C finite-element-method_b200/csrc diag)")
        env["FEMB200_LIB"] = diag
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29611",
                        os.path.join(ROOT, "tests", "mp_slab_check.py")], capture_output=True, text=True, timeout=180,
                       env=env)
    assert "SLAB_OK" in r.stdout and ("COMM " + comm) in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.gpu
def test_public_api_on_two_ranks_matches_single_gpu():
    """object.Object cuts the box into slabs under torchrun; neo_hookean_1_grad / kinematic /
    implicit_solver_neo_hookean / advect_implicit run unchanged (main.py:101-113) and match the 1-GPU run."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29613",
                        os.path.join(ROOT, "tests", "mp_public_api_check.py")], capture_output=True, text=True, timeout=300)
    assert "PUBLIC_SLAB_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.gpu
def test_main_py_under_torchrun(tmp_path):
    """`torchrun --nproc-per-node 2 main.py --config config/cube_1m.json --frames 2`: the driver runs on 2 GPUs with no new calls."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    pkg = os.path.join(ROOT, "finite-element-method_b200")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29615", os.path.join(pkg, "main.py"),
                        "--config", os.path.join(pkg, "config", "cube_1m.json"), "--frames", "2", "--no-render",
                        "--out", str(tmp_path)], capture_output=True, text=True, timeout=300, cwd=pkg)
    assert r.returncode == 0 and "element-updates/s" in r.stdout and "2 rank(s)" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
