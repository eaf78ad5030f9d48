"""debug: 2-rank implicit solve, with a traceback dump if it stalls"""
import faulthandler, os, sys, time
faulthandler.dump_traceback_later(40, exit=True)
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "finite-element-method_b200"), ROOT]
from femb200 import _capi as c, dist as fdist, meshgen
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
def P(*a):
    print("[r%d %.2f]" % (rank, time.time() % 1000), *a, file=sys.stderr, flush=True)
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
nx, ny, nz, hh = 9, 6, 5, 0.4 / 55
center = np.array([0.3, 0.5, 0.3], np.float32)
mu, lam = 4e4 / 2 / 1.4, 4e4 * 0.4 / 1.4 / 0.2
plane = (ny + 1) * (nz + 1)
i0, i1 = fdist.slab_range(nx, rank, world)
verts, tets = meshgen.kuhn_box(nx, ny, nz, hh, x_cells=(i0, i1))
halo = fdist.SlabHalo(rank, world, i1 - i0, ny, nz)
h = c.Handle(verts, tets, center, 1000.0, mu, lam, 10.0, implicit=True, tile_elems=96, halo_verts=halo.vertex_ids, halo_owned=halo.owned)
h.set_stream(stream.cuda_stream)
halo.attach(h)
P("attached")
jit = np.random.default_rng(0).uniform(-1, 1, ((nx + 1) * plane, 3)).astype(np.float32)[i0 * plane:(i1 + 1) * plane]
h.set(c.F_POS, (verts + center).astype(np.float32) + np.float32(0.1 * hh) * jit)
h.solve_begin(1e-4, 1, 0, 0, -1.0)
for step in range(40):
    k = h.solve_advance()
    P("advance ->", k)
    if k == 1: halo.exchange(); torch.cuda.synchronize(); P("exchanged")
    elif k == 2: halo.allreduce(); torch.cuda.synchronize(); P("allreduced", float(halo.allred[0]))
    elif k == 3:
        P("iter end, poll", h.solve_poll())
        if h.solve_poll()[0]: break
P("finish", h.solve_finish())
for name, m, pre, dt, g in [("cg-eager", 1, 0, 1e-4, False), ("ata-graph", 1, 1, 5e-4, True), ("ata-graph2", 1, 1, 5e-4, True), ("jacobi", 0, 0, 1e-4, True)]:
    P(name, halo.implicit_solve(h, dt, m, pre, use_graph=g))
torch.cuda.synchronize()
P("DONE")
halo.close()
h.close()
fdist.shutdown()
P('shutdown ok')
