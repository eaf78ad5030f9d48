# coding=utf-8
"""Headless driver with the reference's command line (main.py:32-133):

    python main.py [--config file.json] [--frames N] [--out DIR] [--checkpoint FILE] [--resume FILE]

Same import order (config -> sys_init -> solver modules, main.py:36-49), same step loop
(main.py:101-113: sim_count x [fem | compute_energy ; kinematic | advect_implicit] per frame), same
virtual-time bookkeeping for frame export (main.py:115-124).  The reference's window loop
(`while widget.running`) becomes a frame count: there is no graphics stack on the box, so the GUI /
GGUI renderer is replaced by a headless Render that writes frames to files (render/render.py).
Extras: per-frame energy / inverted-element / NaN diagnostics and .npz checkpoints.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

import utils


def fem(soft_obj):
    # main.py:8-12
    if constants.use_explicit_method:
        neo_hookean_1_grad(soft_obj)
    else:
        implicit_solver_neo_hookean(soft_obj)


def step(soft_obj, circle_blocks):
    """One simulation step, main.py:103-112."""
    if constants.auto_diff == 0:
        fem(soft_obj)
    else:
        compute_energy(soft_obj)           # the reference wraps this in ti.ad.Tape(loss=soft_obj.U)
    if constants.use_explicit_method or constants.auto_diff == 1:
        ki.kinematic(soft_obj, circle_blocks)
    else:
        advect_implicit(soft_obj, circle_blocks)


def save_checkpoint(path, soft_objects, frame, virtual_time, ply_cnt=0, output_frame_cnt=0):
    data = {"frame": frame, "virtual_time": virtual_time, "n_objects": len(soft_objects), "ply_cnt": ply_cnt,
            "output_frame_cnt": output_frame_cnt}
    for k, o in enumerate(soft_objects):
        data["pos%d" % k] = o.particles.pos.to_numpy()
        data["vel%d" % k] = o.particles.vel.to_numpy()
        if o.implicit:
            data["vel_g%d" % k] = o.particles.vel_g.to_numpy()
            data["past_x%d" % k] = o.past_vec_x.to_numpy()   # stale-by-design Jacobi state (implicit.py:231-232)
    np.savez(path, **data)


def load_checkpoint(path, soft_objects):
    z = np.load(path)
    if int(z["n_objects"]) != len(soft_objects):
        raise ValueError("checkpoint holds %d objects, config has %d" % (int(z["n_objects"]), len(soft_objects)))
    for k, o in enumerate(soft_objects):
        o.particles.pos.from_numpy(z["pos%d" % k])
        o.particles.vel.from_numpy(z["vel%d" % k])
        if o.implicit and ("vel_g%d" % k) in z:
            o.particles.vel_g.from_numpy(z["vel_g%d" % k])
            o.past_vec_x.from_numpy(z["past_x%d" % k])
    counters = (int(z["ply_cnt"]) if "ply_cnt" in z else None, int(z["output_frame_cnt"]) if "output_frame_cnt" in z else None)
    return int(z["frame"]), float(z["virtual_time"]), counters


def diagnostics(soft_obj):
    """Total elastic energy, inverted-element count and NaN count of the current state.  Uses a scratch
    evaluation of compute_energy; pos.grad is overwritten (harmless outside auto_diff mode, and in
    auto_diff mode the next step recomputes it before use)."""
    U = soft_obj._h.energy_grad(fetch=True)
    d = soft_obj._h.diagnostics()      # three counters from the device; no field leaves HBM
    return {"U": float(U), "nan_elements": d["nan_elements"], "nan_vertices": d["nan_vertices"],
            "inverted_elements": d["inverted_elements"]}


if __name__ == '__main__':
    parser = argparse.ArgumentParser(description='FEM on B200 (headless)')
    parser.add_argument('--config', help="Please input a config json file.", type=str, default='default.json')
    parser.add_argument('--frames', type=int, default=60, help="rendered frames to run (the reference runs until the window closes)")
    parser.add_argument('--out', type=str, default='./output')
    parser.add_argument('--checkpoint', type=str, default=None, help="write an .npz checkpoint here at the end")
    parser.add_argument('--resume', type=str, default=None, help="restore pos/vel/vel_g/past_x from this .npz first")
    parser.add_argument('--diag-every', type=int, default=0, help="print energy / inversion diagnostics every N frames")
    parser.add_argument('--no-render', action='store_true')
    args = parser.parse_args()
    config = utils.read_config(args.config)
    utils.sys_init(config)
    # extension: under torchrun (one process per GPU) structured 3D boxes are cut into x-slabs by object.Object
    from femb200 import dist as fdist
    rank, world = fdist.ensure_process_group()
    if world > 1:
        from femb200 import _capi
        _capi.set_host_threads(max(1, (os.cpu_count() or 1) // world))   # torchrun exports OMP_NUM_THREADS=1

    import constants
    from render.render import Render
    import solver.kinematic as ki
    from solver.implicit import implicit_solver_neo_hookean, advect_implicit
    from solver.explicit_auto_diff import compute_energy
    from solver.explicit import neo_hookean_1_grad
    from object import Object
    from circle_blocks import CircleBlocks

    os.makedirs(args.out, exist_ok=True)
    render = None if (args.no_render or rank != 0 and world > 1) else Render(config, output_dir=args.out)
    sim_count = config.get('sim_count')
    circle_blocks = CircleBlocks(config.get('blocks'))
    soft_objects = [Object(i) for i in config.get('objects')]

    is_output_obj = config.get('is_output_obj')
    output_fps = config.get('output_fps', 60)
    frame_time = 1.0 / output_fps
    virtual_time = 0.0
    dt = config.get("delta_time")
    ply_cnt = 0
    frame0 = 0
    if args.resume:
        frame0, virtual_time, (ck_ply, ck_out) = load_checkpoint(args.resume, soft_objects)
        # the export counters continue where the checkpointed run stopped (no frame is written twice)
        ply_cnt = ck_ply if ck_ply is not None else int(virtual_time / frame_time) + 1
        if render is not None and ck_out is not None:
            render.output_frame_cnt = ck_out
    if constants.use_explicit_method:
        print('Simulation method: explicit method. Auto-diff {}'.format(bool(constants.auto_diff)))
    elif constants.implicit_method == constants.JACOBIN_METHOD:
        print('Simulation method: implicit method. System Solver: jacobian iteration.')
    elif constants.implicit_method == constants.CONJUGATE_GRADIENT_METHOD:
        print('Simulation method: implicit method. System Solver: conjugate gradient. Preconditioned: {}'.format(
            bool(constants.preconditioned)))

    t0 = time.perf_counter()
    for frame_cnt in range(frame0, frame0 + args.frames):
        for soft_obj in soft_objects:
            for i in range(sim_count):
                step(soft_obj, circle_blocks)
            virtual_time += sim_count * dt
            # main.py:115-124: export at output_fps of *virtual* time
            if virtual_time / frame_time > ply_cnt:
                if is_output_obj and constants.dim == 3:
                    soft_obj.update_obj()
                    soft_obj.save_obj(os.path.join(args.out, 'obj_{:06d}.obj'.format(ply_cnt)))   # main.py:121 name
                ply_cnt += 1
        if render is not None:
            msgs = []
            if not render.is_output_gif:                 # main.py:124-127
                msgs.append("frame_cnt: {}".format(frame_cnt + 1))
                msgs.append("time: {:.4f}".format(virtual_time))
            render.render(soft_objects, circle_blocks, virtual_time, msgs)
        if args.diag_every and (frame_cnt + 1) % args.diag_every == 0:
            for k, o in enumerate(soft_objects):
                print('frame {} object {}: {}'.format(frame_cnt + 1, k, json.dumps(diagnostics(o))))
    for o in soft_objects:
        o.sync()
    wall = time.perf_counter() - t0
    n_elem = sum(o.element_cnt for o in soft_objects)
    print('{} frames x {} steps, {} elements{}: {:.3f} s wall, {:.3e} element-updates/s'.format(
        args.frames, sim_count, n_elem, ' on this rank of {} rank(s)'.format(world) if world > 1 else '', wall,
        args.frames * sim_count * n_elem / max(wall, 1e-9)))
    if render is not None:
        render.finish()
    if args.checkpoint:
        save_checkpoint(args.checkpoint, soft_objects, frame0 + args.frames, virtual_time, ply_cnt,
                        render.output_frame_cnt if render is not None else 0)
        print('checkpoint written to {}'.format(args.checkpoint))
    sys.stdout.flush()
    if world > 1:
        fdist.shutdown()
