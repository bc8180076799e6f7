"""Driver of the two-GPU tests (tests/test_gpu_multi.py), one process per GPU under torch.distributed.run:
every rank renders ITS part of a frame through the C ABI, the C entries vx_cuda_frame_gather /
vx_cuda_frame_gather_inplace assemble it on rank 0, and rank 0 compares it word for word with the frame it
renders alone on its own GPU.  torch.distributed only carries the communicator id and the verdict.
usage: python -m torch.distributed.run --nproc-per-node N tests/_multi_driver.py rt|raster"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main(kind):
    import torch
    import torch.distributed as dist

    from skybox_rt_b200 import vortex as vx
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    os.environ["VX_CUDA_DEVICE"] = str(local)
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")  # id exchange only: the data path is NCCL inside libvortex.so
    def new_id():  # one NCCL id per communicator, made on rank 0
        ident = [vx.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        return ident[0]

    bad = 0
    if kind == "rt":
        from skybox_rt_b200 import raytrace as rt
        nodes, tris = rt.build_bvh2(rt.scene_terrain(80, 60))
        for (w, h, spp, rpb) in ((320, 200, 1, 8), (203, 77, 2, 4), (64, 5, 1, 8)):  # the last: fewer row blocks than ranks
            with rt.RayTracer(nodes, tris) as tracer:
                dev = tracer.dev
                comm = vx.Comm(dev, new_id(), rank, world)
                arg = rt.make_arg(w, h, 2, spp=spp, seed=17, scene="terrain", rank=rank, num_ranks=world, rows_per_block=rpb)
                tracer.start(arg)
                frame = dev.alloc(w * h * 4) if rank == 0 else None
                comm.frame_gather(tracer.cbuf, frame, w, h, rpb, 0)
                dev.ready_wait()
                if rank == 0:
                    got = np.empty((h, w), np.uint32)
                    frame.read(got)
                    ref = tracer.render(rt.make_arg(w, h, 2, spp=spp, seed=17, scene="terrain"))
                    bad += int((got != ref).sum())
                    frame.free()
                comm.close()
    else:
        import _oracle_draw3d as od
        from skybox_rt_b200 import draw3d as d3
        for name, w, h in (("vase", 256, 200), ("tekkaman", 330, 517)):
            t = d3.Trace(blob=od.trace_blob(name))
            with vx.Device() as dev:
                comm = vx.Comm(dev, new_id(), rank, world)
                app = d3.Draw3D(w, h, flags=7 | d3.WIDE_TILES, rank=rank, num_ranks=world, device=dev)
                app.clear()
                app.render(t)
                comm.frame_gather_inplace(app.color_buffer_handle(), w, h, 32, 0)
                dev.ready_wait()
                if rank == 0:
                    got = app.read().copy()
                    solo = d3.Draw3D(w, h, flags=7 | d3.WIDE_TILES, device=dev)
                    ref = solo.frame(t)
                    bad += int((got != ref).sum())
                    solo.close()
                app.close()
                comm.close()
    verdict = [bad]
    dist.broadcast_object_list(verdict, src=0)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print(f"{kind}: {verdict[0]} words differ from the single-GPU frame")
    sys.exit(1 if verdict[0] else 0)


if __name__ == "__main__":
    main(sys.argv[1])
