"""Two ranks on one GPU (gloo), root/worker form: print the root's trajectory."""
import os, socket, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

def _worker(rank, world, port, batched, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    import torch, torch.distributed as dist
    torch.cuda.set_device(0)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import hypatia_b200  # noqa
    from hypatia_b200 import instances
    from hypatia_b200.gpu import make_gpu_solver
    s = make_gpu_solver(0, dist=dist, trace=True, mode="rootworker", batched_dirs=batched)
    log = []
    s.iter_callback = lambda sv: log.append((sv.stepper.last_num_ref, sv.stepper.last_num_cand, sv.worst_dir_res, sv.mu))
    s.load(instances.linearopt(m=60, n=200, seed=3)); s.solve()
    if rank == 0:
        out.put((s.status, s.num_iters, s.primal_obj, log))
    dist.destroy_process_group()

if __name__ == "__main__":
    import torch.multiprocessing as mp
    for batched in (True, False, True):
        ctx = mp.get_context("spawn"); out = ctx.Queue()
        with socket.socket() as sk:
            sk.bind(("127.0.0.1", 0)); port = sk.getsockname()[1]
        ps = [ctx.Process(target=_worker, args=(r, 2, port, batched, out)) for r in range(2)]
        [p.start() for p in ps]
        st, it, obj, log = out.get(timeout=300)
        [p.join() for p in ps]
        print("batched" if batched else "unbatched", st, it, obj)
        for i, l in enumerate(log): print("  ", i, l)
