"""CC3 on N GPUs (one process per GPU, torchrun): the golden o6v13 solve against the reference's energy, then a
synthetic (o, v) problem timed per iteration.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        tools/cc3_multi_gpu.py [o v]
Prints one JSON line on rank 0."""
import contextlib, io, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import pycc_b200
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("nccl")
group = "world" if world > 1 else None
here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
z = np.load(os.path.join(here, "tests", "golden", "o6v13.npz"))
zm = np.load(os.path.join(here, "tests", "golden", "models", "o6v13.npz"))
sys.path.insert(0, here)
from tests.conftest import Golden
g = Golden("o6v13")
out = dict(n_gpus=world)
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(g.F, dict(g.blocks), g.no, model="CC3", group=group)
    e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
out["golden_dE"] = abs(e - float(zm["cc3_solve_ecc"]))
out["golden_niter"] = [len(w.iteration_log), int(zm["cc3_solve_niter"])]
out["golden_dt2"] = float(np.abs(w.t2.cpu().numpy() - zm["cc3_solve_t2"]).max())
o, v = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (10, 60)
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
blocks = device_blocks(B, o, torch.device("cuda", torch.cuda.current_device()))
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), blocks, o, model="CC3", group=group)
    d = w.make_diis(8)
    for n in range(1, 3):
        w._iteration(d, n, 1)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.time()
    for n in range(3, 6):
        ecc, rms = w._iteration(d, n, 1)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = (time.time() - t0) / 3
out.update(o=o, v=v, s_per_iteration=dt, ecc_after_5=ecc, rms=rms)
if rank == 0:
    print(json.dumps(out))
if world > 1:
    dist.destroy_process_group()
