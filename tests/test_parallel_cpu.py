"""Multi-rank host logic on the CPU: two processes over ``gloo`` (127.0.0.1), each driving the
engine's sharded plans through the NumPy interpreter back end.  Checks that the virtual-index
sharding of the CCSD residual (+ allreduce) and the pair dealing of (T) reproduce the
reference-generated golden vectors, rank by rank."""
import contextlib
import io
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.conftest import Golden
        from tests.interp_backend import InterpBackend
        import pycc_b200
        from pycc_b200.parallel import virtual_slice
        g = Golden(name)
        lo, hi = virtual_slice(g.nv, rank, world)
        blocks = dict(g.blocks)
        if rank == 1:
            # one rank is handed only its own rows of <ab|ef>, the other the full block
            blocks["vvvv"] = np.ascontiguousarray(blocks["vvvv"][lo:hi])
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, blocks, g.no, model="CCSD(T)", group="world",
                                            _backend=InterpBackend())
            r1, r2 = w.residuals(g.F, g["gen_t1"], g["gen_t2"])
            d_r1 = float(np.abs(r1.numpy() - g["gen_r1"]).max())
            d_r2 = float(np.abs(r2.numpy() - g["gen_r2"]).max())
            w._load(None, g["gen_t1"], g["gen_t2"])
            d_et_gen = abs(pycc_b200.t_tjl(w) - float(g["gen_et_tjl"]))
            # fresh object: solve from the MP2 guess
            w = pycc_b200.ccwfn.from_arrays(g.F, blocks, g.no, model="CCSD(T)", group="world",
                                            _backend=InterpBackend())
            e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        # host <-> device amplitude shards: store this rank's slice, wipe, reload through the allgather
        slo, shi = w.amplitude_shard()
        hs = torch.empty(shi - slo, dtype=torch.float64)
        w.store_amplitude_shard(hs)
        keep1, keep2 = w.t1.clone(), w.t2.clone()
        w.t1.zero_()
        w.t2.zero_()
        w.load_amplitude_shard(hs)
        d_shard = float(max((w.t1 - keep1).abs().max(), (w.t2 - keep2).abs().max()))
        out = dict(rank=rank, vslice=(lo, hi), d_r1=d_r1, d_r2=d_r2, d_et_gen=d_et_gen, d_shard=d_shard,
                   shard=(slo, shi),
                   niter=len(w.iteration_log), d_e=abs(e - float(g["solve_eccsd_t_total"])),
                   d_t2=float(np.abs(w.t2.numpy() - g["solve_t2"]).max()),
                   vvvv_rows=int(w.engine.vvvv_loc.shape[0]), npairs=len(pycc_b200.cctriples._engine(w).my_pairs()),
                   reduced=w.comm.bytes_reduced, gathered=getattr(w.comm, 'bytes_gathered', 0),
                   diis_chunk=w.diis.chunk, diis_n=w.diis.n1 + w.diis.n2,
                   diis_pool=max(int(x.numel()) for x in w.diis.diis_vals + w.diis.diis_errors))
        q.put(out)
    except Exception as exc:      # surface the failure in the parent
        import traceback
        q.put(dict(rank=rank, error=traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def _sp_worker(rank, world, port, name, q):
    """Mixed-precision mode on two ranks: the plain-matrix ring-term statements with the sliced index."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.conftest import Golden
        from tests.interp_backend import InterpBackend
        import pycc_b200
        g = Golden(name)
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, precision="SP", group="world", _backend=InterpBackend())
            r1, r2 = w.residuals(g.F, g["gen_t1"], g["gen_t2"])
        q.put(dict(rank=rank, plain=bool(w.engine.sp_plain), rows=int(w.engine.tauWPn.shape[1]),
                   d_r1=float(np.abs(r1.double().numpy() - g["gen_r1"]).max()),
                   d_r2=float(np.abs(r2.double().numpy() - g["gen_r2"]).max())))
    except Exception:
        import traceback
        q.put(dict(rank=rank, error=traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_mixed_precision():
    from tests.conftest import Golden
    g = Golden("o6v13")
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sp_worker, args=(r, world, port, "o6v13", q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for d in res:
        assert "error" not in d, d.get("error")
        assert d["plain"] and d["d_r1"] < 2e-6 and d["d_r2"] < 2e-6
    assert sum(d["rows"] for d in res) == g.nv


def _lambda_worker(rank, world, port, name, model, q):
    """HBAR + Lambda with the v^4 / o v^3 blocks sharded by rows over two ranks."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.interp_backend import InterpBackend
        from tests import lambda_checks as lc
        import pycc_b200
        g, z = lc.load(name)
        tag = model.lower()
        t1 = np.zeros_like(g["gen_t1"]) if model == "CCD" else g["gen_t1"]
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model=model, group="world", full_ovvv=True,
                                            _backend=InterpBackend())
            w._load(None, t1, g["gen_t2"])
            hb = pycc_b200.cchbar(w)
            lam = pycc_b200.cclambda(w, hb)
            r1, r2 = lam.residuals(g.F, t1, g["gen_t2"], z[tag + "_l1"], z[tag + "_l2"])
            lo, hi = w.engine.vlo, w.engine.vhi
            d_hvvvo = float(np.abs(hb.Hvvvo.numpy() - z[tag + "_Hvvvo"][:, :, lo:hi]).max())
            out = dict(rank=rank, rows=int(hb.engine.HvT.shape[0]), d_hvvvo=d_hvvvo,
                       d_r1=float(np.abs(r1.numpy() - z[tag + "_rl1"]).max()),
                       d_r2=float(np.abs(r2.numpy() - z[tag + "_rl2"]).max()))
            w._load(None, z[tag + "_solve_t1"], z[tag + "_solve_t2"])
            hb = pycc_b200.cchbar(w)
            lam = pycc_b200.cclambda(w, hb)
            lecc = lam.solve_lambda(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        out.update(niter=len(lam.iteration_log), d_e=abs(lecc - float(z[tag + "_solve_lecc"])),
                   d_l2=float(np.abs(lam.l2.numpy() - z[tag + "_solve_l2"]).max()),
                   niter_ref=int(z[tag + "_solve_niter"]))
        q.put(out)
    except Exception:
        import traceback
        q.put(dict(rank=rank, error=traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name,model", [("o4v8", "CCSD"), ("o3v5_offdiag", "CCSD"), ("o4v8", "CCD"), ("o4v8", "CC2"),
                                        ("o3v5_offdiag", "CC2")])
def test_two_ranks_gloo_lambda(name, model):
    from tests.conftest import Golden
    g = Golden(name)
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_lambda_worker, args=(r, world, port, name, model, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for d in res:
        assert "error" not in d, d.get("error")
    assert sum(d["rows"] for d in res) == g.nv
    for d in res:
        assert d["d_hvvvo"] < 1e-12 and d["d_r1"] < 1e-12 and d["d_r2"] < 1e-12
        assert d["niter"] == d["niter_ref"] and d["d_e"] < 1e-11 and d["d_l2"] < 1e-9


@pytest.mark.parametrize("name", ["o4v8", "o5v19_offdiag"])
def test_two_ranks_gloo(name):
    from tests.conftest import Golden
    g = Golden(name)
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    res.sort(key=lambda d: d["rank"])
    for d in res:
        assert "error" not in d, d.get("error")
    assert res[0]["vslice"][1] == res[1]["vslice"][0] and res[1]["vslice"][1] == g.nv
    assert sum(d["vvvv_rows"] for d in res) == g.nv
    assert res[0]["shard"][0] == 0 and res[0]["shard"][1] == res[1]["shard"][0] \
        and res[1]["shard"][1] == g.no * g.nv + g.no ** 2 * g.nv ** 2
    assert sum(d["npairs"] for d in res) == g.no * (g.no + 1) // 2
    for d in res:
        assert d["d_r1"] < 1e-12 and d["d_r2"] < 1e-12
        assert d["d_et_gen"] < 1e-12
        assert d["niter"] == g["solve_iter_log"].shape[0]
        assert d["d_e"] < 1e-11 and d["d_t2"] < 1e-10
        assert d["reduced"] > 0
        assert d["d_shard"] == 0.0
        # DIIS history is sliced over the ranks: half-length vectors, one allgather per step
        # (by whole (i,j) blocks of t2 - the blocks the rank also updates - plus the t1 segment)
        n1, blk = g.no * g.nv, g.nv * g.nv
        assert d["diis_chunk"] * world >= d["diis_n"] and d["diis_chunk"] <= d["diis_n"] // 2 + n1 + blk
        assert d["diis_pool"] == d["diis_chunk"] and d["gathered"] > 0


def _cc3_worker(rank, world, port, name, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.conftest import Golden, GOLDEN_DIR
        from tests.interp_backend import InterpBackend
        import pycc_b200
        from pycc_b200.parallel import virtual_slice
        g = Golden(name)
        z = np.load(os.path.join(GOLDEN_DIR, "models", name + ".npz"))
        lo, hi = virtual_slice(g.nv, rank, world)
        blocks = dict(g.blocks)
        blocks["vvvv"] = np.ascontiguousarray(blocks["vvvv"][lo:hi])      # every rank: only its rows of <ab|ef>
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, blocks, g.no, model="CC3", group="world", _backend=InterpBackend())
            r1, r2 = w.residuals(g.F, g["gen_t1"], g["gen_t2"])
            c = w.engine.cc3
            out = dict(rank=rank, npairs=len(c.my_pairs()),
                       d_Wabei=float(np.abs(c.Wv.permute(1, 2, 3, 0).numpy() - z["cc3_Wabei"]).max()),
                       d_r1=float(np.abs(r1.numpy() - z["cc3_r1"]).max()),
                       d_r2=float(np.abs(r2.numpy() - z["cc3_r2"]).max()))
            w = pycc_b200.ccwfn.from_arrays(g.F, blocks, g.no, model="CC3", group="world", _backend=InterpBackend())
            e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        out.update(niter=len(w.iteration_log), niter_ref=int(z["cc3_solve_niter"]),
                   d_e=abs(e - float(z["cc3_solve_ecc"])), d_t2=float(np.abs(w.t2.numpy() - z["cc3_solve_t2"]).max()))
        q.put(out)
    except Exception:
        import traceback
        q.put(dict(rank=rank, error=traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8"])
def test_two_ranks_gloo_cc3(name):
    """CC3 on two ranks: each holds its rows of <ab|ef>, forms its slice of the dressed <ab|ei> (allreduce), and
    evaluates half of the occupied pairs; the partial triples terms are summed with the residual."""
    from tests.conftest import Golden
    g = Golden(name)
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_cc3_worker, args=(r, world, port, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for d in res:
        assert "error" not in d, d.get("error")
    assert sum(d["npairs"] for d in res) == g.no ** 2
    for d in res:
        assert d["d_Wabei"] < 1e-12 and d["d_r1"] < 1e-12 and d["d_r2"] < 1e-12
        assert d["niter"] == d["niter_ref"] and d["d_e"] < 1e-11 and d["d_t2"] < 1e-10


def _complex_worker(rank, world, port, name, model, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.conftest import Golden, GOLDEN_DIR
        from tests.interp_backend import InterpBackend
        import pycc_b200
        from pycc_b200.parallel import virtual_slice
        g = Golden(name)
        z = np.load(os.path.join(GOLDEN_DIR, "complex", name + ".npz"))
        lo, hi = virtual_slice(g.nv, rank, world)
        blocks = dict(g.blocks)
        blocks["vvvv"] = np.ascontiguousarray(blocks["vvvv"][lo:hi])
        tag = model.lower()
        t1 = np.zeros_like(z["t1"]) if model == "CCD" else z["t1"]
        out = dict(rank=rank)
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, blocks, g.no, model=model, group="world", full_ovvv=True,
                                            _backend=InterpBackend())
            w._load(None, 0 * g["gen_t1"] if model == "CCD" else g["gen_t1"], g["gen_t2"])     # the object's own amplitudes
            hb = pycc_b200.cchbar(w)
            lam = pycc_b200.cclambda(w, hb)
            l1 = np.zeros_like(z["l1"]) if model == "CCD" else z["l1"]
            for ftag, F in (("field", z["F_field"]), ("complex", z["F_complex"])):
                r1, r2 = w.residuals(F, t1, z["t2"], real_time=True)
                out["d1_" + ftag] = float(np.abs(r1.numpy() - z["%s_%s_r1" % (tag, ftag)]).max())
                out["d2_" + ftag] = float(np.abs(r2.numpy() - z["%s_%s_r2" % (tag, ftag)]).max())
                # the Lambda residuals of the same call (rtcc.f, pycc/rt/rtcc.py:132): HBAR rebuilt from the complex
                # amplitudes with HvT / Hs sharded by rows, one allreduce per residual
                q1, q2 = lam.residuals(F, t1, z["t2"], l1, z["l2"])
                out["dl1_" + ftag] = float(np.abs(q1.numpy() - z["%s_%s_rl1" % (tag, ftag)]).max())
                out["dl2_" + ftag] = float(np.abs(q2.numpy() - z["%s_%s_rl2" % (tag, ftag)]).max())
        q.put(out)
    except Exception:
        import traceback
        q.put(dict(rank=rank, error=traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("model", ["CCSD", "CCD", "CC2"])
@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8"])
def test_two_ranks_gloo_complex(name, model):
    """Complex-amplitude residuals and Lambda residuals (the real-time propagator's two calls) on two ranks: the
    plane-expanded statements in their sliced form, partial planes summed by one allreduce per residual."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_complex_worker, args=(r, world, port, name, model, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for d in res:
        assert "error" not in d, d.get("error")
        for k in ("d1_field", "d2_field", "d1_complex", "d2_complex", "dl1_field", "dl2_field", "dl1_complex", "dl2_complex"):
            assert d[k] < 1e-12, (k, d[k])


def _complex_cc3_worker(rank, world, port, name, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.conftest import Golden, GOLDEN_DIR
        from tests.interp_backend import InterpBackend
        import pycc_b200
        from pycc_b200.parallel import virtual_slice
        g = Golden(name)
        z = np.load(os.path.join(GOLDEN_DIR, "rtcc3_complex", name + ".npz"))
        lo, hi = virtual_slice(g.nv, rank, world)
        blocks = dict(g.blocks)
        blocks["vvvv"] = np.ascontiguousarray(blocks["vvvv"][lo:hi])
        out = dict(rank=rank)
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, blocks, g.no, model="CC3", group="world", _backend=InterpBackend())
            for call in range(2):            # twice: the gathered rows of the first call must not leak into the second
                for tag in ("field", "mixed"):
                    r1, r2 = w.residuals(z["F_" + tag], z["t1"], z["t2"], real_time=True)
                    out["d1_%s_%d" % (tag, call)] = float(np.abs(r1.numpy() - z[tag + "_r1"]).max())
                    out["d2_%s_%d" % (tag, call)] = float(np.abs(r2.numpy() - z[tag + "_r2"]).max())
            # the Lambda residuals of a real-time wave function: pairs of both triples loops dealt over the ranks, the
            # dressed <ab|ef> by rows
            zl = np.load(os.path.join(GOLDEN_DIR, "rtcc3_complex", name + "_lambda.npz"))
            w.real_time = True
            hb = pycc_b200.cchbar(w)
            lam = pycc_b200.cclambda(w, hb)
            for call in range(2):
                for tag in ("field", "mixed"):
                    q1, q2 = lam.residuals(zl["F_" + tag], zl["t1"], zl["t2"], zl["l1"], zl["l2"])
                    out["dl1_%s_%d" % (tag, call)] = float(np.abs(q1.numpy() - zl[tag + "_rt_r1"]).max())
                    out["dl2_%s_%d" % (tag, call)] = float(np.abs(q2.numpy() - zl[tag + "_rt_r2"]).max())
        q.put(out)
    except Exception:
        import traceback
        q.put(dict(rank=rank, error=traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8"])
def test_two_ranks_gloo_complex_cc3(name):
    """CC3 residuals with complex amplitudes on two ranks: occupied pairs dealt over the ranks, the v^4 term of the
    dressed <ab|ei> from each rank's rows of <ab|ef> (one allreduce), partial planes summed with the builders'."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_complex_cc3_worker, args=(r, world, port, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for d in res:
        assert "error" not in d, d.get("error")
        for k, val in d.items():
            if k.startswith("d"):
                assert val < 1e-12, (k, val)


def _t3density_worker(rank, world, port, name, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.interp_backend import InterpBackend
        from tests import t3density_checks as tc
        import pycc_b200
        g, z = tc.load(name)
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CCSD(T)", group="world", full_ovvv=True,
                                            _backend=InterpBackend())
            w._load(None, g["sym_t1"], g["sym_t2"])
            et = w.t3_density()
            out = dict(rank=rank, d_et=abs(et - float(z["sym_ET"])))
            for f in tc.FIELDS:
                out["d_" + f] = float(np.abs(getattr(w, f).numpy() - z["sym_" + f]).max())
            # the chain solve_cc(make_t3_density) -> cchbar -> solve_lambda on two ranks
            w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CCSD(T)", group="world", full_ovvv=True,
                                            make_t3_density=True, _backend=InterpBackend())
            ecc = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
            hb = pycc_b200.cchbar(w)
            lam = pycc_b200.cclambda(w, hb)
            lecc = lam.solve_lambda(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        out.update(d_ecc=abs(ecc - float(z["solve_ecc_total"])), d_lecc=abs(lecc - float(z["solve_lecc"])),
                   niter=len(lam.iteration_log), niter_ref=int(z["solve_niter"]),
                   d_l2=float(np.abs(lam.l2.numpy() - z["solve_l2"]).max()))
        q.put(out)
    except Exception:
        import traceback
        q.put(dict(rank=rank, error=traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8"])
def test_two_ranks_gloo_t3_density(name):
    """ccwfn.t3_density with the occupied pairs dealt over two ranks (one allreduce per accumulator), and the
    CCSD(T) Lambda chain on top of the sharded HBAR / Lambda engines."""
    from tests import t3density_checks as tc
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_t3density_worker, args=(r, world, port, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for d in res:
        assert "error" not in d, d.get("error")
        assert d["d_et"] < 1e-12
        for f in tc.FIELDS:
            assert d["d_" + f] < 1e-12, (f, d["d_" + f])
        assert d["d_ecc"] < 1e-10 and d["d_lecc"] < 1e-10 and d["niter"] == d["niter_ref"] and d["d_l2"] < 1e-9


def _cc3_lambda_worker(rank, world, port, name, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from tests.interp_backend import InterpBackend
        from tests import lambda_cc3_checks as l3
        import pycc_b200
        from pycc_b200.parallel import virtual_slice
        g, z = l3.load(name)
        lo, hi = virtual_slice(g.nv, rank, world)
        blocks = dict(g.blocks)
        blocks["vvvv"] = np.ascontiguousarray(blocks["vvvv"][lo:hi])
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, blocks, g.no, model="CC3", group="world", _backend=InterpBackend())
            w._load(None, z["t1"], z["t2"])
            hb = pycc_b200.cchbar(w)
            lam = pycc_b200.cclambda(w, hb)
            lam._load(z["l1"], z["l2"])
            r1, r2 = lam.engine.residuals(symmetrize=True)
            out = dict(rank=rank, d_r1=float(np.abs(r1.numpy() - z["r1"]).max()),
                       d_r2=float(np.abs(r2.numpy() - z["r2"]).max()), rows=int(lam.engine.cc3.Wab.shape[0]))
            w = pycc_b200.ccwfn.from_arrays(g.F, blocks, g.no, model="CC3", group="world", _backend=InterpBackend())
            ecc = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
            hb = pycc_b200.cchbar(w)
            lam = pycc_b200.cclambda(w, hb)
            lecc = lam.solve_lambda(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        out.update(d_ecc=abs(ecc - float(z["solve_ecc"])), d_lecc=abs(lecc - float(z["solve_lecc"])),
                   niter=len(lam.iteration_log), niter_ref=int(z["solve_niter"]),
                   d_l2=float(np.abs(lam.l2.numpy() - z["solve_l2"]).max()))
        q.put(out)
    except Exception:
        import traceback
        q.put(dict(rank=rank, error=traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8"])
def test_two_ranks_gloo_cc3_lambda(name):
    """The CC3 Lambda equations on two ranks: pairs of the t3 / l3 loops dealt over the ranks, the dressed <ab|ef> held
    by rows, Z intermediates and Y2 completed by allreduces."""
    from tests.conftest import Golden
    g = Golden(name)
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_cc3_lambda_worker, args=(r, world, port, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=900) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for d in res:
        assert "error" not in d, d.get("error")
    assert sum(d["rows"] for d in res) == g.nv
    for d in res:
        assert d["d_r1"] < 1e-11 and d["d_r2"] < 1e-11
        assert d["d_ecc"] < 1e-10 and d["d_lecc"] < 1e-10 and d["niter"] == d["niter_ref"] and d["d_l2"] < 1e-9


def test_partition_helpers():
    from pycc_b200.parallel import virtual_slice, deal_pairs
    for nv in (1, 7, 19, 150, 250, 400):
        for world in (1, 2, 4, 8):
            edges = [virtual_slice(nv, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == nv
            for a, b in zip(edges, edges[1:]):
                assert a[1] == b[0]
            if nv % 2 == 0 and nv // 2 >= world:
                assert all(lo % 2 == 0 for lo, _ in edges)
    from pycc_b200.parallel import flat_chunk
    for n in (1, 2, 7, 1000, 9003000, 257123457):
        for world in (1, 2, 3, 8):
            cuts = [flat_chunk(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n and len({c[2] for c in cuts}) == 1
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            assert all(hi - lo <= ch for lo, hi, ch in cuts) and cuts[0][2] * world >= n
            if world > 1:
                assert cuts[0][2] % 2 == 0
    for no in (1, 4, 20, 30):
        for world in (1, 2, 8):
            allp = sorted(p for r in range(world) for p in deal_pairs(no, r, world))
            assert allp == [(i, j) for i in range(no) for j in range(i + 1)]
            loads = [sum(j + 1 for _, j in deal_pairs(no, r, world)) for r in range(world)]
            assert max(loads) - min(loads) <= no
