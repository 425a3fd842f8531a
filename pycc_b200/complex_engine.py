"""
Complex-amplitude residuals: the other caller of the hot path is
the real-time propagator, ``rtcc.f`` (pycc/rt/rtcc.py:101-143), which evaluates
``ccwfn.residuals(F, t1, t2)`` and ``cclambda.residuals(F, t1, t2, l1, l2)`` with complex
amplitudes four times per RK4 step.  FP64 tensor cores have no complex mode, so a complex tensor
is held as two real planes, X[0] = Re X, X[1] = Im X, and every statement of the real engines
(``engine.CCSDEngine``, ``hbar_engine.HbarEngine`` / ``LambdaEngine``) is expanded mechanically
into real statements over the planes, planned by the same ``dsl.plan`` onto the same kernels:

  * complex x real (integrals):  ``C[z..] += c * A[z..] * B[..]`` - the plane index z joins the
    GEMM's M (or batch) index, so e.g. the o^2 v^4 ladder reads <ab|ef> once for both planes;
  * complex x complex:           ``Cr += c (Ar Br - Ai Bi)``, ``Ci += c (Ar Bi + Ai Br)`` - two
    launches of two K-concatenated segments each;
  * a real addend goes to the real plane; '=' statements zero the output first.

The Fock matrix is treated as complex as well (rtcc adds a real field, but ``residuals`` takes
whatever it is given).  Both residual calls shard over several GPUs like the real engines (virtual
slice, HvT / Hs by rows, one allreduce per residual).
"""
import re

import torch

from . import dsl
from .dsl import TRef, plan

_NAME = re.compile(r"([A-Za-z_][A-Za-z_0-9]*)\[")


def _fmt(c):
    return repr(abs(float(c)))


def _join(target, idx, terms):
    """terms: [(coef, 'X[..]*Y[..]' or 'X[..]')] -> 'target[idx] += ...'"""
    out = []
    for n, (c, body) in enumerate(terms):
        sign = "-" if c < 0 else ("+" if n else "")
        out.append("%s %s*%s" % (sign, _fmt(c), body))
    return "%s[%s] +=%s" % (target, idx, " ".join(out))


def expand(stmt, is_complex, z="z"):
    """One statement over complex tensors -> list of actions: ('zero', name) or ('plan', text),
    where the texts use ``X`` (both planes, leading index z), ``X__r`` / ``X__i`` (one plane) and the
    unchanged names of real tensors."""
    out_name, out_idx, accumulate, terms = dsl.parse(stmt)
    if not is_complex(out_name):
        raise dsl.PlanError("output %r of %r must be a complex tensor" % (out_name, stmt))
    both, re_, im_ = [], [], []
    for coef, factors in terms:
        if len(factors) == 1:
            (d, di), = factors
            if is_complex(d):
                both.append((coef, "%s[%s%s]" % (d, z, di)))
            else:
                re_.append((coef, "%s[%s]" % (d, di)))
            continue
        (a, ai), (b, bi) = factors
        ca, cb = is_complex(a), is_complex(b)
        if ca and cb:
            re_.append((coef, "%s__r[%s]*%s__r[%s]" % (a, ai, b, bi)))
            re_.append((-coef, "%s__i[%s]*%s__i[%s]" % (a, ai, b, bi)))
            im_.append((coef, "%s__r[%s]*%s__i[%s]" % (a, ai, b, bi)))
            im_.append((coef, "%s__i[%s]*%s__r[%s]" % (a, ai, b, bi)))
        elif ca:
            both.append((coef, "%s[%s%s]*%s[%s]" % (a, z, ai, b, bi)))
        elif cb:
            both.append((coef, "%s[%s]*%s[%s%s]" % (a, ai, b, z, bi)))
        else:
            re_.append((coef, "%s[%s]*%s[%s]" % (a, ai, b, bi)))
    actions = []
    if not accumulate:
        actions.append(("zero", out_name))
    if both:
        actions.append(("plan", _join(out_name, z + out_idx, both)))
    if re_:
        actions.append(("plan", _join(out_name + "__r", out_idx, re_)))
    if im_:
        actions.append(("plan", _join(out_name + "__i", out_idx, im_)))
    return actions


class PlaneSet(object):
    """Plane buffers for the complex tensors plus the real constants, and the planner over them."""

    def __init__(self, be, dtype, dcode, vslice=None):
        self.be, self.dtype, self.dcode = be, dtype, dcode
        self.cplx = {}     # name -> tensor (2, ...)
        self.real = {}     # name -> tensor
        self.vslice = vslice   # (vlo, vl) on several GPUs: ``X_sK`` = X narrowed to the rank's virtual slice along dim K
        self.local = {}        # names that ARE the rank's rows already (vvvv_s0, ovvv_s1)

    @staticmethod
    def _split(name):
        """'t2_s3' -> ('t2', 3); 't2' -> ('t2', None)."""
        m = re.match(r"^(.*)_s([0-9])$", name)
        return (m.group(1), int(m.group(2))) if m else (name, None)

    def add_complex(self, name, shape=None, view=None):
        self.cplx[name] = view if view is not None else self.be.zeros((2,) + tuple(shape), self.dtype)
        return self.cplx[name]

    def add_real(self, name, t):
        self.real[name] = t

    def is_complex(self, name):
        if name in self.local:
            return False
        base, _ = self._split(name)
        if base in self.cplx:
            return True
        if base in self.real:
            return False
        raise dsl.PlanError("unknown tensor %r" % name)

    def _ref(self, name):
        if name in self.local:
            return TRef.of(self.local[name], name)
        plane = None
        full = name
        if name.endswith("__r") or name.endswith("__i"):
            plane, name = (0 if name.endswith("__r") else 1), name[:-3]
        base, dim = self._split(name)
        if base in self.cplx:
            t = self.cplx[base]
            off = 1                      # the plane index leads
            if plane is not None:
                t, off = t[plane], 0
        else:
            t, off = self.real[base], 0
        if dim is not None and self.vslice is not None:
            t = t.narrow(dim + off, self.vslice[0], self.vslice[1])
        return TRef.of(t, full)

    def compile(self, stmts, rename=lambda n: n):
        """Statements -> list of ('zero', tensor) / ('launch', Launch).  ``rename`` maps the
        statement's tensor names to the registered ones (drops the sharding suffixes)."""
        prog = []
        for s in stmts:
            s = _NAME.sub(lambda m: rename(m.group(1)) + "[", s)
            for kind, arg in expand(s, self.is_complex):
                if kind == "zero":
                    prog.append(("zero", self.cplx[arg]))
                else:
                    T = {n: self._ref(n) for n in set(_NAME.findall(arg))}
                    for L in plan(arg, T, self.dcode):
                        prog.append(("launch", L))
        return prog

    def run(self, prog):
        for kind, x in prog:
            if kind == "zero":
                x.zero_()
            else:
                self.be.contract(x.desc)

    def load(self, name, x):
        """Complex (or real) array / tensor -> planes."""
        dst = self.cplx[name]
        x = torch.as_tensor(x)
        if x.is_complex():
            dst[0].copy_(x.real.to(dst.device, dst.dtype))
            dst[1].copy_(x.imag.to(dst.device, dst.dtype))
        else:
            dst[0].copy_(x.to(dst.device, dst.dtype))
            dst[1].zero_()

    def fetch(self, name):
        p = self.cplx[name]
        return torch.complex(p[0], p[1])


def _strip(name):
    """'t2_s3' -> 't2' (one GPU: a sharded view is the whole tensor); 'vvvv_s0' / 'ovvv_s1' likewise."""
    m = re.match(r"^(.*)_s[0-9]$", name)
    return m.group(1) if m else name


class ComplexResiduals(object):
    """Complex CCSD residuals, and HBAR + Lambda residuals, for the models the real engines carry
    (CCSD, CCD, CC2).  Built lazily from a ``CCSDEngine`` (which supplies the ERI blocks)."""

    PREP = ["tau[ijab] = t2[ijab] + t1[ia]*t1[jb]",
            "tauh[ijab] = t2[ijab] + 0.5*t1[ia]*t1[jb]",
            "tauW[ijab] = 0.5*t2[ijab] + t1[ia]*t1[jb]"]

    def __init__(self, eng):
        self.sharded = eng.comm.world > 1
        if eng.model not in ('CCD', 'CC2', 'CCSD', 'CC3'):
            raise ValueError("complex-amplitude residuals: models CCD, CC2, CCSD, CC3 (got %r)" % (eng.model,))
        self.cc3 = None
        self.eng, self.be = eng, eng.be
        o, v = eng.no, eng.nv
        n = o + v
        # several GPUs: the statements keep their sliced form (engine.STATEMENTS: upper-case A / B = the rank's slice
        # of the virtual index, X_sK = X narrowed along K), every rank accumulates partial r1 / r2 planes and one
        # allreduce per residual sums them - the real engine's scheme, plane index included
        P = self.P = PlaneSet(eng.be, eng.dtype, eng.dcode, (eng.vlo, eng.vl) if self.sharded else None)
        if self.sharded:
            P.local["vvvv_s0"] = eng.vvvv_loc
            P.local["ovvv_s1"] = eng.ovvv_loc
        b = eng.blk
        for k in ("oooo", "ooov", "oovv", "ovov"):
            P.add_real(k, b[k])
        P.add_real("ovvv", eng.ovvv_full)
        P.add_real("vvvv", eng.vvvv_loc)
        P.add_real("Loovv", eng.Loovv)
        P.add_real("Looov", eng.Looov)
        Fc = P.add_complex("F", (n, n))
        P.add_complex("Foo", view=Fc[:, :o, :o])
        P.add_complex("Fov", view=Fc[:, :o, o:])
        P.add_complex("Fvv", view=Fc[:, o:, o:])
        vl = eng.vl       # intermediates carrying the sliced index are local-sized, as in the real engine
        shapes = dict(t1=(o, v), t2=(o, o, v, v), tau=(o, o, v, v), tauh=(o, o, v, v), tauW=(o, o, v, v),
                      t2s=(o, o, v, v), Fae=(vl, v), Fmi=(o, o), Fme=(o, v), Fae2=(vl, v), Fmi2=(o, o),
                      Wmnij=(o, o, o, o), Wr1=(o, vl, o, v), Wr2=(o, vl, o, v), Zr=(o, o, o, vl),
                      Y1=(o, o, o, vl), Y2=(o, o, vl, o), r1=(o, v), r2=(o, o, v, v))
        if eng.model == 'CC2':
            shapes.update(T11=(o, o, v, v), Yw=(o, o, o, vl), Xv=(vl, v, o, v))
        for k, s in shapes.items():
            P.add_complex(k, s)
        st = eng._statements(plain=False)
        self.prep = P.compile(self.PREP + (["T11[ijab] = t1[ia]*t1[jb]"] if eng.model == 'CC2' else []))
        self.progs = {name: P.compile(st[name], (lambda n_: n_) if self.sharded else _strip)
                      for name in eng.builder_order()}
        self._lam = None

    # ---- CCSD ---------------------------------------------------------------------------------
    def _prepare(self):
        P, be, eng = self.P, self.be, self.eng
        P.run(self.prep)
        for z in range(2):      # t2s = 2 t2 - t2^T is linear: plane by plane
            be.pair_combine(P.cplx["t2s"][z], P.cplx["t2"][z], 2.0, -1.0, None, None, 0.0, eng.no, eng.nv)

    def load(self, F, t1, t2):
        P = self.P
        P.load("F", F if F is not None else self.eng.F)
        if self.eng.model == 'CCD':
            P.cplx["t1"].zero_()
        else:
            P.load("t1", t1)
        P.load("t2", t2)

    def residuals(self, F, t1, t2, real_time=False):
        """Complex r1, r2 (pycc/ccwfn.py:331-406 evaluated with complex amplitudes)."""
        P, be, eng = self.P, self.be, self.eng
        be.bind_stream()
        self.load(F, t1, t2)
        self._prepare()
        P.cplx["r1"].zero_()
        P.cplx["r2"].zero_()
        for name in eng.builder_order():
            P.run(self.progs[name])
        if eng.model == 'CC3':       # triples contributions X1, X2 (:371-401) before r2 + r2^T
            if self.cc3 is None:
                from .complex_cc3 import ComplexCC3
                self.cc3 = ComplexCC3(self)
            self.cc3.contribute(real_time)
        if self.sharded:
            eng.comm.allreduce_(P.cplx["r1"])
            eng.comm.allreduce_(P.cplx["r2"])
        for z in range(2):
            be.pair_symmetrize(P.cplx["r2"][z], eng.no, eng.nv)
        return P.fetch("r1"), P.fetch("r2")

    # ---- HBAR + Lambda ---------------------------------------------------------------------------
    def _lambda_setup(self):
        from .hbar_engine import HbarEngine, LambdaEngine
        P, eng = self.P, self.eng
        o, v, vl = eng.no, eng.nv, eng.vl
        # several GPUs: as in hbar_engine, HvT[A,b,e,f] and Hs[E,i,a,b] (and the transposed integral copies ovA / ovB)
        # hold this rank's rows only, everything else is replicated; the statements keep their sliced form, every
        # rank accumulates its part of the rl1 / rl2 planes and one allreduce per residual sums them
        shapes = dict(Hov=(o, v), Hvv=(v, v), Hoo=(o, o), Hoooo=(o, o, o, o), HvT=(vl, v, v, v), Hvovv=(v, o, v, v),
                      Hooov=(o, o, o, v), Hovvo=(o, v, v, o), Hovov=(o, v, o, v), Hs=(vl, o, v, v), Hovoo=(o, v, o, o),
                      Xa=(o, v, o, v), Xb=(v, o, o, v), t2x=(o, o, v, v),
                      l1=(o, v), l2=(o, o, v, v), rl1=(o, v), rl2=(o, o, v, v), Goo=(o, o), Gvv=(v, v),
                      Ya=(o, v), Yb=(o, v), Fa=(v, v), Fi=(o, o))
        for k, s in shapes.items():
            P.add_complex(k, s)
        ovA, ovB = self.be.empty((o, vl, v, v), eng.dtype), self.be.empty((o, vl, v, v), eng.dtype)
        self.be.permute4(ovA, eng.ovvv_full.permute(0, 3, 2, 1).narrow(1, eng.vlo, vl))
        self.be.permute4(ovB, eng.ovvv_full.permute(0, 2, 3, 1).narrow(1, eng.vlo, vl))
        P.add_real("ovA", ovA)
        P.add_real("ovB", ovB)
        if eng.model == 'CC2':
            # the object's own (real) Fock matrix and amplitudes, which the reference's CC2 Lambda code
            # reads instead of the arguments (see hbar_engine.STATEMENTS_CC2)
            F0 = eng.F0
            P.add_real("F0oo", F0[:o, :o])
            P.add_real("F0ov", F0[:o, o:])
            P.add_real("F0vv", F0[o:, o:])
            P.add_real("t1_0", eng.t1)
            P.add_real("t2_0", eng.t2)
            self.tau_0 = self.be.empty((o, o, v, v), eng.dtype)
            P.add_real("tau_0", self.tau_0)
        hs = dict(HbarEngine.STATEMENTS)
        ls = dict(LambdaEngine.STATEMENTS)
        if eng.model == 'CC2':
            hs.update(HbarEngine.STATEMENTS_CC2)
            ls.update(LambdaEngine.STATEMENTS_CC2)
        if self.sharded:          # keep the X_sK suffixes: r1_s1 -> rl1_s1
            ren = lambda n_: re.sub(r"^r([12])(_s[0-9])?$", r"rl\1\2", n_)
            keep = lambda n_: n_
        else:
            ren = lambda n_: {"r1": "rl1", "r2": "rl2"}.get(_strip(n_), _strip(n_))
            keep = _strip
        self.hbar_prog = [P.compile(hs[name], keep) for name in HbarEngine.ORDER]
        self.lam_prog = {name: P.compile(ls[name], ren) for name in ("G", "r_L1", "r_L2")}
        self._lam = True

    def lambda_residuals(self, F, t1, t2, l1, l2, real_time=False):
        """Complex L1 / L2 residuals with HBAR rebuilt from (F, t1, t2) (pycc/cclambda.py:281-323)."""
        P, be, eng = self.P, self.be, self.eng
        if self.sharded and eng.ovvv_full is None:
            raise ValueError("complex-amplitude Lambda residuals need the full <mb|ef> block on every rank "
                             "(ccwfn(..., full_ovvv=True))")
        be.bind_stream()
        if self._lam is None:
            self._lambda_setup()
        self.load(F, t1, t2)
        self._prepare()
        for z in range(2):      # t2x[mifb] = 2 t2[mifb] - t2[imfb]
            be.permute4(P.cplx["t2x"][z], P.cplx["t2"][z], 2.0, 0.0)
            be.permute4(P.cplx["t2x"][z], P.cplx["t2"][z].permute(1, 0, 2, 3), -1.0, 1.0)
        if eng.model == 'CC2':
            be.pair_combine(self.tau_0, eng.t2, 1.0, 0.0, eng.t1, eng.t1, 1.0, eng.no, eng.nv)
        for prog in self.hbar_prog:
            P.run(prog)
        if eng.model == 'CCD':
            P.cplx["l1"].zero_()
        else:
            P.load("l1", l1)
        P.load("l2", l2)
        P.run(self.lam_prog["G"])
        P.cplx["rl1"].zero_()
        P.cplx["rl2"].zero_()
        if eng.model != 'CCD':
            P.run(self.lam_prog["r_L1"])
        P.run(self.lam_prog["r_L2"])
        if eng.model == 'CC3':       # Y1, Y2 of the triples (cclambda.py:323-440), before rl2 + rl2^T
            if getattr(self, "cc3lam", None) is None:
                from .complex_cc3 import ComplexCC3Lambda
                self.cc3lam = ComplexCC3Lambda(self)
            P.run(self.progs["Fme"])         # Fov = build_Fme(F, t1) of this call
            if self.sharded:                 # the builders' partial planes first: the triples terms are added in full
                eng.comm.allreduce_(P.cplx["rl1"])
                eng.comm.allreduce_(P.cplx["rl2"])
            self.cc3lam.contribute(real_time)
        elif self.sharded:
            eng.comm.allreduce_(P.cplx["rl1"])
            eng.comm.allreduce_(P.cplx["rl2"])
        for z in range(2):
            be.pair_symmetrize(P.cplx["rl2"][z], eng.no, eng.nv)
        return P.fetch("rl1"), P.fetch("rl2")
