"""TEST INFRASTRUCTURE ONLY -- import the *verbatim* reference modules of the hot path.

The reference (oscarkey/safe-exploration @ 32f0582a, constrained-cem-mpc @ 4e207630) is pure Python but
depends on packages that are absent from this image (gpytorch 0.3.2, polytope, casadi, hessian, matplotlib,
...) and on `torch.eig`, which torch >= 1.13 removed.  This module installs *name-only* stand-ins for those
imports so that the reference's own hot-path files execute unmodified from `/root/reference`:

    constrained_cem_mpc.constrained_cem_mpc        (CEM loop, constraints)
    safe_exploration.gp_reachability_pytorch       (batched one-step reachability, polytope check)
    safe_exploration.gp_reachability               (NumPy twin incl. multistep_reachability)
    safe_exploration.utils / utils_ellipsoid       (remainder bound, Minkowski sums)
    safe_exploration.ssm_cem.ssm_cem               (CemSSM ABC)

It exists to (1) validate `oracle/oracle.py` against the real reference code and (2) generate the golden
vectors under `tests/golden/` (script: `tests/golden/make_golden.py`).  `/root/reference` does not exist on
the GPU box, so nothing in the `-m gpu` tests, `smoke()` or `bench.py` may import this module.

Nothing here is part of the product; the product never imports `oracle/`.
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

REFERENCE_ROOT = os.environ.get("SE_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "safe_exploration"))


class _Anything:
    """Attribute sink used for names the hot path imports but never calls."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, item):
        return _Anything()


def _stub_module(name: str, **attrs) -> types.ModuleType:
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    mod.__path__ = []  # behave like a package so `import stub.sub` resolves through sys.modules
    mod.__all__ = []  # `from stub import *` imports nothing

    def _getattr(item):
        if item.startswith("__"):
            raise AttributeError(item)
        return _Anything

    mod.__getattr__ = _getattr  # type: ignore[attr-defined]
    sys.modules[name] = mod
    return mod


class _Polytope:
    """Minimal stand-in for `polytope.Polytope`: only A, b, dim, chebXc are read on the CEM path
    (constrained_cem_mpc.py:19-23)."""

    def __init__(self, A, b):
        self.A = np.asarray(A, dtype=float)
        self.b = np.asarray(b, dtype=float).reshape(-1)
        self.dim = self.A.shape[1]

    @property
    def chebXc(self):
        # Only used by TerminalConstraint's distance penalty; the box centre is exact for box2poly boxes.
        lo = np.full(self.dim, np.nan)
        hi = np.full(self.dim, np.nan)
        for row, bound in zip(self.A, self.b):
            nz = np.nonzero(row)[0]
            if len(nz) == 1:
                j = nz[0]
                if row[j] > 0:
                    hi[j] = bound / row[j]
                else:
                    lo[j] = bound / row[j]
        return 0.5 * (lo + hi)


def _box2poly(box):
    """`polytope.box2poly`: A = [I; -I], b = [hi; -lo]."""
    box = np.asarray(box, dtype=float)
    n = box.shape[0]
    A = np.vstack((np.eye(n), -np.eye(n)))
    b = np.concatenate((box[:, 1], -box[:, 0]))
    return _Polytope(A, b)


def _torch_eig(x, eigenvectors=False):
    """`torch.eig` (removed in torch 1.13): returns ([n,2] (re, im), None)."""
    ev = torch.linalg.eigvals(x)
    return torch.stack((ev.real, ev.imag), dim=1).to(x.dtype), None


_installed = False


def install():
    """Install the stand-ins and put the reference roots on sys.path (idempotent)."""
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")

    for name in ("casadi", "casadi.tools", "hessian", "matplotlib", "matplotlib.pyplot", "matplotlib.axes",
                 "matplotlib.figure", "matplotlib.patches", "gpytorch", "gpytorch.likelihoods",
                 "gpytorch.kernels", "gpytorch.means", "gpytorch.mlls", "gpytorch.models",
                 "gpytorch.distributions", "gpytorch.functions", "gpytorch.lazy", "gpytorch.utils", "GPy",
                 "pygame", "bnn", "sacred", "gym"):
        if name not in sys.modules:
            _stub_module(name)
    # classes the reference subclasses at import time must be real classes
    for modname, clsnames in {
        "gpytorch.models": ("ExactGP",), "gpytorch.means": ("Mean",), "gpytorch.kernels": ("Kernel",),
        "gpytorch": ("Module",),
    }.items():
        for c in clsnames:
            setattr(sys.modules[modname], c, type(c, (torch.nn.Module,), {}))

    for name in list(sys.modules):
        if "." in name and name.split(".")[0] in ("gpytorch", "matplotlib", "casadi"):
            parent, child = name.rsplit(".", 1)
            if parent in sys.modules:
                setattr(sys.modules[parent], child, sys.modules[name])

    poly_pkg = _stub_module("polytope", Polytope=_Polytope, box2poly=_box2poly)
    poly_sub = _stub_module("polytope.polytope", Polytope=_Polytope, box2poly=_box2poly)
    poly_pkg.polytope = poly_sub

    # torch >= 1.13 keeps a `torch.eig` stub that only raises; replace it
    torch.eig = _torch_eig  # type: ignore[attr-defined]
    # constrained_cem_mpc.py:167 samples Normal(means, stds) with stds that may reach 0 (quirk Q5).
    torch.distributions.Distribution.set_default_validate_args(False)

    for p in (REFERENCE_ROOT, os.path.join(REFERENCE_ROOT, "constrained-cem-mpc")):
        if p not in sys.path:
            sys.path.insert(0, p)
    _installed = True


def load():
    """Return a namespace with the verbatim reference modules of the hot path."""
    install()
    ns = types.SimpleNamespace()
    ns.cem = importlib.import_module("constrained_cem_mpc.constrained_cem_mpc")
    ns.utils = importlib.import_module("safe_exploration.utils")
    ns.utils_ellipsoid = importlib.import_module("safe_exploration.utils_ellipsoid")
    ns.ssm_cem = importlib.import_module("safe_exploration.ssm_cem.ssm_cem")
    ns.reach_np = importlib.import_module("safe_exploration.gp_reachability")
    # gp_reachability_pytorch imports GpCemSSM (gpytorch-heavy) only for an isinstance check in an error path.
    ns.reach_pt = importlib.import_module("safe_exploration.gp_reachability_pytorch")
    return ns
