"""Param — host-side mirror of the reference's ``mutable struct Param`` (src/param.jl:33-100).

Defaults are those of src/param.json.  ``Param(par={...})`` overrides fields like the
reference's ``Param(par=Dict(:na => 500, ...))`` (param.jl:82-87); like Julia's ``setfield!``
it does not convert types: an Int field given a float raises.
"""
from . import _abi

# src/param.json values
DEFAULTS = dict(
    gamma=1.95, beta=0.95, R=1.0526315789473684, maxinterest=200.0, na=100, nD=2, a_high=50.0,
    a_low=0.0, a_lowT=0.0, ny=5, nT=25, retage=26, pension=0.0, alpha=0.35, alphaT=0.0, sigma=0.35,
    **{"lambda": 2e-06}, inc0=0.75, inc1=0.04, inc2=0.0002, cfloor=0.001, cfloor_plot=0.01, rho=0.95,
    k=100, nsims=100, initw0=5.0, initw1=20.0, rseed=25032020, delta=0.0, nu=0.0, bbar=7.7,
)
INT_FIELDS = ("na", "ny", "nT", "nD", "k", "retage", "nsims", "rseed")
_ALIASES = {"ρ": "rho", "ν": "nu", "lambda_": "lambda", "lam": "lambda"}


class Param:
    def __init__(self, par=None, **kw):
        for k, v in DEFAULTS.items():
            object.__setattr__(self, "lambda_" if k == "lambda" else k, v)
        self.quad_group = 1  # no reference field: fixes the quadrature summation order (see dcegm.h)
        over = dict(par or {})
        over.update(kw)
        for k, v in over.items():
            k = str(k).lstrip(":")
            k = _ALIASES.get(k, k)
            if k not in DEFAULTS and k != "quad_group":
                raise KeyError(f"type Param has no field {k}")  # setfield! error in the reference
            if k in INT_FIELDS or k == "quad_group":
                if isinstance(v, bool) or not isinstance(v, int):
                    raise TypeError(f"Param.{k} is an Int field; got {type(v).__name__} (param.jl:82-87 does not convert)")
            else:
                if isinstance(v, bool) or not isinstance(v, (int, float)):
                    raise TypeError(f"Param.{k} is a Float64 field; got {type(v).__name__}")
                v = float(v)
            setattr(self, "lambda_" if k == "lambda" else k, v)
        self._derive()

    def _derive(self):  # param.jl:90-93
        self.neg_gamma = (-1.0) * self.gamma
        self.oneminusgamma = 1.0 - self.gamma
        self.oneover_oneminusgamma = 1.0 / self.oneminusgamma if self.oneminusgamma != 0 else float("inf")
        self.neg_oneover_gamma = (-1.0) / self.gamma

    # Julia field names with non-ASCII identifiers
    @property
    def ρ(self):
        return self.rho

    @property
    def ν(self):
        return self.nu

    def to_c(self):
        c = _abi.dcegm_param()
        for n in _abi._DOUBLES:
            setattr(c, n, float(getattr(self, n)))
        for n in _abi._INTS:
            setattr(c, n, int(getattr(self, n)))
        c.rseed = int(self.rseed)
        return c

    def copy(self, **kw):
        d = {k: getattr(self, "lambda_" if k == "lambda" else k) for k in DEFAULTS}
        d["quad_group"] = self.quad_group
        d.update(kw)
        return Param(par=d)

    def __repr__(self):
        return "Param(" + ", ".join(f"{k}={getattr(self, 'lambda_' if k == 'lambda' else k)!r}" for k in DEFAULTS) + ")"


def params_array(params):
    """list of Param / dcegm_param -> contiguous ctypes array of dcegm_param"""
    arr = (_abi.dcegm_param * len(params))()
    for i, p in enumerate(params):
        arr[i] = p.to_c() if isinstance(p, Param) else p
    return arr
