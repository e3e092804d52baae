"""Generate tests/golden/terms_v1.npz from the UNMODIFIED reference (run in the build container):

    python tests/golden/make_golden_terms.py

Imports celerite from baseline/_ref (the pip install of /root/reference).  For a
handful of term sums and random parameter vectors it stores what the reference's
`TermSum.get_all_coefficients()` and `.jitter` return, and a central-difference Jacobian of that map
(the reference's own Jacobian needs autograd, which this image does not have).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
from celerite import terms  # noqa: E402

KINDS = {"jitter": 0, "real": 1, "complex3": 2, "complex4": 3, "sho": 4, "sho_over": 5, "matern32": 6, "product": 7,
         "sum": 8}
SPECS = [
    ["sho"],
    ["sho", "jitter"],
    ["real", "sho", "sho"],
    ["sho_over", "real"],
    ["complex4", "complex3", "jitter", "real"],
    ["matern32", "sho", "jitter"],
    ["sho_over", "sho", "matern32", "complex4", "real", "jitter"],
    [("product", "real", "real"), "jitter"],
    [("product", "sho", "real"), ("product", "complex4", "sho")],
    ["real", ("product", "real", "matern32"), ("product", "complex3", "complex4"), "sho"],
    # nested products, sums inside products, factors with two coefficient sets
    [("product", "sho_over", "sho")],
    [("product", "sho_over", "sho_over"), "jitter"],
    [("product", ("product", "real", "sho"), "complex3"), "real"],
    [("product", ("sum", "real", "sho"), "sho_over"), "jitter"],
    ["sho", ("product", ("sum", "real", "real"), ("sum", "sho", ("product", "matern32", "complex4")))],
]
NPAR = {"jitter": 1, "real": 2, "complex3": 3, "complex4": 4, "sho": 3, "sho_over": 3, "matern32": 2}


def flatten(names):
    """nested spec -> prefix list of names"""
    out = []
    for n in names:
        if isinstance(n, tuple):
            out += [n[0]] + flatten([n[1]]) + flatten([n[2]])
        else:
            out.append(n)
    return out


def draw(rng, name):
    if name == "jitter":
        return [rng.uniform(-3, 0)], 0.0
    if name == "real":
        return [rng.uniform(-1, 1), rng.uniform(-3, 1)], 0.0
    if name == "complex3":
        return [rng.uniform(-1, 1), rng.uniform(-3, 1), rng.uniform(-1, 2)], 0.0
    if name == "complex4":
        la, lc, ld = rng.uniform(-1, 1), rng.uniform(-3, 1), rng.uniform(-1, 2)
        return [la, la + lc - ld - rng.uniform(0.1, 2.0), lc, ld], 0.0      # a c >= b d
    if name == "sho":
        return [rng.uniform(-1, 1), np.log(rng.uniform(0.6, 8.0)), rng.uniform(-1, 3)], 0.0
    if name == "sho_over":
        return [rng.uniform(-1, 1), np.log(rng.uniform(0.05, 0.45)), rng.uniform(-1, 3)], 0.0
    if name == "matern32":
        return [rng.uniform(-1, 1), rng.uniform(-1, 2)], rng.choice([0.01, 0.003])
    raise ValueError(name)


def one(name, q, eps):
    if name == "real":
        return terms.RealTerm(log_a=q[0], log_c=q[1])
    if name == "complex3":
        return terms.ComplexTerm(log_a=q[0], log_c=q[1], log_d=q[2])
    if name == "complex4":
        return terms.ComplexTerm(log_a=q[0], log_b=q[1], log_c=q[2], log_d=q[3])
    if name == "sho":
        return terms.SHOTerm(log_S0=q[0], log_Q=q[1], log_omega0=q[2])
    return terms.Matern32Term(log_sigma=q[0], log_rho=q[1], eps=eps)


def build(names, params, epss):
    """names: nested spec; epss: one value per FLATTENED position."""
    state = {"o": 0, "pos": 0}

    def node(name):
        if isinstance(name, tuple):
            state["pos"] += 1
            k1 = node(name[1])
            k2 = node(name[2])
            return k1 * k2 if name[0] == "product" else k1 + k2
        eps = epss[state["pos"]]
        state["pos"] += 1
        n = NPAR[name]
        q = params[state["o"]:state["o"] + n]
        state["o"] += n
        if name == "jitter":
            return terms.JitterTerm(log_sigma=q[0])
        if name in ("sho", "sho_over"):
            return terms.SHOTerm(log_S0=q[0], log_Q=q[1], log_omega0=q[2])
        return one(name, q, eps)

    ts = [node(name) for name in names]
    k = ts[0]
    for t in ts[1:]:
        k = k + t
    return k


def row_of(names, params, epss):
    k = build(names, params, epss)
    c = k.get_all_coefficients()
    return np.concatenate([[k.jitter]] + [np.atleast_1d(np.asarray(a, dtype=np.float64)) for a in c])


def main():
    rng = np.random.default_rng(20240607)
    out = {}
    for si, names in enumerate(SPECS):
        P, R, Jc = [], [], []
        epss = None
        for w in range(5):
            ps, es = [], []
            for name in flatten(names):
                if name in ("product", "sum"):
                    es.append(0.0)
                    continue
                q, e = draw(rng, name)
                ps += q
                es.append(e)
            if epss is None:
                epss = es
            ps = np.array(ps)
            row = row_of(names, ps, epss)
            jac = []
            for i in range(ps.size):
                d = np.zeros_like(ps)
                d[i] = 1e-6
                jac.append((row_of(names, ps + d, epss) - row_of(names, ps - d, epss)) / 2e-6)
            P.append(ps); R.append(row); Jc.append(np.array(jac))
        out["spec%d_kinds" % si] = np.array([KINDS[n] for n in flatten(names)], dtype=np.int32)
        out["spec%d_eps" % si] = np.array([e if e else 0.01 for e in epss])
        out["spec%d_params" % si] = np.array(P)
        out["spec%d_rows" % si] = np.array(R)
        out["spec%d_jac" % si] = np.array(Jc)
    out["nspec"] = np.array(len(SPECS))
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "terms_v1.npz"), **out)
    print("wrote terms_v1.npz with", len(SPECS), "term sums")


if __name__ == "__main__":
    main()
