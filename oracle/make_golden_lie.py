#!/usr/bin/env python
"""TEST INFRASTRUCTURE ONLY -- freezes outputs of the reference's own lietorch (dpvo/lietorch/src/lietorch_cpu.cpp + include/*.h,
built unmodified as oracle/_ref/ref_lietorch.so by build_ref.py) as tests/golden/lie_ref.npz: every m.def of lietorch.cpp:286-316,
all four groups, fp64, including the small-angle Taylor branches and un-normalised quaternion inputs.  Runs only in the authoring
container; the fixture is what tests/test_lietorch.py compares wv_lie_forward / wv_lie_backward against on the GPU box."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import build_ref  # noqa: E402

DIMS = {1: (3, 4), 2: (4, 5), 3: (6, 7), 4: (7, 8)}


def cases(g, n=24, seed=0):
    K, N = DIMS[g]
    gen = torch.Generator().manual_seed(seed + g)
    R = lambda *s: torch.randn(*s, generator=gen, dtype=torch.float64)  # noqa: E731
    a = 0.5 * R(n, K); a[0] = 0; a[1] = 1e-8; a[2] *= 1e-3; a[3] *= 4.0
    return dict(a=a, b=R(n, K), xa=0.7 * R(n, K), ya=0.7 * R(n, K), p3=R(n, 3), p4=R(n, 4), gN=R(n, N), gK=R(n, K), g3=R(n, 3), g4=R(n, 4))


def main():
    ref = build_ref.load("ref_lietorch")
    out = {}
    for g, (K, N) in DIMS.items():
        c = cases(g)
        X = ref.expm(g, c["xa"]); Y = ref.expm(g, c["ya"])
        X2 = X.clone(); qo = 0 if g < 3 else 3; X2[:, qo:qo + 4] *= 1.7      # the constructors re-normalise (so3.h:31-37)
        c.update(X=X, Y=Y, X2=X2)
        calls = [("expm", ("a",)), ("logm", ("X",)), ("logm", ("X2",)), ("inv", ("X2",)), ("mul", ("X2", "Y")), ("adj", ("X", "b")),
                 ("adjT", ("X", "b")), ("act", ("X2", "p3")), ("act4", ("X2", "p4")), ("as_matrix", ("X2",)), ("projector", ("X2",)),
                 ("Jinv", ("X", "b")), ("expm_backward", ("gN", "a")), ("logm_backward", ("gK", "X")), ("inv_backward", ("gN", "X2")),
                 ("mul_backward", ("gN", "X2", "Y")), ("adj_backward", ("gK", "X", "b")), ("adjT_backward", ("gK", "X", "b")),
                 ("act_backward", ("g3", "X2", "p3")), ("act4_backward", ("g4", "X2", "p4"))]
        for k, v in c.items():
            out[f"g{g}_in_{k}"] = v.numpy()
        for i, (name, args) in enumerate(calls):
            r = getattr(ref, name)(g, *[c[k].clone() for k in args])
            r = [r] if torch.is_tensor(r) else list(r)
            for j, t in enumerate(r):
                out[f"g{g}_call{i}_{name}__{'_'.join(args)}__out{j}"] = t.numpy()
    path = os.path.join(os.path.dirname(HERE), "tests", "golden", "lie_ref.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes", len(out), "arrays")


if __name__ == "__main__":
    main()
