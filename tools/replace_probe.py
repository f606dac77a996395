import sys, numpy as np
sys.path.insert(0, '/root/repo')
from conjugategradient_b200 import generators as G, solver as S
ctx = S.Context(0)
v, c, r = G.poisson2d_5pt(128)
b = G.default_rhs(r.size - 1); nb = np.linalg.norm(b)
for rel in (1e-12, 1e-14, 1e-16):
    for k in (0, 10, 20, 50):
        cg = S.CG(ctx, residual_replace_every=k); cg.setMatrix(v, c, r); cg.setTarget(b)
        st = cg.solve(rel * nb, max_iter=3000)
        print(f"rel={rel:g} replace={k}: it={st['iterations']} conv={st['converged']} reported={np.sqrt(st['rr'])/nb:.3e} true={np.sqrt(cg.residual()[0])/nb:.3e}")
