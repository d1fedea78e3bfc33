import numpy as np, torch, sys
sys.path.insert(0,'.')
from gapflow_b200.fv import FVSolver
from tests.cases import build_case, inputs
name = sys.argv[1] if len(sys.argv)>1 else 'journal1d_readme'
d, extra = inputs(name)
o = build_case(name)
s = FVSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=o.topo, extra=extra)
s.set_q(o.q); s.pre_run(False)
s.step(1)
torch.cuda.synchronize()
q = s.get_q(); o.update()
print('equal', np.array_equal(q,o.q), np.abs(q-o.q).max())
