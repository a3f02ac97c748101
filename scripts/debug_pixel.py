"""debug helper: a few pixel stacks through the GPU path and the oracle"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import siril_b200 as sb
from oracle.oracle import Oracle
orc = Oracle("port")
for rej in [int(x) for x in sys.argv[1].split(",")]:
    for N, W in [(8, 1), (12, 1), (12, 40), (12, 300)]:
        rng = np.random.default_rng(3)
        fr = (0.1 + 0.01 * rng.standard_normal((N, 1, 1, W))).astype(np.float32); fr[2] += 0.4
        a = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=torch.from_numpy(fr).cuda(), type_of_rejection=rej, sig=(3.0, 3.0), create_rejmaps=True)
        rv = sb.main_stack(a); torch.cuda.synchronize()
        o = orc.stack(fr, rejection=rej, sig=(3.0, 3.0), create_rejmaps=True)
        gl = a.result.rejmap_high.view(torch.int16).cpu().numpy()
        print(f"rej={rej} N={N} W={W} rv={rv} gpu irej={a.result.irej[0].tolist()} maps_sum={int(gl.sum())} | orc irej={o['irej'][0].tolist()} maps_sum={int(o['rej_high'].sum())} out_equal={np.array_equal(a.result.data.cpu().numpy(), o['out'])}")
