"""Builds one synthetic cfg3 MSA REPS times on one context and reports how many different results came out (one, if
all is well): python tools/det_check.py [scale] [reps].  A soak / determinism check for kernels with asynchronous
copies or atomics -- it is what showed the missing proxy fence in the bulk-copy experiment of DESIGN.md section 8."""
import hashlib
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, founderblockgraphs_b200 as pkg
from founderblockgraphs_b200 import synth
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.5
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
msa, c = synth.make_config('cfg3', scale)
m, n = msa.shape
ctx = pkg.Context(0); ctx.set_options(elastic=True, gap_limit=1, ignore_chars=b'')
dev = torch.from_numpy(msa).cuda()
res = []
for r in range(reps):
    try:
        nb, score, info = ctx.run_xgfa_ptr(dev.data_ptr(), m, n, device=True)
        h = hashlib.sha256(ctx.read_xgfa(0, min(info['n_bytes'], 64 << 20))).hexdigest()[:12]
        res.append((nb, score, info['n_bytes'], h))
    except Exception as e:
        res.append(('ERR', str(e)[:60]))
print('%d builds, %d distinct results' % (reps, len(set(res))), res[:2], [x for x in res if x != res[0]][:3])
sys.exit(0 if len(set(res)) == 1 and res[0][0] != 'ERR' else 1)
