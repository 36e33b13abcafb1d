import sys, os, hashlib
sys.path.insert(0, '/root/repo')
import torch, founderblockgraphs_b200 as pkg
from founderblockgraphs_b200 import synth
scale = float(sys.argv[1]); reps = int(sys.argv[2])
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
print(os.environ.get('FBG_BS_DBG'), os.environ.get('FBG_BS_STAGED'), len(set(res)), res[:3], [x for x in res if x != res[0]][:3])
