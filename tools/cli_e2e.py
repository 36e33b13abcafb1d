"""The drop-in executable end to end (FASTA parse -> GPU path -> xGFA bytes on disk), with its
own per-stage wall clock (FBG_TIMING=1):  python tools/cli_e2e.py cfg3 0.2"""
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from founderblockgraphs_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
msa, c = synth.make_config(name, scale)
m, n = msa.shape
tmp = os.environ.get("TMPDIR", "/tmp")
fa, out = os.path.join(tmp, "e2e_%s.fasta" % name), os.path.join(tmp, "e2e_%s.xgfa" % name)
with open(fa, "wb") as fh:
    for i in range(m):
        fh.write(b">h%d\n" % i)
        fh.write(msa[i].tobytes())
        fh.write(b"\n")
exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "founderblockgraphs_b200", "founderblockgraph")
flags = ["--elastic", "--gfa"] if c["elastic"] else ["--gap-limit=1", "--nonelastic-gfa"]
if c["ignore"]:
    flags.append("--ignore-chars=" + c["ignore"].decode())
for rep in range(3):
    t = time.perf_counter()
    pr = subprocess.run([exe, "--input", fa, "--output", out] + flags, env=dict(os.environ, FBG_TIMING="1"), capture_output=True)
    dt = time.perf_counter() - t
    line = [l for l in pr.stderr.decode(errors="replace").splitlines() if l.startswith("[fbg] stages")]
    print("rep %d: exit %d, process wall %.3f s, FASTA %.1f MB, xGFA %.1f MB, %s" % (
        rep, pr.returncode, dt, os.path.getsize(fa) / 1e6, os.path.getsize(out) / 1e6 if os.path.exists(out) else 0, line[-1] if line else pr.stderr.decode(errors="replace")[-300:]))
print("cells %d" % (m * n))
