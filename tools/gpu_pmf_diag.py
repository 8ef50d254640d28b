import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
from pyvspg import api
import orc
GOLD = os.path.join(ROOT, "tests", "golden", "pbrt")
ctx = api.Context(0)
np.set_printoptions(precision=9, linewidth=200)
for name in sys.argv[1:]:
    hs = api.HostScene(path=os.path.join(GOLD, name + ".pbrt")); g = np.load(os.path.join(GOLD, name + ".npz"))
    ctx.set_scene(hs); ctx.set_integrator(hs.ip)
    c9, light = g["li_ctx"], g["li_light"]
    pick, pmf, pmf_of = ctx.debug_light_sample(c9[:, :3], c9[:, 6:9], g["ls_u"], light)
    op, opmf, opmf_of = orc.light_sample(hs, hs.ip, c9[:, :3], c9[:, 6:9], g["ls_u"], light)
    bad = np.where(pmf_of.view(np.uint32) != g["ls_pmf_of"].view(np.uint32))[0]
    print(name, "pmf_of mismatches", len(bad), "of", len(pmf_of), "; oracle-vs-golden mismatches", (opmf_of.view(np.uint32) != g["ls_pmf_of"].view(np.uint32)).sum())
    for i in bad[:6]: print("   ", i, c9[i], light[i], "gpu", pmf_of[i], "ref", g["ls_pmf_of"][i], "orc", opmf_of[i])
