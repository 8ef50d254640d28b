"""Golden for the storeISGBuffer pre-pass (tests/test_isg_store.py): the reference renderer itself (oracle/_ref/pbrt, built
by oracle/build_ref_pbrt.sh from /root/reference; its OIDN is the identity stub, oracle/stubs/OpenImageDenoise) renders the
guided scene of tests/ref_scenes.py "c2" WITHOUT a loaded ISG buffer and with "bool storeISGBuffer" true: the image-space
buffer is then trained on the fly (Update() after waves 1, 2, 4) and stored by the integrator's destructor.  The layers
the test compares are copied into tests/golden/isg_train_<criterion>.npz.

    python tests/golden/make_golden_isg.py        (needs /root/reference built into oracle/_ref and oracle/libportm.so)"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "vspg-pbrt-v4_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from pyvspg import api, pbrt_writer  # noqa: E402
import ref_scenes  # noqa: E402

SPP = 5   # waves 1..5: updates after 1, 2, 4; wave 5 only reaches the records


def isg_config(variance):
    cfg, files = ref_scenes._guided("isg", light_sampler=2, vsp_criterion_variance=1 if variance else 0)
    cfg.spp = SPP
    return cfg, files


def main():
    shim = os.path.join(ROOT, "oracle", "libportm.so")
    pbrt = os.path.join(ROOT, "oracle", "_ref", "pbrt")
    for variance in (1, 0):
        cfg, files = isg_config(variance)
        with tempfile.TemporaryDirectory() as tmp:
            sub = os.path.join(tmp, "scene"); os.makedirs(sub)
            for f in ("field_small.pgl",):
                os.symlink(os.path.join(HERE, f), os.path.join(tmp, f))
            path = pbrt_writer.write_scene(cfg, os.path.join(sub, "isg.pbrt"), outfile="isg.pfm", seed=0, cache_file="../field_small.pgl",
                                           store_isg_file="isg_out.exr")
            r = subprocess.run([pbrt, "--nthreads", "4", path], cwd=sub, env=dict(os.environ, LD_PRELOAD=shim), stdout=subprocess.PIPE,
                               stderr=subprocess.STDOUT, text=True)
            out = os.path.join(sub, "isg_out.exr")
            if r.returncode != 0 or not os.path.exists(out):
                raise SystemExit("reference run failed:\n" + r.stdout[-3000:])
            tag = "2nd" if variance else ""
            ch = {"vsp": "VSP.Y", "surf_spp": "surfSpp.Y", "vol_spp": "volSpp.Y"}
            data = {k: api.exr_read_channel(out, v) for k, v in ch.items()}
            for side in ("surf", "vol"):
                data[side + "_raw"] = np.stack([api.exr_read_channel(out, f"{side}Contrib{tag}Raw.{c}") for c in "RGB"], axis=-1)
            np.savez_compressed(os.path.join(HERE, f"isg_train_{'variance' if variance else 'contribution'}.npz"), **data)
            print("variance" if variance else "contribution", {k: (v.shape, float(np.nanmean(v))) for k, v in data.items()})


if __name__ == "__main__":
    main()
