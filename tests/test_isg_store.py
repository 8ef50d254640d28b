"""storeISGBuffer: the image-space VSP buffer trained on the fly (vspg.cpp:161-178, 230-258, 613-622;
openpgl/imagespace/ImageSpaceGuidingBuffer.h:632-797).  Golden: the reference renderer itself stored the buffer for the
guided scene of tests/ref_scenes.py (tests/golden/make_golden_isg.py; its OIDN is the identity stub, so the stored estimate
is the un-denoised one this path produces).  Oracle (CPU) and CUDA path (GPU, through the C ABI) must reproduce the
records and the VSP layer bit for bit: the per-pixel float running means are updated once per wave, in wave order."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden_isg as G  # noqa: E402  (scene definition shared with the generator; nothing of the reference is run here)
from pyvspg import api, synthetic  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def bits(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = a.view(np.uint32).copy()
    b[np.isnan(a)] = 0x7FC00000
    return b


def setup(variance):
    cfg, _ = G.isg_config(variance)
    hs = api.HostScene(path=None, text=__import__("pyvspg.pbrt_writer", fromlist=["x"]).scene_text(
        cfg, outfile="x.exr", seed=0, cache_file=os.path.join(GOLD, "field_small.pgl")), base_dir=GOLD)
    gold = np.load(os.path.join(GOLD, f"isg_train_{'variance' if variance else 'contribution'}.npz"))
    return cfg, hs, gold


def check(vsp, rec, gold):
    assert np.array_equal(rec[..., 0], gold["surf_spp"]) and np.array_equal(rec[..., 1], gold["vol_spp"])
    assert (gold["vol_spp"] > 0).any() and (gold["surf_spp"] > 0).any() and np.array_equal(gold["vol_spp"] + gold["surf_spp"], np.full((24, 40), float(G.SPP), np.float32))
    assert np.array_equal(bits(rec[..., 2:5]), bits(gold["surf_raw"]))
    assert np.array_equal(bits(rec[..., 5:8]), bits(gold["vol_raw"]))
    assert np.array_equal(bits(vsp), bits(gold["vsp"]))
    ok = gold["vsp"] >= 0
    assert ok.any() and (gold["vsp"][ok] <= 1).all()


@pytest.mark.parametrize("variance", [1, 0])
def test_oracle_trains_the_isg_buffer_like_the_reference(orc, variance):
    cfg, hs, gold = setup(variance)
    film, vsp, rec = orc.render_isg(hs, hs.guiding, hs.ip, G.SPP, seed=0, nthreads=4)
    check(vsp, rec, gold)
    assert film[..., 3].min() == G.SPP
    hs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("variance", [1, 0])
def test_gpu_trains_the_isg_buffer_like_the_reference(ctx, variance):
    cfg, hs, gold = setup(variance)
    ctx.set_scene(hs); ctx.set_guiding(hs.guiding); ctx.set_integrator(hs.ip)
    ctx.film_clear()
    ctx.render_training_isg(G.SPP, seed=0)
    vsp, rec = ctx.isg_store_download(40, 24)
    ctx.isg_store_end()
    check(vsp, rec, gold)
    assert ctx.film_download()[..., 3].min() == G.SPP
    hs.close()


@pytest.mark.gpu
def test_gpu_isg_store_state_rules(ctx):
    cfg, hs, gold = setup(1)
    ctx.set_scene(hs); ctx.set_guiding(hs.guiding); ctx.set_integrator(hs.ip)
    with pytest.raises(api.VspgError):
        ctx.isg_store_update()                       # not begun
    ctx.isg_store_begin()
    with pytest.raises(api.VspgError) as e:
        ctx.render(0, 2, seed=0)                     # two sample indices in one render
    assert "one sample index" in str(e.value)
    with pytest.raises(api.VspgError):
        ctx.isg_store_download(40, 24)               # no update yet -> no VSP layer
    ctx.isg_store_end()
    ctx.film_clear(); ctx.render(0, 2, seed=0)       # back to normal
    hs.close()


@pytest.mark.gpu
def test_cli_store_isg_buffer_writes_the_file_load_isg_buffer_reads(tmp_path):
    """vspg-pbrt on a scene with "storeISGBuffer": the wave loop trains the buffer and the stored EXR's VSP layer is the
    reference's; read back through the loader that serves "loadISGBuffer" it is the same array."""
    import subprocess
    from pyvspg import pbrt_writer
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "vspg-pbrt-v4_b200", "vspg-pbrt")
    cfg, _ = G.isg_config(1)
    text = pbrt_writer.scene_text(cfg, outfile="pre.exr", seed=0, cache_file=os.path.join(GOLD, "field_small.pgl"), store_isg_file="isg.exr")
    (tmp_path / "pre.pbrt").write_text(text)
    r = subprocess.run([exe, "--quiet", str(tmp_path / "pre.pbrt")], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    gold = np.load(os.path.join(GOLD, "isg_train_variance.npz"))
    vsp = api.exr_read_channel(str(tmp_path / "isg.exr"), "VSP.Y")
    assert np.array_equal(bits(vsp), bits(gold["vsp"]))
    assert np.array_equal(api.exr_read_channel(str(tmp_path / "isg.exr"), "volSpp.Y"), gold["vol_spp"])
    # second pass: the same scene now loads the buffer it stored
    text2 = pbrt_writer.scene_text(cfg, outfile="x.exr", seed=0, cache_file=os.path.join(GOLD, "field_small.pgl"), isg_file=str(tmp_path / "isg.exr"))
    hs = api.HostScene(text=text2, base_dir=str(tmp_path))
    got = np.ctypeslib.as_array(hs.guiding.desc.isg_vsp, (24, 40))
    assert np.array_equal(bits(got), bits(gold["vsp"]))
    hs.close()
