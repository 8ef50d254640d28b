"""The storeTrBuffer pre-pass (cpu/guidedvolpathvspgintegrator.cpp:190-193, 727-728, cpu/trbuffer.h): every
depth-0 traversal of the resampling routine adds its ratio-tracking transmittance estimate to a per-pixel mean, the
buffer is written as float EXR "Transmittance.R/G/B" and read back by "loadTrBuffer" for NDS+."""
import os
import subprocess

import numpy as np
import pytest

import helpers as H
from pyvspg import _abi as A, api, pbrt_writer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cfg():
    cfg = H.small("C2", res=40, spp=6, grid_n=24)
    cfg.ip.use_nee = 0          # the buffer only depends on the camera segments; keeps the oracle pass short
    return cfg


def test_oracle_tr_buffer_is_the_mean_transmittance_of_the_primary_segment(orc):
    """Ratio tracking is an unbiased estimate of exp(-tau): the stored mean must match a brute-force optical-depth
    integration along the camera rays, and pixels whose camera ray misses the medium hold no sample."""
    cfg = _cfg()
    cfg.spp = 48
    film, cnt, tr, n = orc.render_tr(cfg.scene, cfg.guiding, cfg.ip, 0, cfg.spp, seed=2, nthreads=8)
    assert tr.shape == (cfg.height, cfg.width, 3) and n.max() >= cfg.spp and n.min() == 0
    assert ((tr >= 0) & (tr <= 1)).all()
    assert np.array_equal(tr[n == 0], np.zeros_like(tr[n == 0]))
    # gray medium: three equal channels
    assert np.array_equal(tr[..., 0], tr[..., 1]) and np.array_equal(tr[..., 0], tr[..., 2])
    # brute force through the pixel centres with the oracle's own density lookup
    med = cfg.scene.desc.media[0]
    sig_t = (med.sigma_a[0] + med.sigma_s[0]) * med.sigma_scale
    dens_grid = next(e for k, e in cfg.scene.log if k == "medium")["density"].astype(np.float64)      # [nz][ny][nx] on world [0,1]^3

    def lookup(pw):      # SampledGrid::Lookup (util/containers.h:804-818): voxel-centre convention, 0 outside
        nz, ny, nx = dens_grid.shape
        inside = ((pw >= 0) & (pw <= 1)).all(axis=1)
        ps = pw * np.array([nx, ny, nz]) - 0.5
        pi = np.floor(ps).astype(int)
        d = ps - pi
        out = np.zeros(len(pw))
        for dz in (0, 1):
            for dy in (0, 1):
                for dx in (0, 1):
                    ix, iy, iz = pi[:, 0] + dx, pi[:, 1] + dy, pi[:, 2] + dz
                    ok = (ix >= 0) & (ix < nx) & (iy >= 0) & (iy < ny) & (iz >= 0) & (iz < nz)
                    v = np.where(ok, dens_grid[np.clip(iz, 0, nz - 1), np.clip(iy, 0, ny - 1), np.clip(ix, 0, nx - 1)], 0.0)
                    out += v * np.where(dx, d[:, 0], 1 - d[:, 0]) * np.where(dy, d[:, 1], 1 - d[:, 1]) * np.where(dz, d[:, 2], 1 - d[:, 2])
        return np.where(inside, out, 0.0)

    ys, xs = np.nonzero(n >= cfg.spp)
    pick = np.random.default_rng(0).choice(len(ys), size=40, replace=False)
    err = []
    for k in pick:
        px, py = int(xs[k]), int(ys[k])
        acc = []
        for s in range(8):
            o, d, _ = orc.camera_ray(cfg.scene, px, py, s, seed=2)
            t = np.linspace(0, 6, 3001, dtype=np.float64)
            pts = o[None, :].astype(np.float64) + t[:, None] * d[None, :].astype(np.float64) + cfg.scene.eye[None, :]
            tau = np.trapezoid(lookup(pts) * sig_t, t)
            acc.append(np.exp(-tau))
        err.append(tr[py, px, 0] - np.mean(acc))
    err = np.array(err)
    assert abs(err.mean()) < 0.03 and np.abs(err).max() < 0.25, (err.mean(), np.abs(err).max())


def test_exr_channel_writer_round_trips_named_layers(tmp_path):
    rng = np.random.default_rng(1)
    ch = {f"Transmittance.{c}": rng.random((7, 11), dtype=np.float32) for c in "RGB"}
    p = str(tmp_path / "tr.exr")
    api.exr_write_channels(p, ch)
    for name, want in ch.items():
        assert np.array_equal(api.exr_read_channel(p, name), want)


def test_front_end_accepts_store_tr_buffer(tmp_path):
    cfg = H.small("C1")
    text = pbrt_writer.scene_text(cfg).replace('"string lightsampler"', '"bool storeTrBuffer" true "string trBufferFileName" "out/tr.exr" "string lightsampler"')
    hs = api.HostScene(text=text, base_dir=str(tmp_path))
    assert hs.store_tr_file == str(tmp_path / "out/tr.exr")
    hs.close()
    with pytest.raises(api.VspgError) as e:
        api.HostScene(text=pbrt_writer.scene_text(cfg).replace('"string lightsampler"', '"bool storeTrBuffer" true "string lightsampler"'))
    assert "trBufferFileName" in str(e.value)


@pytest.mark.gpu
def test_gpu_tr_store_matches_the_oracle(ctx, orc):
    cfg = _cfg()
    ctx.set_scene(cfg.scene); ctx.set_guiding(cfg.guiding); ctx.set_integrator(cfg.ip)
    ctx.tr_store_begin()
    ctx.film_clear()
    ctx.render(0, cfg.spp, seed=9)
    tr_g, n_g = ctx.tr_store_download()
    ctx.tr_store_end()
    film_o, cnt, tr_o, n_o = orc.render_tr(cfg.scene, cfg.guiding, cfg.ip, 0, cfg.spp, seed=9, nthreads=8)
    assert np.array_equal(n_g, n_o)                       # every depth-0 resampling traversal, pixel by pixel
    assert n_g.max() >= cfg.spp and (n_g == 0).any()
    # fp64 sum / count on the device, float running mean on the CPU: equal to rounding
    assert np.abs(tr_g - tr_o).max() <= 2e-6, np.abs(tr_g - tr_o).max()
    # after tr_store_end the render no longer touches the accumulators
    ctx.render(0, 1, seed=10)
    with pytest.raises(api.VspgError):
        ctx.tr_store_download()


@pytest.mark.gpu
def test_cli_store_tr_buffer_writes_the_file_load_tr_buffer_reads(tmp_path, orc):
    """vspg-pbrt with "storeTrBuffer" writes the float EXR "Transmittance.R/G/B" (TrBuffer::Store); read back through the
    loader that serves "loadTrBuffer" (tests/test_frontend.py renders NDS+ from such a file) it equals the oracle's buffer."""
    exe = os.path.join(ROOT, "vspg-pbrt-v4_b200", "vspg-pbrt")
    cfg = H.small("C2", res=32, spp=4, grid_n=16)
    cfg.ip.surface_guiding = cfg.ip.volume_guiding = cfg.ip.vsp_guiding = 0
    cfg.guiding = None
    scene = pbrt_writer.scene_text(cfg, out_dir=str(tmp_path), outfile="pre.exr", seed=4, raw_density=True)
    scene = scene.replace('"string lightsampler"', '"bool storeTrBuffer" true "string trBufferFileName" "tr.exr" "string lightsampler"')
    (tmp_path / "pre.pbrt").write_text(scene)
    r = subprocess.run([exe, "--quiet", str(tmp_path / "pre.pbrt")], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    tr = np.stack([api.exr_read_channel(str(tmp_path / "tr.exr"), f"Transmittance.{c}") for c in "RGB"], axis=-1)
    film_o, _, tr_o, n_o = orc.render_tr(cfg.scene, None, cfg.ip, 0, cfg.spp, seed=4, nthreads=8)
    assert np.abs(tr - tr_o).max() <= 2e-6
    assert (n_o > 0).mean() > 0.2
