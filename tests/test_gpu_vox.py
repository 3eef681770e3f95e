"""GPU parity: voxelisation (bit-exact voxel order, coords, counts and point slots) and pillar decoration
(<= 1e-5 relative, north_star's float standard) through the C ABI vs the oracle / reference golden vectors."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import voxelize as ov
from triton_client_b200 import synth

pytestmark = pytest.mark.gpu

KITTI = ([0, -39.68, -3, 69.12, 39.68, 1], [0.16, 0.16, 4])
NUSC = ([-51.2, -51.2, -5.0, 51.2, 51.2, 3.0], [0.2, 0.2, 8])
SHIPPED = ([0, -40, -3, 70.4, 40, 1], [0.05, 0.05, 0.1])


def _gpu_vox(engine, pts, rng, vs, p, mv, **kw):
    v, c, n, m = engine.voxelize(torch.from_numpy(pts).cuda(), rng, vs, p, mv, **kw)
    torch.cuda.synchronize()
    m = int(m.item())
    return v[:m].cpu().numpy(), c[:m].cpu().numpy(), n[:m].cpu().numpy()


def _check(got, want, batch_col=True):
    v, c, n = got
    wv, wc, wn = want
    assert v.shape == wv.shape, (v.shape, wv.shape)
    assert np.array_equal(c[:, 1:] if batch_col else c, wc) and (not batch_col or (c[:, 0] == 0).all())
    assert np.array_equal(n, wn)
    assert np.array_equal(v, wv)


@pytest.mark.parametrize("cfg,p,mv,overflow", [
    (KITTI, 32, 12000, "continue"), (KITTI, 32, 40000, "continue"), (KITTI, 32, 12000, "break"),
    (SHIPPED, 5, 40000, "continue"), (KITTI, 3, 500, "break"), (KITTI, 40, 100000, "continue"),
])
def test_c4_bit_exact(engine, cfg, p, mv, overflow):
    pts = synth.lidar_sweep(synth.seed_for(4, 1))
    want = ov.points_to_voxel(pts, cfg[0], cfg[1], p, mv, overflow)
    _check(_gpu_vox(engine, pts, cfg[0], cfg[1], p, mv, overflow=overflow), want)


def test_c5_nuscenes_break_mode_f32_and_f64(engine):
    pts = synth.lidar_multisweep(synth.seed_for(5, 1))
    for f64 in (False, True):
        want = ov.points_to_voxel(pts, NUSC[0], NUSC[1], 20, 60000, "break", f64_math=f64)
        _check(_gpu_vox(engine, pts, NUSC[0], NUSC[1], 20, 60000, overflow="break", f64_math=f64), want)
    want = ov.points_to_voxel(pts, NUSC[0], NUSC[1], 20, 5000, "break")   # overflow actually hit
    _check(_gpu_vox(engine, pts, NUSC[0], NUSC[1], 20, 5000, overflow="break"), want)


def test_range_mask_conditioning_and_strided_rows(engine):
    raw = synth.lidar_sweep(77, beams=32, azimuths=600)
    raw[:, 3] *= 37.0
    wide = np.concatenate([raw, np.full((len(raw), 2), 9.0, np.float32)], axis=1)  # 6 columns, 4 used
    cond = ov.condition_points(raw, 1.5)
    d = ov.filter_pc_pointpillar(cond, *KITTI, 32, 40000)
    v, c, n, m = engine.voxelize(torch.from_numpy(wide).cuda(), KITTI[0], KITTI[1], 32, 40000, range_mask=True, z_offset=1.5,
                                 intensity_col=3, nfeat=4)
    m = int(m.item())
    assert np.array_equal(c[:m].cpu().numpy(), d["voxel_coords"]) and np.array_equal(n[:m].cpu().numpy(), d["voxel_num_points"])
    assert np.array_equal(v[:m].cpu().numpy(), d["voxels"])


def test_dense_pillars_more_points_than_slots(engine):
    rng = np.random.default_rng(3)
    pts = np.concatenate([rng.uniform([10, 0, -1, 0], [10.15, 0.15, 0, 1], (5000, 4)),      # one pillar with 5000 points
                          rng.uniform([20, -5, -1, 0], [22, 5, 0, 1], (3000, 4))]).astype(np.float32)
    pts = pts[rng.permutation(len(pts))]
    want = ov.points_to_voxel(pts, *KITTI, 32, 12000)
    assert want[2].max() == 32
    _check(_gpu_vox(engine, pts, *KITTI, 32, 12000), want)


def test_empty_and_all_outside(engine):
    got = _gpu_vox(engine, np.zeros((0, 4), np.float32), *KITTI, 32, 100)
    assert got[0].shape == (0, 32, 4)
    far = np.full((100, 4), 1e4, np.float32)
    assert _gpu_vox(engine, far, *KITTI, 32, 100)[0].shape[0] == 0
    nan = np.full((10, 4), np.nan, np.float32)
    assert _gpu_vox(engine, nan, *KITTI, 32, 100)[0].shape[0] == 0


def test_golden_pillar_decoration(engine):
    g = golden("pillar_vfe")
    for k in range(2):
        rng, vs, p = g[f"range{k}"].tolist(), g[f"vsize{k}"].tolist(), int(g[f"maxpts{k}"])
        v, c, n, m = engine.voxelize(torch.from_numpy(g[f"points{k}"]).cuda(), rng, vs, p, 400)
        feats = engine.pillar_decorate(v, c, n, vs, rng, m=m)
        mm = int(m.item())
        want = g[f"feats{k}"]
        assert mm == want.shape[0]
        np.testing.assert_allclose(feats[:mm].cpu().numpy(), want, rtol=1e-5, atol=2e-5)


def test_decorate_full_size_vs_oracle(engine):
    pts = synth.lidar_sweep(synth.seed_for(4, 2))
    v, c, n, m = engine.voxelize(torch.from_numpy(pts).cuda(), *KITTI, 32, 12000)
    feats = engine.pillar_decorate(v, c, n, KITTI[1], KITTI[0], m=m)
    mm = int(m.item())
    want = ov.pillar_decorate(v[:mm].cpu().numpy(), c[:mm].cpu().numpy(), n[:mm].cpu().numpy(), KITTI[1], np.array(KITTI[0], np.float32))
    np.testing.assert_allclose(feats[:mm].cpu().numpy(), want, rtol=1e-5, atol=2e-5)
    # property: padded slots are exactly zero, kept slots carry the raw features unchanged
    f = feats[:mm].cpu().numpy()
    slot = np.arange(32)[None, :]
    pad = slot >= n[:mm].cpu().numpy()[:, None]
    assert (f[pad] == 0).all() and np.array_equal(f[..., :4][~pad], v[:mm].cpu().numpy()[~pad])


def test_dynamic_pillar_vfe_c5(engine):
    """b7: DynamicPillarVFE decoration on the 10-sweep nuScenes-like cloud, batch of 2."""
    a, b = synth.lidar_multisweep(51), synth.lidar_multisweep(52)[:150000]
    pts = np.concatenate([np.concatenate([np.zeros((len(a), 1), np.float32), a], 1),
                          np.concatenate([np.ones((len(b), 1), np.float32), b], 1)])
    want_f, want_inv, want_c = ov.dynamic_pillar_vfe(pts, NUSC[1], NUSC[0], [512, 512, 1])
    f, inv, c, counts = engine.dynamic_pillar(torch.from_numpy(pts).cuda(), NUSC[0], NUSC[1], batch_size=2)
    nv, nu = (int(v) for v in counts.cpu())
    assert nv == want_f.shape[0] and nu == want_c.shape[0]
    assert np.array_equal(inv[:nv].cpu().numpy(), want_inv)
    assert np.array_equal(c[:nu].cpu().numpy(), want_c)
    got = f[:nv].cpu().numpy()
    assert np.array_equal(got[:, :5], want_f[:, :5]) and np.array_equal(got[:, 8:], want_f[:, 8:])   # raw + centre offsets: exact
    np.testing.assert_allclose(got[:, 5:8], want_f[:, 5:8], rtol=1e-5, atol=2e-5)                  # cluster offsets: mean in f64 vs f32


def test_dynamic_pillar_empty(engine):
    f, inv, c, counts = engine.dynamic_pillar(torch.full((10, 6), 1e4).cuda(), NUSC[0], NUSC[1])
    assert counts.cpu().tolist() == [0, 0]


def test_dynamic_pillar_vfe_reference_golden(engine):
    """DynamicPillarVFE.forward run unmodified (golden_dynamic_vfe): inverse index and voxel coords identical, raw features
    and centre offsets identical, cluster offsets within 1e-5 (the mean is a float32 running sum there)."""
    g = golden("dynamic_vfe")
    f, inv, c, counts = engine.dynamic_pillar(torch.from_numpy(g["points"]).cuda(), g["range"].tolist(), g["vsize"].tolist(), batch_size=2)
    nv, nu = (int(v) for v in counts.cpu())
    assert nv == g["feats"].shape[0] and nu == g["voxel_coords"].shape[0]
    assert np.array_equal(inv[:nv].cpu().numpy(), g["unq_inv"]) and np.array_equal(c[:nu].cpu().numpy(), g["voxel_coords"])
    got = f[:nv].cpu().numpy()
    assert np.array_equal(got[:, :5], g["feats"][:, :5]) and np.array_equal(got[:, 8:], g["feats"][:, 8:])
    np.testing.assert_allclose(got[:, 5:8], g["feats"][:, 5:8], rtol=1e-5, atol=2e-5)


def _batch_of(sweeps):
    offs = np.concatenate([[0], np.cumsum([len(s) for s in sweeps])]).astype(int).tolist()
    return np.concatenate(sweeps).astype(np.float32), offs


@pytest.mark.parametrize("cfg,p,mv,overflow,f", [
    (KITTI, 32, 12000, "continue", 4), (KITTI, 32, 2000, "break", 4), (SHIPPED, 5, 40000, "continue", 4),
    (NUSC, 20, 60000, "break", 5), (NUSC, 20, 3000, "break", 5), (NUSC, 20, 3000, "continue", 5),
])
def test_batched_sweeps_equal_their_single_calls_and_the_oracle(engine, cfg, p, mv, overflow, f):
    """b sweeps in one call (ragged lengths, an empty sweep, one sweep entirely outside the range): every sweep's voxels,
    coords, counts and order are those of the sequential scan of that sweep alone, the batch column is its index —
    the layout of voxel_coords [b,z,y,x] (clients/preprocess/preprocess_3d.py:50)."""
    mk = (lambda seed, **kw: synth.lidar_sweep(seed, **kw)) if f == 4 else \
         (lambda seed, **kw: synth.lidar_multisweep(seed, sweeps=3, beams=kw.get("beams", 32), azimuths=kw.get("azimuths", 940)))
    sweeps = [mk(8101), mk(8102, beams=16, azimuths=700), np.zeros((0, f), np.float32), mk(8103, beams=40, azimuths=1000)[:1025],
              np.full((300, f), 1e4, np.float32), mk(8104, beams=8, azimuths=333)]
    pts, offs = _batch_of(sweeps)
    v, c, n, m = engine.voxelize(torch.from_numpy(pts).cuda(), cfg[0], cfg[1], p, mv, overflow=overflow, offsets=offs)
    torch.cuda.synchronize()
    assert v.shape == (len(sweeps), mv, p, f) and m.shape == (len(sweeps),)
    for s, sw in enumerate(sweeps):
        ms = int(m[s])
        want = ov.points_to_voxel(sw, cfg[0], cfg[1], p, mv, overflow)
        got = (v[s, :ms].cpu().numpy(), c[s, :ms].cpu().numpy(), n[s, :ms].cpu().numpy())
        assert (got[1][:, 0] == s).all()
        assert got[0].shape == want[0].shape and np.array_equal(got[1][:, 1:], want[1]) and np.array_equal(got[2], want[2])
        assert np.array_equal(got[0], want[0])
        one = _gpu_vox(engine, sw, cfg[0], cfg[1], p, mv, overflow=overflow)
        assert np.array_equal(one[0], got[0]) and np.array_equal(one[1][:, 1:], got[1][:, 1:]) and np.array_equal(one[2], got[2])
    assert int(m[2]) == 0 and int(m[4]) == 0


def test_batched_conditioning_uses_each_sweeps_own_maximum(engine):
    a, b = synth.lidar_sweep(8201, beams=16, azimuths=500), synth.lidar_sweep(8202, beams=16, azimuths=400)
    a[:, 3] *= 37.0
    b[:, 3] *= 5.0
    pts, offs = _batch_of([a, b])
    v, c, n, m = engine.voxelize(torch.from_numpy(pts).cuda(), KITTI[0], KITTI[1], 32, 40000, range_mask=True, z_offset=1.5,
                                 intensity_col=3, offsets=offs)
    for s, sw in enumerate((a, b)):
        d = ov.filter_pc_pointpillar(ov.condition_points(sw, 1.5), *KITTI, 32, 40000)
        ms = int(m[s])
        assert np.array_equal(c[s, :ms, 1:].cpu().numpy(), d["voxel_coords"][:, 1:]) and np.array_equal(v[s, :ms].cpu().numpy(), d["voxels"])


def test_batched_decoration(engine):
    sweeps = [synth.lidar_sweep(8301), synth.lidar_sweep(8302, beams=32, azimuths=900), synth.lidar_sweep(8303, beams=8, azimuths=100)]
    pts, offs = _batch_of(sweeps)
    v, c, n, m = engine.voxelize(torch.from_numpy(pts).cuda(), *KITTI, 32, 12000, offsets=offs)
    feats = engine.pillar_decorate(v, c, n, KITTI[1], KITTI[0], m=m)
    assert feats.shape == (3, 12000, 32, 10)
    for s in range(3):
        ms = int(m[s])
        want = ov.pillar_decorate(v[s, :ms].cpu().numpy(), c[s, :ms].cpu().numpy(), n[s, :ms].cpu().numpy(), KITTI[1], np.array(KITTI[0], np.float32))
        np.testing.assert_allclose(feats[s, :ms].cpu().numpy(), want, rtol=1e-5, atol=2e-5)
        one = engine.pillar_decorate(v[s].contiguous(), c[s].contiguous(), n[s].contiguous(), KITTI[1], KITTI[0], m=m[s:s + 1].contiguous())
        assert torch.equal(one[:ms], feats[s, :ms])
    # five features / 20 slots (the nuScenes pillars): rows of 20 and 44 bytes
    cloud = synth.lidar_multisweep(8304, sweeps=2)
    v, c, n, m = engine.voxelize(torch.from_numpy(cloud).cuda(), *NUSC, 20, 30000)
    feats = engine.pillar_decorate(v, c, n, NUSC[1], NUSC[0], m=m)
    ms = int(m[0])
    want = ov.pillar_decorate(v[:ms].cpu().numpy(), c[:ms].cpu().numpy(), n[:ms].cpu().numpy(), NUSC[1], NUSC[0])
    np.testing.assert_allclose(feats[:ms].cpu().numpy(), want, rtol=1e-5, atol=2e-5)
