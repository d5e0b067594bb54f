"""GPU parity of the Gaussian-surfel rasterizer (through the C-ABI) against the UNMODIFIED reference
CUDA rasterizer built into oracle/_ref/libref_rs.so
(submodules/diff-gaussian-surfel-rasterization-w-pose).

Bars: radii / tile lists / ranges / n_contrib bit-exact; images (RGB, normal,
depth, opac) rtol 1e-4 atol 1e-5; `contributions` (float atomics in both implementations) rtol 1e-4;
gradients within 1e-3 of the per-tensor max.
"""
import pytest
import torch

from pings_b200 import synth

pytestmark = pytest.mark.gpu

IMG_RTOL, IMG_ATOL = 1e-4, 1e-5
GRAD_REL_TO_MAX = 1e-3


def _scene(P, W, H, seed):
    if (W, H) == (1920, 1080):   # BASELINE.json configs[2] (the bench.py C3 scene)
        cam = synth.hd_camera(seed, W, H)
        sc = {k: v.cuda() for k, v in synth.surfels_in_view(cam, P, seed).items()}
    else:
        cam = synth.kitti_camera(seed, W, H)
        sc = {k: v.cuda() for k, v in synth.gaussians_in_view(cam, P, seed, surfel=True).items()}
    mats = [m.cuda() for m in cam.matrices()]
    ct = {k: v.cuda() for k, v in synth.cotangents(cam, seed, normal=True).items()}
    extra = dict(prcppoint=torch.tensor([cam.cx / W, cam.cy / H], device="cuda"),
                 patch_bbox=torch.tensor([0.0, 0.0, H - 1.0, W - 1.0], device="cuda"))
    return cam, sc, mats, ct, extra


def _relmax(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return ((a - b).abs().max() / (b.abs().max() + 1e-30)).item()


CONFIGS = [
    [1, 1, 1, 1, 1],  # PINGS default (GR:137-142): surface, normalised depth, per-pixel depth, front only
    [1, 0, 1, 1, 0],  # un-normalised depth (D + 10 T), back faces kept
    [1, 1, 0, 1, 1],  # no per-pixel depth
    [0, 1, 0, 1, 0],  # not a surface: plain 3D Gaussians through the surfel pipeline
]


@pytest.mark.parametrize("P,W,H,seed,cfg", [(3000, 320, 200, 3, 0), (40000, 640, 360, 1, 0), (20000, 333, 217, 5, 1),
                                             (20000, 320, 200, 6, 2), (10000, 320, 200, 8, 3),
                                             (1000000, 1920, 1080, 1, 0)])   # C3 at full size
def test_forward_bitexact_and_backward(P, W, H, seed, cfg):
    from oracle.ref_api import RefRasterS, available
    from pings_b200 import _raster_c as RC
    if not available("ref_rs"):
        pytest.fail("oracle/_ref/libref_rs.so missing: run __graft_entry__.build() where /root/reference exists")
    cam, sc, mats, ct, ex = _scene(P, W, H, seed)
    view, full, proj, center = mats
    config = torch.tensor(CONFIGS[cfg], dtype=torch.float32, device="cuda")
    bg = torch.tensor([1.0, 0.5, 0.25], device="cuda")
    e = torch.empty(0)
    res = RC.rasterize_gaussians_surfel(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"],
                                        sc["rotations"], 1.0, e, view, full, proj, ex["prcppoint"], ex["patch_bbox"],
                                        cam.tanfovx, cam.tanfovy, H, W, e, 0, center, False, True, config)
    R, color, normal, depth, opac, radii, contributions, geom, binning, img = res
    ref = RefRasterS()
    r = ref.forward(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, bg, ex["prcppoint"], ex["patch_bbox"],
                    config, campos=center)
    assert R == r["num_rendered"]
    assert torch.equal(radii, r["radii"])
    assert int((radii > 0).sum()) > P // 20
    dv = RC.debug_views(1, P, W, H, R, geom, binning, img)
    vis = radii > 0
    assert torch.equal(dv["records"][vis][:, 2, 0].contiguous().view(torch.int32), r["depths"][vis].view(torch.int32))
    assert torch.equal(dv["records"][vis][:, 0, :2].contiguous().view(torch.int32),
                       r["means2D"][vis].contiguous().view(torch.int32))
    assert torch.equal(dv["records"][vis][:, 1, :].contiguous().view(torch.int32),
                       r["conic_opacity"][vis].contiguous().view(torch.int32))
    assert torch.equal(dv["point_list"], r["point_list"])
    assert torch.equal(dv["ranges"], r["ranges"])
    assert torch.equal(dv["n_contrib"], r["n_contrib"])
    torch.testing.assert_close(color, r["color"], rtol=IMG_RTOL, atol=IMG_ATOL)
    torch.testing.assert_close(normal, r["normal"], rtol=IMG_RTOL, atol=IMG_ATOL)
    torch.testing.assert_close(depth, r["depth"], rtol=IMG_RTOL, atol=2e-5)
    torch.testing.assert_close(opac, r["opac"], rtol=IMG_RTOL, atol=IMG_ATOL)
    torch.testing.assert_close(dv["final_T"], r["final_T"], rtol=IMG_RTOL, atol=IMG_ATOL)
    torch.testing.assert_close(contributions, r["contributions"], rtol=1e-4, atol=1e-5)

    out = RC.rasterize_gaussians_surfel_backward(
        bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, e, view, full, proj,
        ex["prcppoint"], ex["patch_bbox"], cam.tanfovx, cam.tanfovy, ct["color"], ct["normal"], ct["depth"], ct["opac"],
        e, 0, center, geom, R, binning, img, True, config, full_tau=True)
    g = ref.backward(sc, (view, full, proj), cam.tanfovx, cam.tanfovy, bg, ex["prcppoint"], ex["patch_bbox"], config,
                     r["radii"], ct["color"], ct["normal"], ct["depth"], ct["opac"], campos=center)
    # The reference accumulates with per-thread float atomics (RS backward.cu:767-911), so its own
    # gradients vary run to run.  The blend-level sums (means2D, colors, opacity) are well conditioned and
    # must agree to 1e-3 of the tensor max on EVERY row.  The per-surfel chain behind them (conic ->
    # cov2D -> cov3D -> scale/rotation, mean, pose) cancels catastrophically in fp32 (a*c - b*b of the
    # projected covariance) for thin surfels seen edge-on, in the reference exactly as here: both are the same
    # fp32 formulas.  Who is right is decided by the fp64 arbiter (oracle/cpu_rasters64.c: blend + chain in
    # double on the same tile lists, itself pinned by finite differences in tests/test_oracle_rasters64.py):
    #   (a) per tensor, max-norm: |ours - fp64| <= 1.5 x max over 6 reference runs of |ref - fp64| (+3e-4)
    #   (b) chain alone (each side's own fp32 blend sums through the fp64 chain; a single-row max statistic of
    #       pure fp32 rounding, so looser): <= 4 x the reference's + 1e-3
    #   (c) rows with kappa = ac/(ac-b^2) < 10 (85 % of the rows): 99 % of them within 1e-3 of the fp64 row
    #   (d) summed pose gradient: |ours - fp64| <= max(1.5 x worst reference run, 5e-4)
    # At the full C3 size the fp64 blend would take minutes on the CPU, so (b) and (c) run against the fp64
    # chain of our own blend sums and the reference comparison keeps the run-to-run-spread bar.
    runs = [g] + [ref.backward(sc, (view, full, proj), cam.tanfovx, cam.tanfovy, bg, ex["prcppoint"], ex["patch_bbox"],
                               config, r["radii"], ct["color"], ct["normal"], ct["depth"], ct["opac"], campos=center)
                  for _ in range(5)]
    names = ["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations", "tau"]
    ours = dict(zip(names, out))
    nvis = int((radii > 0).sum())
    for n in ("means2D", "colors", "opacity"):
        refs0 = runs[0][n].double().reshape(P, -1)
        mx = refs0.abs().max() + 1e-30
        err = (ours[n].double().reshape(P, -1) - refs0).abs().max(1).values
        assert float(err.max() / mx) < GRAD_REL_TO_MAX, (n, float(err.max() / mx))
    from oracle import cpu_rasters64 as O64
    chain_names = ("means3D", "cov3D", "scales", "rotations", "tau")
    bs = {k: v.clone() for k, v in RC.debug_blend_sums(1, P, geom).items()}
    mats3 = (view, full, proj)
    own64 = O64.chain(sc, mats3, W, H, cam.tanfovx, cam.tanfovy, config, radii, ours["means2D"], bs["conic"],
                      bs["normal"], bs["depths"])
    # OUR backward accumulates with float atomics too (RED.128 per tile): the single-row max statistics below move
    # run to run on the ill-conditioned rows, so (a), (b) and (d) take the best of three of our runs against the
    # worst of six reference runs
    own_runs = [(ours, own64)]
    if P <= 50000:
        for _ in range(2):
            o_ = dict(zip(names, RC.rasterize_gaussians_surfel_backward(
                bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, e, view, full, proj,
                ex["prcppoint"], ex["patch_bbox"], cam.tanfovx, cam.tanfovy, ct["color"], ct["normal"], ct["depth"],
                ct["opac"], e, 0, center, geom, R, binning, img, True, config, full_tau=True)))
            b_ = RC.debug_blend_sums(1, P, geom)
            own_runs.append((o_, O64.chain(sc, mats3, W, H, cam.tanfovx, cam.tanfovy, config, radii, o_["means2D"],
                                           b_["conic"], b_["normal"], b_["depths"])))
    co = dv["records"][:, 1, :3].double().cpu()
    kappa = (co[:, 0] * co[:, 2]) / (co[:, 0] * co[:, 2] - co[:, 1] ** 2).clamp_min(1e-300)
    easy = (radii.cpu() > 0) & (kappa < 10)
    assert int(easy.sum()) > nvis // 2

    def relmax(a, b):
        b = b.double().cpu().reshape(P, -1)
        return float((a.double().cpu().reshape(P, -1) - b).abs().max() / (b.abs().max() + 1e-300))

    def rowrel(a, b):
        a, b = a.double().cpu().reshape(P, -1), b.double().cpu().reshape(P, -1)
        return (a - b).abs().max(1).values / (b.abs().max(1).values + 1e-4 * b.abs().max())

    if P <= 50000:
        a64 = O64.blend(sc, mats3, W, H, cam.tanfovx, cam.tanfovy, bg, config, radii, dv["point_list"], dv["ranges"], ct)
        for k, img_ in dict(color=color, normal=normal, depth=depth, opac=opac).items():   # fp64 decisions flip a few pixels
            d = (img_.double().cpu() - a64[k].reshape(img_.shape)).abs()
            assert float(d.flatten().quantile(0.999)) <= 1e-4 * float(a64[k].abs().max()) + 1e-6, k
        g64 = O64.chain(sc, mats3, W, H, cam.tanfovx, cam.tanfovy, config, radii, a64["g_means2D"], a64["g_conic"],
                        a64["g_normal"], a64["g_depths"])
        for n in chain_names:
            e_ours = min(relmax(o_[n], g64[n]) for o_, _ in own_runs)
            e_ref = max(relmax(x[n], g64[n]) for x in runs)
            assert e_ours <= 1.5 * e_ref + 3e-4, ("vs fp64 arbiter", n, e_ours, e_ref)
            x0 = runs[0]
            ref_own = O64.chain(sc, mats3, W, H, cam.tanfovx, cam.tanfovy, config, radii, x0["means2D"],
                                x0["conic"].reshape(P, 4)[:, [0, 1, 3]], x0["normal"], x0["depths"])
            c_ours, c_ref = min(relmax(o_[n], w_[n]) for o_, w_ in own_runs), relmax(x0[n], ref_own[n])
            assert c_ours <= 4.0 * c_ref + 1e-3, ("chain alone", n, c_ours, c_ref)
            assert float(rowrel(ours[n], g64[n])[easy].quantile(0.99)) < 1e-3, ("well-conditioned rows", n)
        tau64 = g64["tau"].sum(0)
        t_ours = min(float((o_["tau"].double().sum(0).cpu() - tau64).abs().max() / tau64.abs().max()) for o_, _ in own_runs)
        t_ref = max(float((x["tau"].double().sum(0).cpu() - tau64).abs().max() / tau64.abs().max()) for x in runs)
        assert t_ours <= max(1.5 * t_ref, 5e-4), ("pose gradient vs fp64", t_ours, t_ref)
    else:
        for n in chain_names:
            assert float(rowrel(ours[n], own64[n])[easy].quantile(0.99)) < 1e-3, ("well-conditioned rows (chain)", n)
            refs = torch.stack([x[n].double().reshape(P, -1) for x in runs])
            mx = refs[0].abs().max() + 1e-30
            err = (ours[n].double().reshape(P, -1) - refs[0]).abs().max(1).values
            spread = float((refs.max(0).values - refs.min(0).values).max() / mx)
            rel = err / (refs[0].abs().max(1).values + 1e-3 * mx)
            assert int((rel > 1e-2).sum()) <= max(2, nvis // 100), (n, int((rel > 1e-2).sum()), nvis)
            assert float(err.max() / mx) < max(0.05, 10 * spread), (n, float(err.max() / mx), spread)
    if CONFIGS[cfg][0]:
        assert float(out[6][:, 2].abs().max()) == 0.0  # dL_dscale.z = 0 for surfaces (RS backward.cu:419)
    out2 = RC.rasterize_gaussians_surfel_backward(
        bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, e, view, full, proj,
        ex["prcppoint"], ex["patch_bbox"], cam.tanfovx, cam.tanfovy, ct["color"], ct["normal"], ct["depth"], ct["opac"],
        e, 0, center, geom, R, binning, img, False, config)
    tau_runs = torch.stack([x["tau"].double().sum(0) for x in runs])
    tau_spread = float((tau_runs.max(0).values - tau_runs.min(0).values).max() / tau_runs[0].abs().max())
    assert _relmax(out2[-1].flatten(), tau_runs[0]) < max(1e-4, 5 * tau_spread)


def test_module_api_mark_visible_and_empty():
    from oracle.ref_api import RefRasterS
    from pings_b200.diff_gaussian_surfel_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    W, H = 320, 200
    cam, sc, mats, ct, ex = _scene(5000, W, H, 9)
    view, full, proj, center = mats
    config = torch.tensor([1, 1, 1, 1, 1], dtype=torch.float32, device="cuda")
    rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, torch.zeros(3, device="cuda"), 1.0, view, full,
                                       proj, ex["patch_bbox"], ex["prcppoint"], 0, center, False, False, config)
    rast = GaussianRasterizer(rs)
    vis = rast.markVisible(sc["means3D"])
    assert torch.equal(vis, RefRasterS().mark_visible(sc["means3D"], view, full))
    assert 0 < int(vis.sum()) < 5000
    # autograd through the module: every leaf and the pose deltas receive a gradient
    leaves = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "colors_precomp", "scales", "rotations")}
    theta = torch.zeros(1, 3, device="cuda", requires_grad=True)
    rho = torch.zeros(1, 3, device="cuda", requires_grad=True)
    m2d = torch.zeros(5000, 3, device="cuda", requires_grad=True)
    color, normal, depth, opac, radii, contrib = rast(leaves["means3D"], m2d, leaves["opacities"],
                                                      colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
                                                      rotations=leaves["rotations"], theta=theta, rho=rho)
    assert color.shape == (3, H, W) and normal.shape == (3, H, W) and depth.shape == (1, H, W) and opac.shape == (1, H, W)
    ((color * ct["color"]).sum() + (normal * ct["normal"]).sum() + (depth * ct["depth"]).sum()
     + (opac * ct["opac"]).sum()).backward()
    for k, v in leaves.items():
        assert v.grad is not None and torch.isfinite(v.grad).all(), k
    assert theta.grad.shape == (1, 3) and rho.grad.shape == (1, 3) and float(theta.grad.abs().sum()) > 0
    assert m2d.grad.shape == (5000, 3)
    # P == 0 short-circuit and argument errors (RS __init__.py:218-224)
    z = torch.zeros(0, 3, device="cuda")
    out = rast(z, z, torch.zeros(0, 1, device="cuda"), colors_precomp=z, scales=z, rotations=torch.zeros(0, 4, device="cuda"))
    assert out[0].shape == (3, H, W) and out[4].numel() == 0 and float(out[0].abs().sum()) == 0.0
    with pytest.raises(Exception):
        rast(z, z, torch.zeros(0, 1, device="cuda"))
