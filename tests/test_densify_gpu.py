"""Arena densify / prune (mgs_densify_plan + mgs_densify_apply) against a torch restatement of
GaussianModel.densify_and_prune / densify_and_clone / densify_and_split / prune_points
(gaussian_splatting/scene/gaussian_model.py:470-691), with the split noise supplied explicitly so that the arithmetic
is comparable (the draws of torch.normal themselves can not be pinned)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def _model(P, M, isotropic, seed=0):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    p = {"xyz": r(P, 3), "f_dc": r(P, 1, 3), "opacity": r(P, 1) * 2.5, "scaling": r(P, 1 if isotropic else 3) * 0.8 - 3.0,
         "rotation": r(P, 4)}
    if M > 1:
        p["f_rest"] = r(P, M - 1, 3)
    p = {k: v.cuda().contiguous() for k, v in p.items()}
    m = {k: torch.randn(v.shape, generator=g).cuda() for k, v in p.items()}
    v = {k: torch.rand(v.shape, generator=g).cuda() for k, v in p.items()}
    accum = torch.rand(P, generator=g).cuda() * 2e-3
    denom = torch.randint(0, 4, (P,), generator=g).float().cuda()      # zeros -> NaN grads -> treated as 0
    kf = torch.randint(0, 9, (P,), generator=g, dtype=torch.int32).cuda()
    nobs = torch.randint(0, 5, (P,), generator=g, dtype=torch.int32).cuda()
    return p, m, v, accum, denom, kf, nobs


def _build_rotation(r):   # gaussian_splatting/utils/general_utils.py:114-137
    q = r / torch.sqrt((r * r).sum(1))[:, None]
    R = torch.zeros(q.size(0), 3, 3, device=r.device)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - w * z); R[:, 0, 2] = 2 * (x * z + w * y)
    R[:, 1, 0] = 2 * (x * y + w * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - w * x)
    R[:, 2, 0] = 2 * (x * z - w * y); R[:, 2, 1] = 2 * (y * z + w * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


class _RefModel:
    """The reference's sequence of cat / boolean-mask operations on plain tensors (parameters, Adam moments, ids)."""

    def __init__(self, p, m, v, kf, nobs, percent_dense=0.01):
        self.p, self.m, self.v = dict(p), dict(m), dict(v)
        self.kf, self.nobs, self.percent_dense = kf.clone(), nobs.clone(), percent_dense

    def scaling(self):
        s = torch.exp(self.p["scaling"])
        return s.repeat(1, 3) if s.size(1) == 1 else s

    def opacity(self):
        return torch.sigmoid(self.p["opacity"])

    def postfix(self, new, kf, nobs):          # cat_tensors_to_optimizer + densification_postfix (:529-598)
        for k in self.p:
            self.p[k] = torch.cat((self.p[k], new[k]), 0)
            self.m[k] = torch.cat((self.m[k], torch.zeros_like(new[k])), 0)
            self.v[k] = torch.cat((self.v[k], torch.zeros_like(new[k])), 0)
        self.kf, self.nobs = torch.cat((self.kf, kf)), torch.cat((self.nobs, nobs))

    def prune(self, mask):                     # prune_points + _prune_optimizer (:488-527)
        keep = ~mask
        for k in self.p:
            self.p[k], self.m[k], self.v[k] = self.p[k][keep], self.m[k][keep], self.v[k][keep]
        self.kf, self.nobs = self.kf[keep], self.nobs[keep]

    def clone(self, grads, thr, extent):       # :644-674
        sel = (torch.norm(grads, dim=-1) >= thr) & (self.scaling().max(1).values <= self.percent_dense * extent)
        self.postfix({k: t[sel] for k, t in self.p.items()}, self.kf[sel], self.nobs[sel])

    def split(self, grads, thr, extent, noise, N=2):   # :600-642
        n_init = self.p["xyz"].size(0)
        padded = torch.zeros(n_init, device="cuda")
        padded[: grads.size(0)] = grads.squeeze()
        sel = (padded >= thr) & (self.scaling().max(1).values > self.percent_dense * extent)
        stds = self.scaling()[sel].repeat(N, 1)
        samples = noise * stds                                         # torch.normal(mean=0, std=stds)
        rots = _build_rotation(self.p["rotation"][sel]).repeat(N, 1, 1)
        new = {k: t[sel].repeat(N, *([1] * (t.dim() - 1))) for k, t in self.p.items()}
        new["xyz"] = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + self.p["xyz"][sel].repeat(N, 1)
        s_new = torch.log(self.scaling()[sel].repeat(N, 1) / (0.8 * N))
        new["scaling"] = s_new[:, :1] if self.p["scaling"].size(1) == 1 else s_new
        self.postfix(new, self.kf[sel].repeat(N), self.nobs[sel].repeat(N))
        self.prune(torch.cat((sel, torch.zeros(N * int(sel.sum()), device="cuda", dtype=torch.bool))))
        return int(sel.sum())

    def densify_and_prune(self, accum, denom, max_grad, min_opacity, extent, max_screen_size, noise):   # :676-691
        grads = (accum / denom).view(-1, 1)
        grads[grads.isnan()] = 0.0
        self.clone(grads, max_grad, extent)
        self.split(grads, max_grad, extent, noise)
        mask = (self.opacity() < min_opacity).squeeze()
        if max_screen_size:
            max_radii2D = torch.zeros(self.p["xyz"].size(0), device="cuda")     # reset by densification_postfix (:594)
            mask = mask | (max_radii2D > max_screen_size) | (self.scaling().max(1).values > 0.1 * extent)
        self.prune(mask)


@pytest.mark.parametrize("P,M,isotropic,max_screen", [(5000, 1, False, 20.0), (3000, 4, False, None), (2000, 1, True, 20.0),
                                                       (70000, 16, False, 20.0)])
def test_densify_and_prune_matches_reference_sequence(P, M, isotropic, max_screen):
    from monogs_b200 import densify as DZ
    p, m, v, accum, denom, kf, nobs = _model(P, M, isotropic)
    max_grad, min_opacity, extent = 4e-4, 0.1, 6.0
    # number of split-selected Gaussians, to draw the noise in the reference's layout [N * n_sel, 3]
    grads = accum / denom
    grads[grads.isnan()] = 0.0
    s = torch.exp(p["scaling"])
    smax = (s.repeat(1, 3) if isotropic else s).max(1).values
    n_sel = int(((grads >= max_grad) & (smax > 0.01 * extent)).sum())
    assert 0 < n_sel < P
    noise = torch.randn(2 * n_sel, 3, generator=torch.Generator().manual_seed(3)).cuda()
    ref = _RefModel(p, m, v, kf, nobs)
    ref.densify_and_prune(accum.clone(), denom.clone(), max_grad, min_opacity, extent, max_screen, noise)
    out = DZ.densify_and_prune(p, m, v, accum, denom, max_grad, min_opacity, extent, max_screen, unique_kfIDs=kf, n_obs=nobs,
                               noise=noise)
    torch.cuda.synchronize()
    assert out["counts"]["split_selected"] == n_sel
    P_new = ref.p["xyz"].size(0)
    assert out["params"]["xyz"].size(0) == P_new and P_new != P
    for k in p:
        if k in ("xyz", "scaling"):      # children carry new arithmetic (expf / logf / rotation): float tolerance
            assert torch.allclose(out["params"][k], ref.p[k], rtol=1e-5, atol=1e-6), k
        else:
            assert torch.equal(out["params"][k], ref.p[k]), k
        assert torch.equal(out["exp_avg"][k], ref.m[k]) and torch.equal(out["exp_avg_sq"][k], ref.v[k]), k
    assert torch.equal(out["unique_kfIDs"], ref.kf) and torch.equal(out["n_obs"], ref.nobs)
    # untouched originals keep their bits (also in xyz / scaling): compare the kept prefix exactly
    nk = out["counts"]["kept"] + out["counts"]["cloned"]
    assert torch.equal(out["params"]["xyz"][:nk], ref.p["xyz"][:nk]) and torch.equal(out["params"]["scaling"][:nk], ref.p["scaling"][:nk])


def test_prune_points_matches_reference():
    from monogs_b200 import densify as DZ
    p, m, v, accum, denom, kf, nobs = _model(4000, 4, False, seed=2)
    mask = torch.rand(4000, generator=torch.Generator().manual_seed(9)).cuda() < 0.3
    ref = _RefModel(p, m, v, kf, nobs)
    ref.prune(mask)
    out = DZ.prune_points(p, m, v, mask, unique_kfIDs=kf, n_obs=nobs)
    for k in p:
        assert torch.equal(out["params"][k], ref.p[k]) and torch.equal(out["exp_avg"][k], ref.m[k]) and torch.equal(out["exp_avg_sq"][k], ref.v[k])
    assert torch.equal(out["unique_kfIDs"], ref.kf) and torch.equal(out["n_obs"], ref.nobs)
    # everything pruned / nothing pruned
    none = DZ.prune_points(p, m, v, torch.zeros(4000, dtype=torch.bool).cuda())
    assert torch.equal(none["params"]["xyz"], p["xyz"])
    allp = DZ.prune_points(p, m, v, torch.ones(4000, dtype=torch.bool).cuda())
    assert allp["params"]["xyz"].size(0) == 0
