"""Row f4: per-frame masks, medians and covisibility counts against the reference's torch formulation
(utils/slam_utils.py:7-41, 286-297; utils/camera_utils.py:110-142; utils/slam_frontend.py:1713-1741)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

# the reference's Scharr filter goes through cuDNN convolutions, which may run in TF32 on this GPU; the product computes
# in fp32, so the restatement below is pinned to exact fp32 convolutions
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def _image_gradient(image):          # slam_utils.py:7-24
    c = image.shape[0]
    conv_y = torch.tensor([[3, 0, -3], [10, 0, -10], [3, 0, -3]], dtype=torch.float32, device="cuda")
    conv_x = torch.tensor([[3, 10, 3], [0, 0, 0], [-3, -10, -3]], dtype=torch.float32, device="cuda")
    normalizer = 1.0 / torch.abs(conv_y).sum()
    p_img = F.pad(image, (1, 1, 1, 1), mode="reflect")[None]
    v = normalizer * F.conv2d(p_img, conv_x.view(1, 1, 3, 3).repeat(c, 1, 1, 1), groups=c)
    h = normalizer * F.conv2d(p_img, conv_y.view(1, 1, 3, 3).repeat(c, 1, 1, 1), groups=c)
    return v[0], h[0]


def _image_gradient_mask(image, eps=0.01):   # slam_utils.py:27-41
    c = image.shape[0]
    ones = torch.ones((1, 1, 3, 3), dtype=torch.float32, device="cuda")
    p_img = (torch.abs(F.pad(image, (1, 1, 1, 1), mode="reflect")[None]) > eps).float()
    v = F.conv2d(p_img, ones.repeat(c, 1, 1, 1), groups=c)
    return v[0] == 9, v[0] == 9


def _ref_intensity(img):             # camera_utils.py:113-118
    gray = img.mean(dim=0, keepdim=True)
    gv, gh = _image_gradient(gray)
    mv, mh = _image_gradient_mask(gray)
    return torch.sqrt((gv * mv) ** 2 + (gh * mh) ** 2)


def _img(H, W, seed):
    g = torch.Generator().manual_seed(seed)
    img = torch.rand(3, H, W, generator=g)
    img[:, : H // 6, : W // 3] = 0.0          # a dark corner: exercises the validity mask and the rgb boundary mask
    return img.cuda()


@pytest.mark.parametrize("H,W", [(48, 64), (480, 640), (101, 77)])
def test_grad_intensity_and_masks_match_reference(H, W):
    from monogs_b200 import frame as FR
    img = _img(H, W, 1)
    ref = _ref_intensity(img)
    mine = FR.grad_intensity(img)
    assert torch.allclose(mine, ref[0], rtol=1e-5, atol=2e-6), float((mine - ref[0]).abs().max())
    edge = 1.1
    med = ref.median()
    gm, rm, rmm = FR.frame_masks(img, edge, 0.01)
    # pixels whose intensity sits within rounding of the threshold may flip: compare away from it
    safe = ((ref - med * edge).abs() > 1e-5 * med).float()
    gm_ref = (ref > med * edge).float()
    assert torch.equal(gm * safe, gm_ref * safe) and float(safe.mean()) > 0.999
    rpm = (img.sum(dim=0) > 0.01).view(1, H, W).float()
    assert torch.equal(rmm, rpm) and torch.equal(rm * safe, rpm * gm_ref * safe)


def test_patch_median_masks_match_reference():
    """Replica variant (camera_utils.py:120-129): 32x32 patch medians via unfold / fold."""
    from monogs_b200 import frame as FR
    H, W = 680, 1200
    img = _img(H, W, 2)
    I = _ref_intensity(img).unsqueeze(0)
    I_unf = F.unfold(I, 32, stride=32)
    med, _ = torch.median(I_unf, dim=1, keepdim=True)
    ref = F.fold((I_unf > med * 4.0).float(), I.shape[-2:], 32, stride=32).squeeze(0)
    near = F.fold((((I_unf - med * 4.0).abs() <= 1e-5 * med) & (med > 0)).float(), I.shape[-2:], 32, stride=32).squeeze(0)
    gm, _, _ = FR.frame_masks(img, 4.0, 0.01, patch_median=True)
    safe = 1 - near
    assert torch.equal(gm * safe, ref * safe) and float(safe.mean()) > 0.999
    assert float(gm[:, 672:, :].sum()) == 0 and float(gm[:, :, 1184:].sum()) == 0      # remainder outside whole patches


@pytest.mark.parametrize("n", [1, 2, 5, 1000, 307200])
def test_median_is_torch_lower_median(n):
    from monogs_b200 import frame as FR
    x = torch.randn(n, generator=torch.Generator().manual_seed(n)).cuda() * 3
    x[::7] = float(x[0])                                # ties
    assert torch.equal(FR.median(x)[0], x.median())
    assert torch.equal(FR.median(x.abs())[0], x.abs().median())


def test_median_depth_matches_reference():
    from monogs_b200 import frame as FR
    g = torch.Generator().manual_seed(4)
    depth = (torch.rand(1, 120, 160, generator=g) * 5 - 0.5).cuda()
    opacity = torch.rand(1, 120, 160, generator=g).cuda()
    mask = (torch.rand(1, 120, 160, generator=g) > 0.3).cuda()
    for op, mk in ((opacity, None), (opacity, mask), (None, None)):
        valid = depth > 0
        if op is not None: valid = valid & (op > 0.95)
        if mk is not None: valid = valid & mk
        med, nv = FR.median_depth(depth, op, mk)
        assert int(nv) == int(valid.sum()) and torch.equal(med[0], depth[valid].median())
    med, nv = FR.median_depth(-depth.abs(), None, None)
    assert int(nv) == 0 and bool(torch.isnan(med[0]))


def test_covisibility_counts_match_reference():
    from monogs_b200 import frame as FR
    P, K = 100003, 5
    g = torch.Generator().manual_seed(6)
    cur = torch.randint(0, 3, (P,), generator=g, dtype=torch.int32).cuda()
    kfs = [torch.randint(0, 2 + k, (P,), generator=g, dtype=torch.int32).cuda() for k in range(K)]
    cb = FR.pack_visibility(cur)
    kb = torch.stack([FR.pack_visibility(k) for k in kfs])
    out = FR.covisibility(P, cb, kb).cpu()
    cv = cur > 0
    for k in range(K):
        kv = kfs[k] > 0
        assert int(out[k, 0]) == int(torch.logical_and(cv, kv).count_nonzero())
        assert int(out[k, 1]) == int(torch.logical_or(cv, kv).count_nonzero())
        assert int(out[k, 2]) == int(kv.count_nonzero())
    assert int(out[K, 0]) == int(cv.count_nonzero())
    assert FR.covisibility(P, cb, kb[:0]).shape == (1, 3)
