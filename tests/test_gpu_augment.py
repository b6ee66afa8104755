"""The GPU augmentation (ntx_simclr_augment through simclr_b200.augment.SimCLRTransformsGPU) against torchvision / PIL.

The reference's pipeline (dataset.py:19-50) is random; the kernel writes out the parameters it drew, and the test
replays exactly those through torchvision's functional ops on PIL images -- the code the reference's transforms
call -- so the pixel arithmetic is pinned against the reference's own library.  The sampling rules are checked
through their statistics."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from simclr_b200 import _cabi
    _cabi.load()
    return torch.device("cuda:0")


def _images(b, h=32, w=32, seed=0):
    """Smooth random images with saturated regions (exercise clipping, hue wrap-around, grey pixels)."""
    g = torch.Generator().manual_seed(seed)
    low = torch.rand(b, 3, 5, 5, generator=g)
    img = torch.nn.functional.interpolate(low, size=(h, w), mode="bicubic", align_corners=False)
    img = (img * 1.4 - 0.2).clamp(0, 1)
    img[: b // 8] = img[: b // 8].mean(dim=1, keepdim=True)          # some grey images (zero saturation)
    return (img * 255).round().to(torch.uint8).permute(0, 2, 3, 1).contiguous()


def _replay(image_hwc, p, size):
    """torchvision's functional ops on a PIL image with the parameters the kernel drew (dataset.py:19-39)."""
    from PIL import Image
    from torchvision.transforms import functional as F
    img = Image.fromarray(image_hwc.numpy(), "RGB")
    top, left, h, w = (int(p[i]) for i in range(4))
    img = F.resized_crop(img, top, left, h, w, [size, size])        # RandomResizedCrop's default: bilinear
    if p[4] > 0.5:
        img = F.hflip(img)
    if p[5] > 0.5:
        for op in (int(v) for v in p[11:15]):
            if op == 0:
                img = F.adjust_brightness(img, float(p[6]))
            elif op == 1:
                img = F.adjust_contrast(img, float(p[7]))
            elif op == 2:
                img = F.adjust_saturation(img, float(p[8]))
            else:
                img = F.adjust_hue(img, float(p[9]))
    if p[10] > 0.5:
        img = F.rgb_to_grayscale(img, num_output_channels=3)
    return F.to_tensor(img)


@pytest.mark.parametrize("shape,size", [((32, 32), 32), ((48, 40), 32), ((32, 32), 24)])
def test_augmented_views_match_torchvision_replay(shape, size, dev):
    from simclr_b200.augment import SimCLRTransformsGPU
    b = 96
    images = _images(b, *shape, seed=shape[0] + size)
    tf = SimCLRTransformsGPU(strength=0.5, size=size, seed=11)
    v0, v1, params = tf(images.to(dev), return_params=True)
    v = torch.stack([v0, v1], dim=1).cpu()                            # (B, 2, 3, S, S)
    params = params.cpu().numpy()
    exact = total = 0
    worst = 0.0
    for i in range(b):
        for k in range(2):
            want = _replay(images[i], params[i, k], size)
            diff = (v[i, k] - want).abs() * 255.0
            worst = max(worst, float(diff.max()))
            exact += int((diff < 0.5).sum())
            total += diff.numel()
            # a pixel may differ by a few levels where a float rounding lands on the other side of an 8-bit step
            # (hue conversion); anything systematic would move whole images
            assert float((diff > 1.5).float().mean()) < 0.02, (i, k, params[i, k])
    assert exact / total > 0.99, (exact / total, worst)
    assert worst <= 8.0, worst


def test_sampling_rules(dev):
    """torchvision's get_params rules through their statistics over 8192 draws (strength 0.5, 32 x 32 images)."""
    from simclr_b200.augment import SimCLRTransformsGPU
    b = 4096
    images = torch.zeros(b, 32, 32, 3, dtype=torch.uint8, device=dev)
    tf = SimCLRTransformsGPU(strength=0.5, size=32, seed=3)
    _, _, p1 = tf(images, return_params=True)
    _, _, p2 = tf(images, return_params=True)
    assert not torch.equal(p1, p2)                                    # every call draws anew
    p = p1.cpu().numpy().reshape(-1, 16)
    top, left, h, w = p[:, 0], p[:, 1], p[:, 2], p[:, 3]
    assert (h >= 1).all() and (w >= 1).all() and (top >= 0).all() and (left >= 0).all()
    assert (top + h <= 32).all() and (left + w <= 32).all()
    area = h * w / 1024.0
    ratio = w / h
    # U(0.08, 1) of the area minus the rejected boxes: torchvision's own get_params gives mean 0.488 on 32 x 32
    assert 0.07 <= area.min() and area.max() <= 1.0 and abs(area.mean() - 0.488) < 0.015
    assert ratio.min() > 0.6 and ratio.max() < 1.6 and abs(np.log(ratio).mean()) < 0.02    # log-uniform on (3/4, 4/3)
    assert abs(p[:, 4].mean() - 0.5) < 0.03 and abs(p[:, 5].mean() - 0.8) < 0.03 and abs(p[:, 10].mean() - 0.2) < 0.03
    for col in (6, 7, 8):                                             # U(0.6, 1.4)
        assert p[:, col].min() >= 0.6 and p[:, col].max() <= 1.4 and abs(p[:, col].mean() - 1.0) < 0.02
    assert p[:, 9].min() >= -0.1 and p[:, 9].max() <= 0.1 and abs(p[:, 9].mean()) < 0.01   # U(-0.1, 0.1)
    order = p[:, 11:15].astype(int)
    assert (np.sort(order, axis=1) == np.arange(4)).all()             # a permutation, each op equally often first
    assert all(abs((order[:, 0] == k).mean() - 0.25) < 0.03 for k in range(4))
    # the two views of an image are drawn independently
    q = p1.cpu().numpy()
    assert abs(np.corrcoef(q[:, 0, 2], q[:, 1, 2])[0, 1]) < 0.06
    with pytest.raises(TypeError):
        tf(images.float())
    with pytest.raises(RuntimeError):
        tf(images.cpu())
