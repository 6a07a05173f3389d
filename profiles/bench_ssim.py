"""Fused SSIM / mapping colour loss timed on the GPU (CUDA events, L2 flushed), torch's conv2d formulation
(the reference's code path, restated) beside it.  usage: python profiles/bench_ssim.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
import torch
import torch.nn.functional as F
from eags_raster.losses import ssim

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def torch_ssim(img1, img2):    # src/entities/losses.py:88-113 (restated)
    g = torch.tensor([torch.exp(torch.tensor(-(x - 5) ** 2 / 4.5)) for x in range(11)])
    g = (g / g.sum()).unsqueeze(1)
    w = (g @ g.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().cuda()
    mu1, mu2 = F.conv2d(img1, w, padding=5, groups=3), F.conv2d(img2, w, padding=5, groups=3)
    s11 = F.conv2d(img1 * img1, w, padding=5, groups=3) - mu1 * mu1
    s22 = F.conv2d(img2 * img2, w, padding=5, groups=3) - mu2 * mu2
    s12 = F.conv2d(img1 * img2, w, padding=5, groups=3) - mu1 * mu2
    return (((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s11 + s22 + 9e-4))).mean()


def timed(fn, reps=30):
    for _ in range(3): fn()
    t = 0.0
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        t += e0.elapsed_time(e1)
    return t / reps


for (H, W) in [(480, 640), (680, 1200), (1168, 1752)]:
    gt = torch.rand(3, H, W, device="cuda")
    img = (gt + 0.1 * torch.randn_like(gt)).clamp(0, 1).requires_grad_(True)

    def fused():
        img.grad = None
        ssim(img, gt).backward()

    def ref():
        img.grad = None
        torch_ssim(img[None], gt[None]).backward()

    n = 3 * H * W
    ms_f, ms_r = timed(fused), timed(ref)
    # algorithmic bytes: fwd reads 2 images, writes 3 partial maps; bwd reads 3 maps + 2 images, writes 1
    print(json.dumps({"op": "ssim fwd+bwd", "image": f"3x{H}x{W}", "fused_ms": ms_f, "torch_conv2d_ms": ms_r,
                      "algorithmic_GBps": 11 * 4 * n / ms_f / 1e6}))
