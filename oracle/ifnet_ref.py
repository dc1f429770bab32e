"""CPU oracle: RIFE v4.25 IFNet restated in plain PyTorch fp32.

TEST INFRASTRUCTURE ONLY.  Nothing on the product path may import this file;
only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
arm do, and only as the checker / the CPU baseline.

PARITY UNPINNED: the arithmetic lives in hzwer/Practical-RIFE (submodule pin
17d8c7a1005b37f4c97bfee04e316aaec7fdc536, release v4.25 train_log.zip), which
is absent from the reference tree (.SUBMODULES.json:9-15 status "missing").
This file restates its published architecture per SURVEY.md Appendix A and is
anchored on the reference's own call sites:
    models/sota/rife_wrapper.py:93,102-105,146   Model().load_model/eval/device/inference
    scripts/run_experiments.py:264-268,385        same surface via train_log.RIFE_HDv3
The only structural check available is the parameter count 5 664 776.
"""
import math
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

PARAM_COUNT = 5_664_776
BLOCK_C = (192, 128, 96, 64, 32)
WEIGHT_SEED = 1234


def _conv(i, o, k=3, s=1, p=1):
    # Appendix A.2: conv() = Conv2d(bias) + LeakyReLU(0.2)
    return nn.Sequential(nn.Conv2d(i, o, k, s, p, bias=True), nn.LeakyReLU(0.2, True))


class Head(nn.Module):
    """3->16 (s2) -> 16 -> 16 -> deconv 16->4 (x2); LReLU after the first three only."""

    def __init__(self):
        super().__init__()
        self.cnn0 = nn.Conv2d(3, 16, 3, 2, 1)
        self.cnn1 = nn.Conv2d(16, 16, 3, 1, 1)
        self.cnn2 = nn.Conv2d(16, 16, 3, 1, 1)
        self.cnn3 = nn.ConvTranspose2d(16, 4, 4, 2, 1)
        self.relu = nn.LeakyReLU(0.2, True)

    def forward(self, x):
        x = self.relu(self.cnn0(x))
        x = self.relu(self.cnn1(x))
        x = self.relu(self.cnn2(x))
        return self.cnn3(x)


class ResConv(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, 1, 1)
        self.beta = nn.Parameter(torch.ones((1, c, 1, 1)))
        self.relu = nn.LeakyReLU(0.2, True)

    def forward(self, x):
        return self.relu(self.conv(x) * self.beta + x)


class IFBlock(nn.Module):
    def __init__(self, in_planes, c):
        super().__init__()
        self.conv0 = nn.Sequential(_conv(in_planes, c // 2, 3, 2, 1), _conv(c // 2, c, 3, 2, 1))
        self.convblock = nn.Sequential(*[ResConv(c) for _ in range(8)])
        self.lastconv = nn.Sequential(nn.ConvTranspose2d(c, 4 * 13, 4, 2, 1), nn.PixelShuffle(2))

    def forward(self, x, flow=None, scale=1):
        x = F.interpolate(x, scale_factor=1.0 / scale, mode="bilinear", align_corners=False)
        if flow is not None:
            flow = F.interpolate(flow, scale_factor=1.0 / scale, mode="bilinear",
                                 align_corners=False) * (1.0 / scale)
            x = torch.cat((x, flow), 1)
        feat = self.convblock(self.conv0(x))
        tmp = self.lastconv(feat)
        tmp = F.interpolate(tmp, scale_factor=scale, mode="bilinear", align_corners=False)
        return tmp[:, :4] * scale, tmp[:, 4:5], tmp[:, 5:]


def warp(x, flow):
    """grid_sample(bilinear, border, align_corners=True) at (j+u, i+v)  (Appendix A.2)."""
    n, _, h, w = x.shape
    gx = torch.linspace(-1.0, 1.0, w, dtype=x.dtype).view(1, 1, 1, w).expand(n, -1, h, -1)
    gy = torch.linspace(-1.0, 1.0, h, dtype=x.dtype).view(1, 1, h, 1).expand(n, -1, -1, w)
    grid = torch.cat([gx, gy], 1)
    fl = torch.cat([flow[:, 0:1] / ((w - 1.0) / 2.0), flow[:, 1:2] / ((h - 1.0) / 2.0)], 1)
    g = (grid + fl).permute(0, 2, 3, 1)
    return F.grid_sample(x, g, mode="bilinear", padding_mode="border", align_corners=True)


class IFNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.block0 = IFBlock(7 + 8, BLOCK_C[0])
        self.block1 = IFBlock(8 + 4 + 8 + 8, BLOCK_C[1])
        self.block2 = IFBlock(8 + 4 + 8 + 8, BLOCK_C[2])
        self.block3 = IFBlock(8 + 4 + 8 + 8, BLOCK_C[3])
        self.block4 = IFBlock(8 + 4 + 8 + 8, BLOCK_C[4])
        self.encode = Head()

    def forward(self, x, timestep=0.5, scale_list=(16, 8, 4, 2, 1), return_all=False):
        img0, img1 = x[:, :3], x[:, 3:6]
        if not torch.is_tensor(timestep):
            timestep = (x[:, :1] * 0 + 1) * timestep
        else:
            timestep = timestep.view(-1, 1, 1, 1).to(x.dtype) * (x[:, :1] * 0 + 1)
        f0, f1 = self.encode(img0), self.encode(img1)
        blocks = [self.block0, self.block1, self.block2, self.block3, self.block4]
        flow = mask = feat = None
        w0, w1 = img0, img1
        flows = []
        for i, blk in enumerate(blocks):
            if flow is None:
                flow, mask, feat = blk(torch.cat((img0, img1, f0, f1, timestep), 1), None,
                                       scale=scale_list[i])
            else:
                wf0 = warp(f0, flow[:, :2])
                wf1 = warp(f1, flow[:, 2:4])
                fd, m0, feat = blk(torch.cat((w0, w1, wf0, wf1, timestep, mask, feat), 1), flow,
                                   scale=scale_list[i])
                mask = m0
                flow = flow + fd
            w0 = warp(img0, flow[:, :2])
            w1 = warp(img1, flow[:, 2:4])
            flows.append(flow)
        m = torch.sigmoid(mask)
        merged = w0 * m + w1 * (1 - m)
        if return_all:
            return merged, flow, mask, (f0, f1), flows
        return merged


def make_ifnet(seed=WEIGHT_SEED):
    """Random-init weights as SURVEY.md §8(d) specifies (default torch init, beta = 1)."""
    g = torch.random.get_rng_state()
    torch.manual_seed(seed)
    net = IFNet().eval()
    torch.random.set_rng_state(g)
    n = sum(p.numel() for p in net.parameters())
    assert n == PARAM_COUNT, f"IFNet parameter count {n} != {PARAM_COUNT}"
    return net


class Model:
    """The 4-method surface every reference loader uses (train_log/RIFE_HDv3.py shape)."""

    def __init__(self, local_rank=-1):
        self.flownet = make_ifnet()
        self.version = 4.25

    def load_model(self, path, rank=0):
        f = os.path.join(str(path), "flownet.pkl")
        if os.path.exists(f):
            sd = torch.load(f, map_location="cpu")
            sd = {k.replace("module.", "", 1) if k.startswith("module.") else k: v for k, v in sd.items()}
            self.flownet.load_state_dict(sd, strict=False)

    def eval(self):
        self.flownet.eval()

    def device(self):
        pass  # oracle stays on the CPU

    @torch.no_grad()
    def inference(self, img0, img1, timestep=0.5, scale=1.0):
        imgs = torch.cat((img0, img1), 1)
        sl = [16 / scale, 8 / scale, 4 / scale, 2 / scale, 1 / scale]
        return self.flownet(imgs, timestep, sl)


# ---- the reference wrapper's host-side prologue / epilogue, restated -------------------

def to_tensor(frame):
    """models/base.py:175-186."""
    return (torch.from_numpy(frame).permute(2, 0, 1).float() / 255.0).unsqueeze(0)


def pad_reflect(t, multiple):
    """models/base.py:203-225 (bottom/right reflect pad)."""
    _, _, h, w = t.shape
    ph = (multiple - h % multiple) % multiple
    pw = (multiple - w % multiple) % multiple
    if ph == 0 and pw == 0:
        return t
    return F.pad(t, (0, pw, 0, ph), mode="reflect")


def to_numpy_u8(t):
    """models/base.py:188-201: (x*255).clip(0,255).astype(uint8) -- truncation."""
    import numpy as np
    if t.dim() == 4:
        t = t.squeeze(0)
    fr = t.permute(1, 2, 0).cpu().numpy()
    return (fr * 255).clip(0, 255).astype(np.uint8)


def pad_multiple_for(scale):
    """SURVEY Appendix C decision: max(32, 64/scale)."""
    return int(max(32, math.ceil(64 / scale)))


@torch.no_grad()
def interpolate_ref(model, frame0, frame1, timestamps, scale=1.0, return_float=False):
    """RIFEModel.interpolate (rife_wrapper.py:110-154) on the oracle Model."""
    h, w = frame0.shape[:2]
    m = pad_multiple_for(scale)
    i0 = pad_reflect(to_tensor(frame0), m)
    i1 = pad_reflect(to_tensor(frame1), m)
    outs, outs_f = [], []
    for t in timestamps:
        mid = model.inference(i0, i1, timestep=t, scale=scale)[:, :, :h, :w]
        outs.append(to_numpy_u8(mid))
        outs_f.append(mid[0].permute(1, 2, 0).contiguous().numpy())
    return (outs, outs_f) if return_float else outs
