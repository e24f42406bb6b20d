"""SlowFast-R50 (8x8, alpha=4, beta=1/8) in stock PyTorch -- CONTEXT ONLY, not part of the path.

The reference takes this model from ``torch.hub.load("facebookresearch/pytorchvideo", "slowfast_r50")``
and swaps its head for ``create_res_basic_head(in_features=2304, ...)`` (``run.py:105-110``).  Neither
the hub (no network) nor pytorchvideo is available here, and BASELINE.json configs[4] wants an
end-to-end training step for context, so this file restates the published architecture
(Feichtenhofer et al., "SlowFast Networks for Video Recognition", Table 1; pytorchvideo's
``create_slowfast`` defaults) with random initialisation.  What matters to the preprocessing path is
only its input contract: ``[slow B x 3 x 8 x S x S, fast B x 3 x 32 x S x S]``.
"""
from __future__ import annotations

from typing import List

import torch
import torch.nn as nn


def conv_bn_relu(cin, cout, kernel, stride, padding, relu=True):
    layers = [nn.Conv3d(cin, cout, kernel, stride, padding, bias=False), nn.BatchNorm3d(cout)]
    if relu:
        layers.append(nn.ReLU(inplace=True))
    return nn.Sequential(*layers)


class Bottleneck(nn.Module):
    """1x1x1 (temporal kernel kt) -> 1x3x3 (spatial stride) -> 1x1x1, residual."""

    def __init__(self, cin, cinner, cout, kt, stride):
        super().__init__()
        self.a = conv_bn_relu(cin, cinner, (kt, 1, 1), 1, (kt // 2, 0, 0))
        self.b = conv_bn_relu(cinner, cinner, (1, 3, 3), (1, stride, stride), (0, 1, 1))
        self.c = conv_bn_relu(cinner, cout, 1, 1, 0, relu=False)
        self.short = None
        if cin != cout or stride != 1:
            self.short = conv_bn_relu(cin, cout, 1, (1, stride, stride), 0, relu=False)
        self.relu = nn.ReLU(inplace=True)

    def forward(self, x):
        y = self.c(self.b(self.a(x)))
        return self.relu(y + (x if self.short is None else self.short(x)))


def stage(cin, cinner, cout, depth, kt, stride):
    blocks = [Bottleneck(cin, cinner, cout, kt, stride)]
    blocks += [Bottleneck(cout, cinner, cout, kt, 1) for _ in range(depth - 1)]
    return nn.Sequential(*blocks)


class FuseFastToSlow(nn.Module):
    """Lateral connection: time-strided conv on the fast pathway, concatenated onto the slow one."""

    def __init__(self, cfast, alpha):
        super().__init__()
        self.conv = conv_bn_relu(cfast, 2 * cfast, (7, 1, 1), (alpha, 1, 1), (3, 0, 0))

    def forward(self, slow, fast):
        return torch.cat([slow, self.conv(fast)], dim=1), fast


class SlowFastR50(nn.Module):
    def __init__(self, num_classes: int = 700, alpha: int = 4, beta_inv: int = 8, dropout: float = 0.5):
        super().__init__()
        cs, cf = 64, 64 // beta_inv
        self.slow_stem = nn.Sequential(conv_bn_relu(3, cs, (1, 7, 7), (1, 2, 2), (0, 3, 3)),
                                       nn.MaxPool3d((1, 3, 3), (1, 2, 2), (0, 1, 1)))
        self.fast_stem = nn.Sequential(conv_bn_relu(3, cf, (5, 7, 7), (1, 2, 2), (2, 3, 3)),
                                       nn.MaxPool3d((1, 3, 3), (1, 2, 2), (0, 1, 1)))
        self.fuse = nn.ModuleList([FuseFastToSlow(cf, alpha)])
        depths, slow_kt = (3, 4, 6, 3), (1, 1, 3, 3)
        self.slow_stages, self.fast_stages = nn.ModuleList(), nn.ModuleList()
        cin_s, cin_f = cs + 2 * cf, cf
        for i, depth in enumerate(depths):
            inner_s, out_s = 64 * 2 ** i, 256 * 2 ** i
            inner_f, out_f = inner_s // beta_inv, out_s // beta_inv
            stride = 1 if i == 0 else 2
            self.slow_stages.append(stage(cin_s, inner_s, out_s, depth, slow_kt[i], stride))
            self.fast_stages.append(stage(cin_f, inner_f, out_f, depth, 3, stride))
            if i < len(depths) - 1:
                self.fuse.append(FuseFastToSlow(out_f, alpha))
                cin_s, cin_f = out_s + 2 * out_f, out_f
            else:
                cin_s, cin_f = out_s, out_f
        self.pool = nn.AdaptiveAvgPool3d(1)
        self.dropout = nn.Dropout(dropout)
        self.proj = nn.Linear(cin_s + cin_f, num_classes)  # 2048 + 256 = 2304 (run.py:109)

    def forward(self, x: List[torch.Tensor]) -> torch.Tensor:
        slow, fast = self.slow_stem(x[0]), self.fast_stem(x[1])
        slow, fast = self.fuse[0](slow, fast)
        for i, (ss, fs) in enumerate(zip(self.slow_stages, self.fast_stages)):
            slow, fast = ss(slow), fs(fast)
            if i + 1 < len(self.fuse):
                slow, fast = self.fuse[i + 1](slow, fast)
        feat = torch.cat([self.pool(slow).flatten(1), self.pool(fast).flatten(1)], dim=1)
        return self.proj(self.dropout(feat))


if __name__ == "__main__":
    m = SlowFastR50(700)
    print("parameters: %.1f M" % (sum(p.numel() for p in m.parameters()) / 1e6))
    with torch.no_grad():
        print(m([torch.zeros(1, 3, 8, 64, 64), torch.zeros(1, 3, 32, 64, 64)]).shape)
