"""Stand-in for `build_segmentor(cfg.model)` of configs/config_pretrain.py (mmseg/mmcv are not
installable in this image).  OUT OF SCOPE for the kernels: it stays on
PyTorch/cuDNN and is timed separately by bench.py.

Same layer layout and parameter shapes as the reference's segmentor in contrast mode:
  backbone     ResNet-50, strides (1,2,2,1), dilations (1,1,1,2), contract_dilation=True
               (configs/config_pretrain.py:8-19) == torchvision resnet50 with
               replace_stride_with_dilation=[False, False, True]  -> output stride 16
  decode_head  ASPPHead(in 2048, channels 512, dilations (1,6,12,18), contrast=True)
               (mmsegmentation/mmseg/models/decode_heads/aspp_head.py:65-131): image pool +
               4 ASPP branches -> 3x3 bottleneck -> contrast_conv 512 -> 512 -> ReLU -> dim
               `conv_seg` (512 -> num_classes) exists but is unused in contrast mode, exactly as
               in the reference (hence find_unused_parameters=True at main.py:456-461).
66.0 M parameters, like the mmseg model.
"""
import torch
import torch.nn as nn


def _conv_bn_relu(cin, cout, k, dilation=1):
    pad = 0 if k == 1 else dilation
    return nn.Sequential(nn.Conv2d(cin, cout, k, padding=pad, dilation=dilation, bias=False),
                         nn.BatchNorm2d(cout), nn.ReLU(inplace=True))


class _ResNet50Stages(nn.Module):
    """`.backbone(x)` -> 4 stage outputs (out_indices=(0,1,2,3)); probed at builder.py:398-401."""

    def __init__(self):
        super().__init__()
        from torchvision.models import resnet50
        net = resnet50(weights=None, replace_stride_with_dilation=[False, False, True])
        self.stem = nn.Sequential(net.conv1, net.bn1, net.relu, net.maxpool)
        self.layer1, self.layer2, self.layer3, self.layer4 = net.layer1, net.layer2, net.layer3, net.layer4

    def forward(self, x):
        x = self.stem(x)
        c1 = self.layer1(x)
        c2 = self.layer2(c1)
        c3 = self.layer3(c2)
        c4 = self.layer4(c3)
        return c1, c2, c3, c4

    def init_weights(self):      # ImageNet weights are not available offline (builder.py:375-378)
        pass


class ASPPContrastHead(nn.Module):
    def __init__(self, in_channels=2048, channels=512, dilations=(1, 6, 12, 18), dim=128, num_classes=2):
        super().__init__()
        self.image_pool = nn.Sequential(nn.AdaptiveAvgPool2d(1), _conv_bn_relu(in_channels, channels, 1))
        self.aspp_modules = nn.ModuleList(
            [_conv_bn_relu(in_channels, channels, 1 if d == 1 else 3, dilation=d) for d in dilations])
        self.bottleneck = _conv_bn_relu((len(dilations) + 1) * channels, channels, 3)
        self.conv_seg = nn.Conv2d(channels, num_classes, 1)         # unused in contrast mode
        self.contrast_conv = nn.Sequential(nn.Conv2d(channels, channels, 1), nn.ReLU(),
                                           nn.Conv2d(channels, dim, 1))

    def forward(self, x):
        # bilinear resize of a 1x1 map (aspp_head.py:105-110) is a broadcast (equal up to the 1-ulp rounding of w0*x + w1*x)
        outs = [self.image_pool(x).expand(-1, -1, x.shape[2], x.shape[3])]
        outs.extend(m(x) for m in self.aspp_modules)
        return self.contrast_conv(self.bottleneck(torch.cat(outs, dim=1)))


class DeepLabV3Encoder(nn.Module):
    """forward(img [B,3,H,W]) -> [B,dim,H/16,W/16]."""

    def __init__(self, dim=128):
        super().__init__()
        self.backbone = _ResNet50Stages()
        self.decode_head = ASPPContrastHead(dim=dim)

    def forward(self, x):
        return self.decode_head(self.backbone(x)[3])

    def graph_stages(self):
        """The chain forward() runs, stage by stage (for MODEL.graph_encoders(query_stages=...))."""
        b = self.backbone
        return [b.stem, b.layer1, b.layer2, b.layer3, b.layer4, self.decode_head]


# ------------------------------------------------------------------------------------------------------------
# Exact rewrite of strongly dilated 3x3 convolutions on small feature maps (no counterpart in the reference).
# A 3x3 convolution with dilation d and padding d reads, for the tap (ky, kx), the input at offset
# ((ky-1)*d, (kx-1)*d): on an H x W map that offset lands in the zero padding for every output position outside
# a (H - |dy|) x (W - |dx|) window, and for |dy| >= H or |dx| >= W the tap never sees data at all.  At the 14 x 14
# map CP2 trains at (224 x 224 views, output stride 16) the ASPP branch with dilation 18 IS a 1x1 convolution
# with the centre weights, and the dilation-12 branch touches 2-pixel borders only (the degeneration the DeepLabv3
# paper describes for large rates) — yet cuDNN runs both as full 3x3 implicit GEMMs (118 GFLOP each at B=32).
# folded_dilated_conv2d applies every tap as a 1x1 convolution on exactly the window where it sees data:
# the same sums (taps that would multiply zeros are the only ones dropped), the same parameters / checkpoint keys,
# gradients to the full 3x3 weight included.
def _tap_windows(H, W, d):
    """[(ky, kx, y0, y1, x0, x1)]: output window [y0:y1, x0:x1] on which tap (ky, kx) reads inside the map."""
    out = []
    for ky in range(3):
        for kx in range(3):
            dy, dx = (ky - 1) * d, (kx - 1) * d
            y0, y1, x0, x1 = max(0, -dy), min(H, H - dy), max(0, -dx), min(W, W - dx)
            if y1 > y0 and x1 > x0:
                out.append((ky, kx, y0, y1, x0, x1))
    return out


def folded_dilated_conv2d(x, weight, d, max_cover=0.15):
    """conv2d(x, weight [O,I,3,3], stride 1, padding d, dilation d) computed tap by tap on the windows where a tap
    sees data.  Returns None when the taps cover more than `max_cover` of the 9*H*W products (the plain
    convolution is the better kernel then).  Measured on a B200 at B=32, 2048 -> 512 channels, 14 x 14: dilation 18
    (cover 0.11: the centre tap alone) 1.84 ms per training step faster than cuDNN's 3x3 kernel; dilation 12 (cover
    0.18: eight extra 2-pixel windows) 0.47 ms SLOWER tap by tap, dilation 6 (0.51) slower still: hence 0.15."""
    H, W = x.shape[-2:]
    wins = _tap_windows(H, W, d)
    if sum((y1 - y0) * (x1 - x0) for _, _, y0, y1, x0, x1 in wins) > max_cover * 9 * H * W:
        return None
    import torch.nn.functional as F
    out = F.conv2d(x, weight[:, :, 1:2, 1:2])                 # the centre tap sees the whole map
    for ky, kx, y0, y1, x0, x1 in wins:
        if ky == 1 and kx == 1:
            continue
        dy, dx = (ky - 1) * d, (kx - 1) * d
        out[:, :, y0:y1, x0:x1] += F.conv2d(x[:, :, y0 + dy:y1 + dy, x0 + dx:x1 + dx], weight[:, :, ky:ky + 1, kx:kx + 1])
    return out


def fold_dilated_convs(module, max_cover=0.15):
    """Patch every eligible nn.Conv2d below `module` (3x3, stride 1, groups 1, padding == dilation > 1, zero
    padding) so that its forward uses folded_dilated_conv2d whenever the input map is small enough for it to
    pay; other inputs run the stock convolution.  Parameters and state_dict are untouched.  -> number of
    patched convolutions."""
    import types
    n = 0
    for m in module.modules():
        if (isinstance(m, nn.Conv2d) and m.kernel_size == (3, 3) and m.stride == (1, 1) and m.groups == 1
                and m.dilation[0] == m.dilation[1] and m.dilation[0] > 1 and m.padding == m.dilation
                and m.padding_mode == "zeros" and not hasattr(m, "_cp2_stock_forward")):
            m._cp2_stock_forward = m.forward

            def forward(self, x, _cover=max_cover):
                y = folded_dilated_conv2d(x, self.weight, self.dilation[0], _cover) if x.dim() == 4 else None
                if y is None:
                    return self._cp2_stock_forward(x)
                return y if self.bias is None else y + self.bias.view(1, -1, 1, 1)
            m.forward = types.MethodType(forward, m)
            n += 1
    return n
