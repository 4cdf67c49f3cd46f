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
