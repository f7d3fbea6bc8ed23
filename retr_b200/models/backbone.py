"""ResNet backbone parameter tree (reference models/backbone.py + torchvision ResNet).

The modules here only *hold* parameters under the reference's state_dict names
(``backbone.body.layer3.7.conv2.weight`` ...).  The arithmetic — conv + folded frozen BN +
ReLU/residual, max-pool, mask sub-sampling — runs in the native sm_100a library
(``csrc/conv_*.cu``), driven by ``retr_b200.runtime.Engine``.

Initialisation replays torchvision's RNG consumption order exactly (construction order of
the convolutions, the throw-away classifier, then kaiming-normal over ``modules()``), so
``torch.manual_seed(s); build_model(cfg)`` yields bit-identical weights on both sides
(checked in tests/test_init_parity.py against the fixture made from the live reference).
"""
import torch
from torch import nn

from . import utils

# name -> (block counts per stage, bottleneck?)
RESNET_SPECS = {
    "ResNet18": ((2, 2, 2, 2), False),
    "ResNet34": ((3, 4, 6, 3), False),
    "ResNet50": ((3, 4, 6, 3), True),
    "ResNet101": ((3, 4, 23, 3), True),
    "ResNet152": ((3, 8, 36, 3), True),
}


class FrozenBatchNorm2d(nn.Module):
    """Four fixed per-channel buffers; y = x*scale + shift with
    scale = w*rsqrt(rv+1e-5), shift = b - rm*scale (reference models/backbone.py:41-51).
    The native library folds scale into the conv weights and shift into the epilogue."""
    EPS = 1e-5

    def __init__(self, n):
        super().__init__()
        self.register_buffer("weight", torch.ones(n))
        self.register_buffer("bias", torch.zeros(n))
        self.register_buffer("running_mean", torch.zeros(n))
        self.register_buffer("running_var", torch.ones(n))

    def _load_from_state_dict(self, state_dict, prefix, *rest):
        state_dict.pop(prefix + "num_batches_tracked", None)   # torchvision checkpoints carry it
        super()._load_from_state_dict(state_dict, prefix, *rest)

    def scale_shift(self):
        scale = self.weight * torch.rsqrt(self.running_var + self.EPS)
        return scale, self.bias - self.running_mean * scale


def _conv(cin, cout, k, stride=1, dilation=1):
    pad = dilation * (k // 2)
    return nn.Conv2d(cin, cout, k, stride=stride, padding=pad, dilation=dilation, bias=False)


class Bottleneck(nn.Module):
    """1x1 -> 3x3 (carries stride/dilation) -> 1x1(x4) + identity/downsample, ReLU after the add
    (torchvision/models/resnet.py:108-165)."""
    expansion = 4

    def __init__(self, inplanes, planes, stride, dilation, downsample):
        super().__init__()
        self.conv1 = _conv(inplanes, planes, 1)
        self.bn1 = FrozenBatchNorm2d(planes)
        self.conv2 = _conv(planes, planes, 3, stride, dilation)
        self.bn2 = FrozenBatchNorm2d(planes)
        self.conv3 = _conv(planes, planes * 4, 1)
        self.bn3 = FrozenBatchNorm2d(planes * 4)
        self.downsample = downsample


class BasicBlock(nn.Module):
    """3x3 (carries the stride) -> 3x3 + identity/downsample, ReLU after the add (torchvision/models/resnet.py:59-105;
    ResNet-18/34).  torchvision refuses dilation > 1 here, so these backbones need ``config.dilation = False``."""
    expansion = 1

    def __init__(self, inplanes, planes, stride, dilation, downsample):
        super().__init__()
        if dilation > 1:
            raise NotImplementedError("Dilation > 1 not supported in BasicBlock")     # torchvision's own error
        self.conv1 = _conv(inplanes, planes, 3, stride)
        self.bn1 = FrozenBatchNorm2d(planes)
        self.conv2 = _conv(planes, planes, 3)
        self.bn2 = FrozenBatchNorm2d(planes)
        self.downsample = downsample


class ResNetBody(nn.Module):
    """conv1/bn1/maxpool + layer1..layer4; what IntermediateLayerGetter keeps of torchvision's
    ResNet when only layer4 is returned (reference models/backbone.py:61-65)."""

    def __init__(self, blocks, dilate_last, bottleneck=True):
        super().__init__()
        Block = Bottleneck if bottleneck else BasicBlock
        ex = Block.expansion
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)
        self.bn1 = FrozenBatchNorm2d(64)
        inplanes, dilation = 64, 1
        for i, (planes, n) in enumerate(zip((64, 128, 256, 512), blocks)):
            stride = 1 if i == 0 else 2
            prev_dilation = dilation
            if i == 3 and dilate_last:          # replace_stride_with_dilation=[False, False, True]
                dilation *= stride
                stride = 1
            down = None
            if stride != 1 or inplanes != planes * ex:
                down = nn.Sequential(_conv(inplanes, planes * ex, 1, stride), FrozenBatchNorm2d(planes * ex))
            stage = [Block(inplanes, planes, stride, prev_dilation, down)]
            inplanes = planes * ex
            stage += [Block(inplanes, planes, 1, dilation, None) for _ in range(1, n)]
            setattr(self, f"layer{i + 1}", nn.Sequential(*stage))
        nn.Linear(inplanes, 1000)               # torchvision's classifier head: consumes RNG, then dropped
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def stages(self):
        return [getattr(self, f"layer{i}") for i in range(1, 5)]


class Backbone(nn.Module):
    """ResNet with frozen BatchNorm (reference models/backbone.py:54-93).  conv1+layer1 are
    always frozen; layer2-4 train only when ``train_backbone``."""

    def __init__(self, name: str, train_backbone: bool, return_interm_layers: bool, dilation: bool):
        super().__init__()
        if return_interm_layers:
            raise NotImplementedError("only the layer4 feature map is on the hot path")
        if name not in RESNET_SPECS:
            raise NotImplementedError(f"backbone {name!r}: one of {sorted(RESNET_SPECS)}")
        if utils.is_main_process() and not getattr(Backbone, "_warned", False):
            # reference downloads ImageNet weights here (models/backbone.py:87-88); no network
            # on this box, so random init + load_state_dict is the only route.
            Backbone._warned = True
        blocks, bottleneck = RESNET_SPECS[name]
        self.body = ResNetBody(blocks, dilation, bottleneck)
        self.num_channels = 2048 if bottleneck else 512        # reference models/backbone.py:92
        self.arch = name
        self.dilation = bool(dilation)
        for pname, p in self.body.named_parameters():
            if not train_backbone or not any(f"layer{i}" in pname for i in (2, 3, 4)):
                p.requires_grad_(False)


def build_backbone(config):
    return Backbone(config.backbone, config.lr_backbone > 0, False, config.dilation)
