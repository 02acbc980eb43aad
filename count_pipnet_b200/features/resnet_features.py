"""ResNet feature extractors with the reference's parameter names and stride choices
(features/resnet_features.py of TarasKutsyk/Count_PIPNet): conv1/bn1/maxpool, layer1..layer4 where layer3 and layer4
keep stride 1 (:153-154), no avgpool / fc. 224^2 input -> 28x28 map.

Composed PyTorch modules (autograd path); the eval-mode inference forward of the Bottleneck variants runs through the
fused sm_100a kernels in `count_pipnet_b200.engine` (BatchNorm folded into the GEMM epilogue).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.utils.model_zoo as model_zoo

model_urls = {
    'resnet18': 'https://download.pytorch.org/models/resnet18-5c106cde.pth',
    'resnet34': 'https://download.pytorch.org/models/resnet34-333f7ec4.pth',
    'resnet50': 'https://download.pytorch.org/models/resnet50-19c8e357.pth',
    'resnet101': 'https://download.pytorch.org/models/resnet101-5d3b4d8f.pth',
    'resnet152': 'https://download.pytorch.org/models/resnet152-b121ed2d.pth',
}
model_dir = './pretrained_models'


def conv3x3(in_planes, out_planes, stride=1):
    return nn.Conv2d(in_planes, out_planes, kernel_size=3, stride=stride, padding=1, bias=False)


def conv1x1(in_planes, out_planes, stride=1):
    return nn.Conv2d(in_planes, out_planes, kernel_size=1, stride=stride, bias=False)


class BasicBlock(nn.Module):
    """features/resnet_features.py:32-74."""
    expansion = 1
    num_layers = 2

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = conv3x3(inplanes, planes, stride)
        self.bn1 = nn.BatchNorm2d(planes)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = conv3x3(planes, planes)
        self.bn2 = nn.BatchNorm2d(planes)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return self.relu(out + identity)

    def block_conv_info(self):
        return [3, 3], [self.stride, 1], [1, 1]


class Bottleneck(nn.Module):
    """1x1 -> 3x3(stride) -> 1x1(x4), BN after each, residual add, ReLU (features/resnet_features.py:77-124)."""
    expansion = 4
    num_layers = 3

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = conv1x1(inplanes, planes)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = conv3x3(planes, planes, stride)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = conv1x1(planes, planes * self.expansion)
        self.bn3 = nn.BatchNorm2d(planes * self.expansion)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        return self.relu(out + identity)

    def block_conv_info(self):
        return [1, 3, 1], [1, self.stride, 1], [0, 1, 0]


class ResNet_features(nn.Module):
    """features/resnet_features.py:126-229."""

    def __init__(self, block, layers, num_classes=1000, zero_init_residual=False):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        self.kernel_sizes = [7, 3]
        self.strides = [2, 2]
        self.paddings = [3, 1]
        self.block = block
        self.layers = layers
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=1)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=1)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode='fan_out', nonlinearity='relu')
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)
        if zero_init_residual:
            for m in self.modules():
                if isinstance(m, Bottleneck):
                    nn.init.constant_(m.bn3.weight, 0)
                elif isinstance(m, BasicBlock):
                    nn.init.constant_(m.bn2.weight, 0)

    def _make_layer(self, block, planes, num_blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(conv1x1(self.inplanes, planes * block.expansion, stride),
                                       nn.BatchNorm2d(planes * block.expansion))
        blocks = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        blocks += [block(self.inplanes, planes) for _ in range(1, num_blocks)]
        for b in blocks:
            k, s, p = b.block_conv_info()
            self.kernel_sizes.extend(k)
            self.strides.extend(s)
            self.paddings.extend(p)
        return nn.Sequential(*blocks)

    def _composed(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        return self.layer4(self.layer3(self.layer2(self.layer1(x))))

    def forward(self, x):
        from .. import engine
        if engine.use_fused(self, x):
            return engine.backbone_forward_nchw(self, x)
        return self._composed(x)

    def conv_info(self):
        return self.kernel_sizes, self.strides, self.paddings

    def num_layers(self):
        return self.block.num_layers * sum(self.layers) + 1

    def __repr__(self):
        return 'resnet{}_features'.format(self.num_layers() + 1)


_ARCH = {
    'resnet18': (BasicBlock, [2, 2, 2, 2]),
    'resnet34': (BasicBlock, [3, 4, 6, 3]),
    'resnet50': (Bottleneck, [3, 4, 6, 3]),
    'resnet101': (Bottleneck, [3, 4, 23, 3]),
    'resnet152': (Bottleneck, [3, 8, 36, 3]),
}


def _imagenet_resnet(name: str, pretrained: bool, **kwargs) -> ResNet_features:
    """Factory body shared by resnet{18,34,50,101,152}_features (features/resnet_features.py:231-328): ImageNet
    weights come from the torchvision model zoo URL with the fc layer dropped."""
    block, layers = _ARCH[name]
    model = ResNet_features(block, layers, **kwargs)
    if pretrained:
        sd = model_zoo.load_url(model_urls[name], model_dir=model_dir)
        sd = {k: v for k, v in sd.items() if not k.startswith('fc.')}
        model.load_state_dict(sd, strict=False)
    return model


def resnet18_features(pretrained=False, **kwargs):
    return _imagenet_resnet('resnet18', pretrained, **kwargs)


def resnet34_features(pretrained=False, **kwargs):
    return _imagenet_resnet('resnet34', pretrained, **kwargs)


def resnet50_features(pretrained=False, **kwargs):
    return _imagenet_resnet('resnet50', pretrained, **kwargs)


def resnet101_features(pretrained=False, **kwargs):
    return _imagenet_resnet('resnet101', pretrained, **kwargs)


def resnet152_features(pretrained=False, **kwargs):
    return _imagenet_resnet('resnet152', pretrained, **kwargs)


def resnet50_features_inat(pretrained=False, **kwargs):
    """ResNet-50 with the BBN iNaturalist2017 checkpoint (features/resnet_features.py:274-302): the `cb_block`
    of that checkpoint becomes `layer4.2`, `rb_block` and the classifier are dropped."""
    import os
    model = ResNet_features(Bottleneck, [3, 4, 6, 3], **kwargs)
    if pretrained:
        path = os.path.join('features', 'state_dicts', 'BBN.iNaturalist2017.res50.180epoch.best_model.pth')
        if not os.path.exists(path):
            print("To use Resnet50 pretrained on iNaturalist, create features/state_dicts and download "
                  "BBN.iNaturalist2017.res50.180epoch.best_model.pth into it from "
                  "https://drive.google.com/drive/folders/1yHme1iFQy-Lz_11yZJPlNd9bO_YPKlEU.", flush=True)
        raw = torch.load(path)
        sd = {}
        for k, v in raw.items():
            if k.startswith('module.backbone.cb_block'):
                sd['layer4.2' + k.split('cb_block')[-1]] = v
            elif k.startswith('module.backbone.rb_block') or k.startswith('module.classifier'):
                continue
            elif k.startswith('module.backbone.'):
                sd[k.split('backbone.')[-1]] = v
            else:
                sd[k] = v
        model.load_state_dict(sd, strict=True)
    return model
