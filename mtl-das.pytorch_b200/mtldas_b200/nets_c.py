"""Boundary class for the reference's model C (Inception-V3 multi-classifier).

/root/reference/model/modelC_multiClassifier.py:28-172 builds torchvision's Inception-V3 with a 1-channel stem, 32
classes and no auxiliary head.  This class keeps that constructor signature, parameter tree and state_dict (566 keys,
torchvision naming) so that `utils.py` imports, checkpoints load with strict=True and
`--model multi_classifier` reaches the forward -- where it stops: the Inception kernels (1x7 / 7x1 / 5x5 / valid 3x3
convolutions, 3x3 average and max pools, the 2048->32 linear layer) are SURVEY.md section 8 row f1 ("next") and are not
built yet, and this package has no PyTorch fallback by design.
"""
import torch
import torch.nn as nn

from ._lib import MtldasError


class _ConvBN(nn.Module):
    """conv (no bias) + BatchNorm(eps=1e-3) parameter holder with the reference's key names (`conv`, `bn`)."""

    def __init__(self, cin, cout, **kw):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, bias=False, **kw)
        self.bn = nn.BatchNorm2d(cout, eps=0.001)


class Multi_Classifier(nn.Module):
    def __init__(self, num_classes: int = 32, aux_logits: bool = False, transform_input: bool = False,
                 inception_blocks=None, init_weights=True):
        super().__init__()
        from torchvision.models.inception import InceptionA, InceptionB, InceptionC, InceptionD, InceptionE, InceptionAux
        blocks = inception_blocks or [_ConvBN, InceptionA, InceptionB, InceptionC, InceptionD, InceptionE, InceptionAux]
        if len(blocks) != 7:
            raise AssertionError("inception_blocks must have 7 entries")
        cb, ia, ib, ic, idd, ie, iaux = blocks
        self.aux_logits, self.transform_input = aux_logits, transform_input
        self.Conv2d_1a_3x3 = cb(1, 32, kernel_size=3, stride=2)          # 1-channel stem (modelC:63)
        self.Conv2d_2a_3x3 = cb(32, 32, kernel_size=3)
        self.Conv2d_2b_3x3 = cb(32, 64, kernel_size=3, padding=1)
        self.maxpool1 = nn.MaxPool2d(kernel_size=3, stride=2)
        self.Conv2d_3b_1x1 = cb(64, 80, kernel_size=1)
        self.Conv2d_4a_3x3 = cb(80, 192, kernel_size=3)
        self.maxpool2 = nn.MaxPool2d(kernel_size=3, stride=2)
        self.Mixed_5b = ia(192, pool_features=32)
        self.Mixed_5c = ia(256, pool_features=64)
        self.Mixed_5d = ia(288, pool_features=64)
        self.Mixed_6a = ib(288)
        self.Mixed_6b = ic(768, channels_7x7=128)
        self.Mixed_6c = ic(768, channels_7x7=160)
        self.Mixed_6d = ic(768, channels_7x7=160)
        self.Mixed_6e = ic(768, channels_7x7=192)
        self.AuxLogits = iaux(768, num_classes) if aux_logits else None
        self.Mixed_7a = idd(768)
        self.Mixed_7b = ie(1280)
        self.Mixed_7c = ie(2048)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.dropout = nn.Dropout()
        self.fc = nn.Linear(2048, num_classes)
        if init_weights:
            self._truncnorm_init()

    def _truncnorm_init(self):
        """truncnorm(-2, 2) * stddev for conv / linear weights, BatchNorm (1, 0) -- what modelC:88-100 does, drawn with
        torch's generator instead of scipy's (the reference's init is not reproducible from a torch seed anyway)."""
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                std = float(getattr(m, "stddev", 0.1))
                nn.init.trunc_normal_(m.weight, mean=0.0, std=std, a=-2 * std, b=2 * std)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def forward(self, x):
        raise MtldasError("Multi_Classifier (model C, Inception-V3): the sm_100a kernels for this model are not built "
                          "yet (SURVEY.md section 8 row f1); there is no PyTorch fallback in this package")
