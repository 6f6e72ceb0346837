"""Drop-in for the reference's ``lib/model.py``: same names, constructor arguments, attributes
and state_dict keys; the matching head runs on this library's CUDA kernels, the feature trunk
stays on stock PyTorch/cuDNN (BASELINE.json north_star)."""
from __future__ import annotations

import os
from collections import OrderedDict

import torch
import torch.nn as nn

from . import functional as F_
from .conv4d import Conv4d


def featureL2Norm(feature):
    """x / sqrt(sum_c x^2 + 1e-6)  (reference lib/model.py:14-17)."""
    return F_.feature_l2norm(feature)


class FeatureExtraction(torch.nn.Module):
    """Feature trunk, stock torchvision (reference lib/model.py:19-87).  Out of the hot path; kept
    so ``ImMatchNet`` is self-contained.  Only its trailing L2 normalisation runs on our kernel."""

    def __init__(self, train_fe=False, feature_extraction_cnn='resnet101', feature_extraction_model_file='',
                 normalization=True, last_layer='', use_cuda=True, pretrained=None):
        """``pretrained`` (not in the reference): None = ImageNet weights as the reference's ``pretrained=True``
        (lib/model.py:39: torchvision's IMAGENET1K_V1 set) unless NCNET_TRUNK_WEIGHTS=random is set; False = random
        initialisation (what ``ImMatchNet`` passes when a checkpoint will overwrite the trunk anyway).  A failed
        download raises: a silently random trunk trains and evaluates to garbage."""
        super().__init__()
        import torchvision.models as models
        self.normalization = normalization
        self.feature_extraction_cnn = feature_extraction_cnn
        if pretrained is None:
            pretrained = os.environ.get("NCNET_TRUNK_WEIGHTS", "imagenet").lower() != "random"

        def make(name):
            return getattr(models, name)(weights="IMAGENET1K_V1" if pretrained else None)

        if feature_extraction_cnn == 'vgg':
            names = ['conv1_1', 'relu1_1', 'conv1_2', 'relu1_2', 'pool1', 'conv2_1', 'relu2_1', 'conv2_2', 'relu2_2',
                     'pool2', 'conv3_1', 'relu3_1', 'conv3_2', 'relu3_2', 'conv3_3', 'relu3_3', 'pool3', 'conv4_1',
                     'relu4_1', 'conv4_2', 'relu4_2', 'conv4_3', 'relu4_3', 'pool4', 'conv5_1', 'relu5_1', 'conv5_2',
                     'relu5_2', 'conv5_3', 'relu5_3', 'pool5']
            cut = names.index(last_layer or 'pool4')
            self.model = nn.Sequential(*list(make('vgg16').features.children())[:cut + 1])
        elif feature_extraction_cnn == 'resnet101':
            names = ['conv1', 'bn1', 'relu', 'maxpool', 'layer1', 'layer2', 'layer3', 'layer4']
            net = make('resnet101')
            cut = names.index(last_layer or 'layer3')
            self.model = nn.Sequential(*[getattr(net, n) for n in names[:cut + 1]])
        elif feature_extraction_cnn == 'densenet201':
            self.model = nn.Sequential(*list(make('densenet201').features.children())[:-4])
        else:
            raise ValueError(f"FeatureExtraction: unsupported trunk {feature_extraction_cnn!r}")
        if not train_fe:
            for param in self.model.parameters():
                param.requires_grad = False
        if use_cuda:
            self.model = self.model.cuda()

    def forward(self, image_batch):
        features = self.model(image_batch)
        if self.normalization:
            if torch.is_grad_enabled() and features.requires_grad:      # fine-tuning the trunk: keep autograd
                features = features / (features.pow(2).sum(1, keepdim=True) + 1e-6).pow(0.5)
            else:
                features = featureL2Norm(features)
        return features


class FeatureCorrelation(torch.nn.Module):
    """All-pairs feature correlation (reference lib/model.py:89-120)."""

    def __init__(self, shape='3D', normalization=True):
        super().__init__()
        self.normalization = normalization
        self.shape = shape
        self.ReLU = nn.ReLU()

    def forward(self, feature_A, feature_B):
        return F_.feature_correlation(feature_A, feature_B, shape=self.shape, normalization=self.normalization)


class NeighConsensus(torch.nn.Module):
    """Stack of Conv4d+ReLU layers applied to the volume and its A<->B transpose
    (reference lib/model.py:122-153).  ``self.conv`` keeps the reference's Sequential structure
    (Conv4d at even indices, ReLU at odd ones) so checkpoints load unchanged; ``forward`` runs the
    whole stack through the fused kernels instead of calling the Sequential."""

    def __init__(self, use_cuda=True, kernel_sizes=[3, 3, 3], channels=[10, 10, 1], symmetric_mode=True):
        super().__init__()
        self.symmetric_mode = symmetric_mode
        self.kernel_sizes = kernel_sizes
        self.channels = channels
        mods = []
        ch_in = 1
        for k_size, ch_out in zip(kernel_sizes, channels):
            mods.append(Conv4d(in_channels=ch_in, out_channels=ch_out, kernel_size=k_size, bias=True))
            mods.append(nn.ReLU(inplace=True))
            ch_in = ch_out
        self.conv = nn.Sequential(*mods)
        if use_cuda:
            self.conv.cuda()

    def _params(self):
        layers = [m for m in self.conv if isinstance(m, Conv4d)]
        return [m.weight for m in layers], [m.bias for m in layers]

    def forward_branches(self, x):
        """[B, G*C_last, ...] with G=2 un-summed branch outputs when symmetric (the following
        MutualMatching can sum them while it reads)."""
        ws, bs = self._params()
        return F_.neigh_consensus_branches(x, ws, bs, symmetric=self.symmetric_mode)

    def forward(self, x):
        y = self.forward_branches(x)
        if self.symmetric_mode:
            c = y.shape[1] // 2
            y = y[:, :c] + y[:, c:]
        return y


def MutualMatching(corr4d):
    """Mutual nearest-neighbour soft filtering (reference lib/model.py:155-175)."""
    return F_.mutual_matching(corr4d)


def maxpool4d(corr4d_hres, k_size=4):
    """4D max-pool with window-relative arg-max offsets (reference lib/model.py:177-191)."""
    return F_.maxpool4d(corr4d_hres, k_size=k_size)


class ImMatchNet(nn.Module):
    """Reference lib/model.py:193-282 with the head on this library's kernels."""

    def __init__(self,
                 feature_extraction_cnn='resnet101',
                 feature_extraction_last_layer='',
                 feature_extraction_model_file=None,
                 return_correlation=False,
                 ncons_kernel_sizes=[3, 3, 3],
                 ncons_channels=[10, 10, 1],
                 normalize_features=True,
                 train_fe=False,
                 use_cuda=True,
                 relocalization_k_size=0,
                 half_precision=False,
                 checkpoint=None,
                 ):
        super().__init__()
        if checkpoint is not None and checkpoint != '':
            print('Loading checkpoint...')
            checkpoint = torch.load(checkpoint, map_location=lambda storage, loc: storage, weights_only=False)
            checkpoint['state_dict'] = OrderedDict(
                [(k.replace('vgg', 'model'), v) for k, v in checkpoint['state_dict'].items()])
            print('Using checkpoint parameters: ')
            ncons_channels = checkpoint['args'].ncons_channels
            print('  ncons_channels: ' + str(ncons_channels))
            ncons_kernel_sizes = checkpoint['args'].ncons_kernel_sizes
            print('  ncons_kernel_sizes: ' + str(ncons_kernel_sizes))

        self.use_cuda = use_cuda
        self.normalize_features = normalize_features
        self.return_correlation = return_correlation
        self.relocalization_k_size = relocalization_k_size
        self.half_precision = half_precision

        self.FeatureExtraction = FeatureExtraction(train_fe=train_fe,
                                                   feature_extraction_cnn=feature_extraction_cnn,
                                                   feature_extraction_model_file=feature_extraction_model_file,
                                                   last_layer=feature_extraction_last_layer,
                                                   normalization=normalize_features,
                                                   use_cuda=self.use_cuda,
                                                   pretrained=False if (checkpoint is not None and checkpoint != '') else None)
        self.FeatureCorrelation = FeatureCorrelation(shape='4D', normalization=False)
        self.NeighConsensus = NeighConsensus(use_cuda=self.use_cuda, kernel_sizes=ncons_kernel_sizes,
                                             channels=ncons_channels)

        if checkpoint is not None and checkpoint != '':
            print('Copying weights...')
            for name, param in self.FeatureExtraction.state_dict().items():
                if 'num_batches_tracked' not in name:
                    self.FeatureExtraction.state_dict()[name].copy_(checkpoint['state_dict']['FeatureExtraction.' + name])
            for name, param in self.NeighConsensus.state_dict().items():
                self.NeighConsensus.state_dict()[name].copy_(checkpoint['state_dict']['NeighConsensus.' + name])
            print('Done!')

        self.FeatureExtraction.eval()

        if self.half_precision:
            for p in self.NeighConsensus.parameters():
                p.data = p.data.half()
            for l in self.NeighConsensus.conv:
                if isinstance(l, Conv4d):
                    l.use_half = True

    def match_features(self, feature_A, feature_B):
        """The hot path: correlation -> [maxpool4d] -> MutualMatching -> NeighConsensus ->
        MutualMatching (reference lib/model.py:269-282)."""
        delta4d = None
        fc = self.FeatureCorrelation
        if (self.relocalization_k_size > 1 and fc.shape == '4D'
                and F_.can_fuse_correlation_pool(feature_A, feature_B, self.relocalization_k_size)):
            # correlation and relocalisation pooling in one kernel: the un-pooled volume is never written
            corr4d, max_i, max_j, max_k, max_l = F_.feature_correlation_pooled(
                feature_A, feature_B, self.relocalization_k_size, normalization=fc.normalization)
            delta4d = (max_i, max_j, max_k, max_l)
        else:
            corr4d = fc(feature_A, feature_B)
            if self.relocalization_k_size > 1:
                corr4d, max_i, max_j, max_k, max_l = maxpool4d(corr4d, k_size=self.relocalization_k_size)
                delta4d = (max_i, max_j, max_k, max_l)
        corr4d = MutualMatching(corr4d)
        nc = self.NeighConsensus
        if nc.symmetric_mode and nc.channels[-1] == 1:
            # the second MutualMatching sums the two symmetric branches while reading them
            corr4d = F_.mutual_matching(nc.forward_branches(corr4d), sum_branches=True)
        else:
            corr4d = MutualMatching(nc(corr4d))
        if delta4d is not None:
            return (corr4d, delta4d)
        return corr4d

    def forward(self, tnf_batch):
        feature_A = self.FeatureExtraction(tnf_batch['source_image'])
        feature_B = self.FeatureExtraction(tnf_batch['target_image'])
        if self.half_precision:
            feature_A = feature_A.half()
            feature_B = feature_B.half()
        return self.match_features(feature_A, feature_B)
