"""Host-side mirror of cvpce/models/classification.py for the hot path (same names, arguments and results).

The conv trunks stay torch modules (not rebuilt); pooling + concat + L2 (classification.py:46-51, 79-85)
and the nearest-neighbour search (classification.py:87-95) run in libcvpce_b200.so.
"""
import torch
from torch import nn
from torchvision.models import resnet, vgg

from . import ops

# MACVGG.forward normalisation constants (classification.py:41-44): ImageNet statistics re-expressed for [-1,1] input
NORM_MEAN = [0.485 * 2 - 1, 0.456 * 2 - 1, 0.406 * 2 - 1]
NORM_STD = [0.229 * 2, 0.224 * 2, 0.225 * 2]


def tanh_normalize_affine():
    """Per-channel (a,b) with a*x+b == normalize(scale_to_tanh(x)) for x in [0,1]: folds utils.py:280-281 and
    classification.py:41-44 into the crop kernel's affine."""
    out = []
    for m, s in zip(NORM_MEAN, NORM_STD):
        out += [2.0 / s, (-1.0 - m) / s]
    return out


class MACVGG(nn.Module):
    """classification.py:20-51. VGG16 split after relu4_3 / relu5_3; MAC descriptor of both, concatenated, L2-normalised."""

    embedding_size = 512 * 2

    def __init__(self, config='D', convs_per_block=(2, 2, 3, 3, 3), batch_norm=True, vgg_state_dict=None):
        super().__init__()
        source = vgg.VGG(vgg.make_layers(vgg.cfgs[config], batch_norm=batch_norm), init_weights=vgg_state_dict is None)
        if vgg_state_dict is not None:
            source.load_state_dict(vgg_state_dict)
        per_conv = 3 if batch_norm else 2
        block_lengths = [n * per_conv + 1 for n in convs_per_block]  # + max pool
        split_1 = sum(block_lengths[:-1]) - 1  # through the last ReLU of block 4
        split_2 = sum(block_lengths) - 1       # through the last ReLU of block 5 (no pool5)
        self.block1 = source.features[:split_1]
        self.block2 = source.features[split_1:split_2]
        self.register_buffer('_mean', torch.tensor(NORM_MEAN).view(1, 3, 1, 1), persistent=False)
        self.register_buffer('_std', torch.tensor(NORM_STD).view(1, 3, 1, 1), persistent=False)
        self.regions_levels = 0  # 0: MAC (reference); L>0: R-MAC extension with L scales
        self.fast_trunk = False
        self.own_launches = 0    # launches of library kernels inside the trunk (max-pool), for bench.py's gpu_launches
        self.stage_timer = None  # optional pipeline.StageTimer: records the max-pool launches like the other kernels

    def enable_fast_trunk(self):
        """Drive the (unchanged, not rebuilt) torch trunk the fast way on the GPU: channels-last weights and activations
        (no NCHW<->NHWC transposes around every cuDNN convolution), torch.cudnn_convolution_relu (conv + bias + ReLU in
        one cuDNN call instead of three kernels) and the library's streaming 2x2 max-pool kernel between the convolutions
        (cvpce_maxpool2x2_nhwc instead of at::max_pool_forward_nhwc).  Same parameters, same fp32/TF32 arithmetic as the plain modules
        (measured on B200: 5.2k -> 13.3k crops/s at 256x256, relu5_3 equal to the NCHW run within 3e-4 relative)."""
        self.to(memory_format=torch.channels_last)
        self.fast_trunk = True
        return self

    accepts_padded_input = True  # fast trunk: [B,4,S,S] channels-last crops whose 4th channel is zero (ops.crop_resize pad_channel)

    def _padded_weight(self, conv):
        """conv.weight with one zero input channel appended (channels-last), cached per (storage, version, dtype): cuDNN then
        runs the 3-channel first convolution of VGG16 without its channel-padding pass (nhwcAddPaddingKernel) and on the
        aligned kernel -- 1.51 -> 1.21 ms per 256 crops, the same output bit for bit (the extra products are exact zeros)."""
        w = conv.weight
        key = (w.data_ptr(), w._version, w.dtype, w.device)
        cache = self.__dict__.setdefault("_wpad_cache", {})
        if cache.get("key") != key:
            wp = torch.zeros((w.shape[0], w.shape[1] + 1, w.shape[2], w.shape[3]), dtype=w.dtype, device=w.device)
            wp[:, :w.shape[1]] = w.detach()
            cache["key"], cache["w"] = key, wp.contiguous(memory_format=torch.channels_last)
        return cache["w"]

    def _run_fused(self, layers, x):
        layers = list(layers)
        i = 0
        while i < len(layers):
            l = layers[i]
            if (isinstance(l, nn.Conv2d) and i + 1 < len(layers) and isinstance(layers[i + 1], nn.ReLU) and l.bias is not None
                    and l.padding_mode == 'zeros'):
                w = l.weight
                if x.shape[1] == w.shape[1] + 1 and l.groups == 1:
                    w = self._padded_weight(l)  # 4-channel crops (zero 4th channel) against a zero-padded weight
                x = torch.cudnn_convolution_relu(x, w, l.bias, l.stride, l.padding, l.dilation, l.groups)
                i += 2
            elif (isinstance(l, nn.MaxPool2d) and l.kernel_size in (2, (2, 2)) and l.stride in (2, (2, 2)) and l.padding in (0, (0, 0))
                  and l.dilation in (1, (1, 1)) and not l.ceil_mode and not l.return_indices and ops.maxpool2x2_supported(x)):
                # at::max_pool_forward_nhwc was 28 % of the step; same values bit for bit
                if self.stage_timer is not None:
                    with self.stage_timer.span("maxpool2x2", 1.25 * x.numel() * x.element_size()):
                        x = ops.maxpool2x2(x)
                else:
                    x = ops.maxpool2x2(x)
                self.own_launches += 1
                i += 1
            else:
                x = l(x)
                i += 1
        return x

    def features(self, x, normalized=False, pooled_f4=False):
        """(relu4_3, relu5_3) maps.  pooled_f4=True (fast trunk only): return pool4(relu4_3) in place of relu4_3 when the map
        has even sides -- the 2x2 windows then tile it exactly, so the max over the pooled map IS the max over the map
        (bit for bit) and the MAC kernel reads a quarter of the bytes of a tensor the trunk produces anyway."""
        if x.shape[1] == 4 and not (normalized and self.fast_trunk and x.is_cuda):
            raise RuntimeError("MACVGG.features: 4-channel (zero-padded) input is for the normalised fast-trunk path only")
        if not normalized:
            x = (x - self._mean.to(x.dtype)) / self._std.to(x.dtype)
        if self.fast_trunk and x.is_cuda:
            x = x.contiguous(memory_format=torch.channels_last)
            f4 = self._run_fused(self.block1, x)
            b2 = list(self.block2)
            if (pooled_f4 and isinstance(b2[0], nn.MaxPool2d) and b2[0].kernel_size in (2, (2, 2)) and b2[0].stride in (2, (2, 2))
                    and b2[0].padding in (0, (0, 0)) and not b2[0].ceil_mode and f4.shape[2] % 2 == 0 and f4.shape[3] % 2 == 0):
                p4 = self._run_fused(b2[:1], f4)
                return p4, self._run_fused(b2[1:], p4)
            f5 = self._run_fused(b2, f4)
            return f4, f5
        f4 = self.block1(x)
        f5 = self.block2(f4)
        return f4, f5

    def forward(self, x, eps=1e-8, normalized=False):
        f4, f5 = self.features(x, normalized, pooled_f4=self.regions_levels == 0)
        if f4.dtype not in (torch.float32, torch.bfloat16, torch.float16):
            f4, f5 = f4.float(), f5.float()
        if self.regions_levels > 0:
            regions = [ops.rmac_regions(f.shape[2], f.shape[3], self.regions_levels) for f in (f4, f5)]
            return ops.region_pool_embed([f4, f5], regions=regions, per_region_l2=True, final_l2=True, eps=eps)
        return ops.region_pool_embed([f4, f5], eps=eps)


class MACResNet(nn.Module):
    """classification.py:53-85."""
    layer_output_sizes = [64, 256, 512, 1024, 2048]

    def __init__(self, source_resnet, descriptor_layers=(2, 3)):
        super().__init__()
        stages = [nn.Sequential(source_resnet.conv1, source_resnet.bn1, source_resnet.relu, source_resnet.maxpool),
                  source_resnet.layer1, source_resnet.layer2, source_resnet.layer3, source_resnet.layer4]
        self.blocks = nn.ModuleList()
        start = 0
        for l in descriptor_layers:
            self.blocks.append(nn.Sequential(*stages[start:l + 1]))
            start = l + 1
        self.embedding_size = sum(self.layer_output_sizes[l] for l in descriptor_layers)

    def forward(self, x, eps=1e-8):
        feats = []
        for b in self.blocks:
            x = b(x)
            feats.append(x)
        return ops.region_pool_embed(feats, eps=eps)


# below this many similarity entries the launch-bound fp32 kernel is used; above it the tcgen05 kernel
_TENSOR_CORE_MIN_WORK = 1 << 16


def nearest_neighbors(anchors, queries, k=1, exact=True):
    """classification.py:90-95: indices int64[Q,k] of the k anchors with the smallest cosine distance to each query,
    ascending distance; ties by ascending anchor index (the reference's argsort leaves ties unspecified)."""
    return similarity_topk(anchors, queries, k, exact)[1]


def similarity_topk(anchors, queries, k=1, exact=True, prepared=None):
    """(cosine similarity f32[Q,k], indices i64[Q,k]). `prepared` = PreparedGallery for a gallery searched repeatedly."""
    Q, D = queries.shape
    G = anchors.shape[0] if prepared is None else prepared.G
    k_eff = min(k, G) if G > 0 else 0
    if Q == 0 or k_eff == 0:
        return (torch.empty((Q, k_eff), dtype=torch.float32, device=queries.device),
                torch.empty((Q, k_eff), dtype=torch.int64, device=queries.device))
    # the tcgen05 kernel keeps k + 8 <= 32 candidates per row for the exact re-score: k <= 24 (32 without the re-score)
    use_tc = (D % 64 == 0 and 64 <= D <= 4096 and k_eff <= (24 if exact else 32) and Q * G >= _TENSOR_CORE_MIN_WORK)
    if not use_tc:
        g = anchors if prepared is None else prepared.f32
        return ops.sim_topk_f32(queries, g, k_eff, normalize=True)
    if prepared is None:
        prepared = PreparedGallery(anchors)
    q32 = queries.contiguous().float()
    qb, qinv = ops.rows_prepare(q32, normalize=True)
    return ops.sim_topk(qb, prepared.bf16, k_eff, 0, q32 if exact else None, prepared.f32 if exact else None,
                        qinv if exact else None, prepared.inv_norm if exact else None)


class PreparedGallery:
    """One-time operand preparation of a gallery (Classifier.build_index time): fp32 rows, inverse norms, bf16 unit rows."""

    def __init__(self, embedding, idx_base=0):
        self.f32 = embedding.contiguous().float()
        self.G, self.D = self.f32.shape
        self.idx_base = idx_base
        if self.G > 0 and self.D % 64 == 0:
            self.bf16, self.inv_norm = ops.rows_prepare(self.f32, normalize=True)
        else:
            self.bf16, self.inv_norm = None, None


def macvgg_embedder(model='vgg16_bn', pretrained=True, progress=True):
    """classification.py:97-109."""
    configs = {'vgg16': ('D', False), 'vgg16_bn': ('D', True)}
    if model not in configs:
        raise NotImplementedError(f'MACVGG not implemented for {model}')
    state = None
    if pretrained:
        weights = {'vgg16': vgg.VGG16_Weights.IMAGENET1K_V1, 'vgg16_bn': vgg.VGG16_BN_Weights.IMAGENET1K_V1}[model]
        state = weights.get_state_dict(progress=progress)
    config, bn = configs[model]
    return MACVGG(config, vgg_state_dict=state, batch_norm=bn)


def macresnet_encoder(model='resnet50', pretrained=True, progress=True, batch_norm=True, desc_layers=(2, 3)):
    """classification.py:111-121 (rebuilt on the current torchvision constructor; the reference's call is broken there)."""
    if model != 'resnet50':
        raise NotImplementedError(f'MACResNet not implemented for {model}')
    norm_layer = nn.BatchNorm2d if batch_norm else nn.Identity
    weights = resnet.ResNet50_Weights.IMAGENET1K_V1 if pretrained else None
    return MACResNet(resnet.resnet50(weights=weights, progress=progress, norm_layer=norm_layer), desc_layers)
