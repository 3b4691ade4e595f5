"""Drop-in replacement for the reference's ``models.py`` (lukas-blecher/super-resolution).

Put this directory first on ``sys.path`` and ``from models import *`` (esrgan.py:17, evaluation/eval.py:11,
evaluation/demo.py:17, datasets.py:12) resolves here.  Same exported names, constructor signatures, attribute
names and ``state_dict`` layout as the reference, so its checkpoints load with
``strict=True`` and vice versa.  The modules only *hold* parameters; every forward/backward of the hot path
(GeneratorRRDB, Markovian_Discriminator, SumPool2d) runs hand-written sm_100a kernels through libsrb200.so
(``srb200/``).  There is no CPU path: calling a hot-path module on CPU tensors raises RuntimeError("CUDA ...").
"""
import math
from itertools import product

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from srb200 import _lib as _L
from srb200 import ops as _ops
from srb200.generator import GeneratorEngine as _GeneratorEngine
from srb200 import discriminator as _disc
from srb200 import small_ops as _small


def _conv_block(in_features, filters, non_linearity=True):
    layers = [nn.Conv2d(in_features, filters, 3, 1, 1, bias=True)]
    if non_linearity:
        layers.append(nn.LeakyReLU())
    return nn.Sequential(*layers)


class DenseResidualBlock(nn.Module):
    """Five 3x3 convs over a growing concatenation, ``out5 * res_scale + x`` (reference models.py:9-41).
    Parameter names b{k}.0.{weight,bias}; ``blocks`` is a plain list as in the reference."""

    def __init__(self, filters, res_scale=0.2, drop_rate=0):
        super().__init__()
        self.res_scale = res_scale
        self.filters = filters
        for k in range(1, 6):
            setattr(self, "b%d" % k, _conv_block(k * filters, filters, non_linearity=k < 5))
        self.drop = drop_rate > 0
        if self.drop:
            self.drop1 = nn.Dropout2d(drop_rate)
        self.blocks = [self.b1, self.b2, self.b3, self.b4, self.b5]

    def forward(self, x):
        # stand-alone use (the generator drives its blocks through the fused engine instead)
        return _small.dense_block_forward(self, x)


class ResidualInResidualDenseBlock(nn.Module):
    """Three dense blocks and an outer scaled skip (reference models.py:44-53).  As in the reference the inner
    blocks keep their default res_scale of 0.2; only the outer one follows the argument."""

    def __init__(self, filters, res_scale=0.2, drop_rate=0):
        super().__init__()
        self.res_scale = res_scale
        self.dense_blocks = nn.Sequential(*[DenseResidualBlock(filters, drop_rate=drop_rate) for _ in range(3)])

    def forward(self, x):
        return self.dense_blocks(x).mul(self.res_scale) + x


class _GeneratorFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gen, need, x, *params):
        # grad mode is always off inside Function.forward, so the caller decides whether to keep activations
        ctx.gen = gen
        out, srs = gen._engine().forward(x, power=gen._host_scalar("power"), multiplier=gen._host_scalar("multiplier"),
                                         thres=float(gen.thres), training=gen.training, save_for_backward=need,
                                         drop_masks=gen._drop_masks)
        ctx.n = len(params)
        ctx.fwd_gen = gen._engine()._fwd_gen if need else None
        return out, srs

    @staticmethod
    def backward(ctx, g_out, g_srs):
        gen = ctx.gen
        grads = gen._engine().backward(g_out, g_srs, fwd_gen=ctx.fwd_gen)
        res = []
        for name, p in gen.named_parameters():
            g = grads.get(name) if p.requires_grad else None
            res.append(g)
        return (None, None, None) + tuple(res)


class GeneratorRRDB(nn.Module):
    """RRDB generator (reference models.py:56-135): conv1 -> RRDBs -> conv2 (+skip) -> upsampling -> conv3, with the
    power/multiplier pre- and post-scaling and the hardshrink output head.  ``self.srs`` is set by every forward."""

    def __init__(self, channels=1, filters=64, num_res_blocks=10, num_upsample=1, power=1, multiplier=1, drop_rate=0,
                 res_scale=0.2, use_transposed_conv=False, fully_tconv_upsample=False, num_final_layer_res=0,
                 uniform_init=False):
        super().__init__()
        self.num_final_layer_res = num_final_layer_res
        self.conv1 = nn.Conv2d(channels, filters, kernel_size=3, stride=1, padding=1)
        self.res_blocks = nn.Sequential(*[ResidualInResidualDenseBlock(filters, res_scale=res_scale, drop_rate=drop_rate)
                                          for _ in range(num_res_blocks)])
        self.conv2 = nn.Conv2d(filters, filters, kernel_size=3, stride=1, padding=1)
        up = []
        for ii in range(num_upsample):
            tconv = fully_tconv_upsample or (use_transposed_conv and ii % 2 == 1)
            if use_transposed_conv and fully_tconv_upsample:
                tconv = ii % 2 == 1          # the reference tests use_transposed_conv first (models.py:70-83)
            if tconv:
                up += [nn.ConvTranspose2d(filters, filters, kernel_size=2, stride=2, padding=0), nn.LeakyReLU()]
            else:
                up += [nn.Conv2d(filters, filters * 4, kernel_size=3, stride=1, padding=1), nn.LeakyReLU(),
                       nn.PixelShuffle(upscale_factor=2)]
        self.upsampling = nn.Sequential(*up)
        if num_final_layer_res > 0:
            self.res_blocks_final = nn.Sequential(*[ResidualInResidualDenseBlock(filters, res_scale=res_scale,
                                                                                 drop_rate=drop_rate)
                                                    for _ in range(num_final_layer_res)])
        self.conv3 = nn.Sequential(nn.Conv2d(filters, filters, kernel_size=3, stride=1, padding=1), nn.LeakyReLU(),
                                   nn.Conv2d(filters, channels, kernel_size=3, stride=1, padding=1))
        self.thres = 0
        self.power = nn.Parameter(torch.Tensor([power]), False)
        self.multiplier = nn.Parameter(torch.Tensor([multiplier]), False)
        stages = [("tconv" if isinstance(m, nn.ConvTranspose2d) else "ps", "upsampling.%d" % i)
                  for i, m in enumerate(self.upsampling) if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d))]
        self._cfg = dict(channels=channels, filters=filters, num_res_blocks=num_res_blocks, num_upsample=num_upsample,
                         res_scale=res_scale, up_stages=stages, num_final=num_final_layer_res, drop_rate=drop_rate)
        self._drop_masks = None      # tests: explicit Dropout2d factors {"main": [...], "final": [...]} for the next forwards
        self._eng = None
        self._scalars = {}
        if uniform_init:
            self.init_conv2d()

    def init_conv2d(self):
        for c in self.children():
            if isinstance(c, nn.Conv2d):
                nn.init.xavier_uniform_(c.weight)
                nn.init.constant_(c.bias, 0.)

    # -- engine plumbing ---------------------------------------------------------------------------------------
    def _engine(self):
        if self._eng is None:
            self._eng = _GeneratorEngine(dict(self.named_parameters()), **self._cfg)
        else:
            self._eng.p = dict(self.named_parameters())       # .to()/load_state_dict may have swapped storages
        return self._eng

    def _host_scalar(self, name):
        """power / multiplier live on the device as in the reference (models.py:102-103) but are mirrored on the
        host so a forward needs no .item() sync (the reference pays two per forward, models.py:116,133)."""
        p = getattr(self, name)
        key = (p._version, p.data_ptr())
        hit = self._scalars.get(name)
        if hit is None or hit[0] != key:
            hit = (key, float(p.detach().cpu().item()))
            self._scalars[name] = hit
        return hit[1]

    def out(self, x, pow=torch.ones(1)):
        if self.training:
            return F.hardshrink(x, lambd=self.thres ** pow.item())
        return F.hardshrink(F.relu(x), lambd=self.thres ** pow.item())

    def forward(self, x):
        params = list(self.parameters())
        need = torch.is_grad_enabled() and any(p.requires_grad for p in params)
        out, srs = _GeneratorFn.apply(self, need, x, *params)
        self.srs = srs
        return out


# ------------------------------------------------------------------------------------------------ discriminators
def discriminator_block(in_filters, out_filters, stride=(1, 2)):
    return [nn.Conv2d(in_filters, out_filters, kernel_size=3, stride=stride[0], padding=1),
            nn.LeakyReLU(0.2, inplace=True),
            nn.Conv2d(out_filters, out_filters, kernel_size=3, stride=stride[1], padding=1),
            nn.LeakyReLU(0.2, inplace=True)]


def _patch_layers(in_channels, channels, h, w):
    layers, cin = [], in_channels
    for c in channels:
        layers.extend(discriminator_block(cin, c))
        cin = c
        h, w = int(np.ceil(h / 2)), int(np.ceil(w / 2))
    layers.append(nn.Conv2d(cin, 1, kernel_size=3, stride=1, padding=1))
    return layers, h, w


class Markovian_Discriminator(nn.Module):
    """PatchGAN discriminator (reference models.py:149-174); forward ignores the condition argument."""

    def __init__(self, input_shape, channels=[16, 32, 32, 64]):
        super().__init__()
        self.channels = channels
        self.input_shape = input_shape
        c, h, w = input_shape
        layers, ph, pw = _patch_layers(c, channels, h, w)
        self.output_shape = (1, ph, pw)
        self.model = nn.Sequential(*layers)

    def forward(self, img, *args):
        return _disc.sequential_forward(self.model, img)


class Standard_Discriminator(Markovian_Discriminator):
    def __init__(self, input_shape, channels):
        super().__init__(input_shape, channels)
        self.model = self.model[:-1]
        self.fc = nn.Sequential(nn.Linear(self.channels[-1] * self.output_shape[-2] * self.output_shape[-1], 1024),
                                nn.ReLU(), nn.Linear(1024, 1))
        self.output_shape = (1,)

    def forward(self, img, *args):
        return self.fc(_disc.sequential_forward(self.model, img).view(img.shape[0], -1))


class _TwoBranch(nn.Module):
    def __init__(self, input_shape, channels, num_upsample):
        super().__init__()
        self.channels = channels
        self.input_shape = input_shape
        cin, h, w = input_shape
        hr, cl, end = [], [], []
        for i, c in enumerate(channels):
            if i < num_upsample:
                hr.extend(discriminator_block(cin, c))
                cl.extend(discriminator_block(cin, c, stride=(1, 1)))
            elif i == num_upsample:
                end.extend(discriminator_block(cin * 2, c))
            else:
                end.extend(discriminator_block(cin, c))
            cin = c
            h, w = int(np.ceil(h / 2)), int(np.ceil(w / 2))
        end.append(nn.Conv2d(cin, 1, kernel_size=3, stride=1, padding=1))
        self.output_shape = (1, h, w)
        self.model_hr = nn.Sequential(*hr)
        self.model_c = nn.Sequential(*cl)
        self.endmodel = nn.Sequential(*end)

    def _trunk(self, img, cond):
        a = _disc.sequential_forward(self.model_hr, img)
        b = _disc.sequential_forward(self.model_c, cond)
        return _disc.sequential_forward(self.endmodel, torch.cat([a, b], 1))


class Conditional_Discriminator(_TwoBranch):
    def __init__(self, input_shape, channels=[32, 64, 128, 256], num_upsample=3):
        super().__init__(input_shape, channels, num_upsample)

    def forward(self, img, cond):
        return self._trunk(img, cond)


class Wasserstein_Discriminator(_TwoBranch):
    def __init__(self, input_shape, channels=[32, 64, 128, 256], num_upsample=3):
        super().__init__(input_shape, channels, num_upsample)
        self.out_channels, self.out_height, self.out_width = self.output_shape
        self.fc = nn.Linear(self.out_channels * self.out_height * self.out_width, 1)

    def forward(self, img, cond):
        return self.fc(self._trunk(img, cond).view(-1, self.out_channels * self.out_height * self.out_width))


class Wasserstein_PatchDiscriminator(Markovian_Discriminator):
    def __init__(self, input_shape, channels=[16, 32, 32, 64]):
        super().__init__(input_shape, channels)
        self.out_channels, self.out_height, self.out_width = self.output_shape
        self.fc = nn.Linear(self.out_channels * self.out_height * self.out_width, 1)

    def forward(self, img, *args):
        out = _disc.sequential_forward(self.model, img)
        return self.fc(out.view(-1, self.out_channels * self.out_height * self.out_width))


class SumPool2d(torch.nn.Module):
    """2D sum pooling (reference models.py:297-305)."""

    def __init__(self, k=4, stride=None):
        super().__init__()
        self.pool = torch.nn.AvgPool2d(k, stride=stride)      # kept for attribute parity; not used in forward
        self.k = k
        self.kernel_size = k * k
        self._native = stride is None or stride == k

    def forward(self, x):
        if not self._native:
            raise RuntimeError("CUDA srb200: SumPool2d kernels cover stride == k")
        if not x.is_cuda and torch.cuda.is_available():
            # datasets.py:160-166,247 pool single CPU images inside Dataset.__getitem__: run the kernel on the device
            # and hand the result back on the caller's device (still no CPU arithmetic path)
            return _small.sum_pool(x.cuda(), self.k).to(x.device)
        return _small.sum_pool(x, self.k)


class DiffableHistogram(nn.Module):
    """Differentiable histogram (reference models.py:308-342); off the hot path (lambda_hist = 0 by default):
    plain tensor math on whatever device the input lives on."""

    def __init__(self, bins, min=0, max=1, sigma=25, batchwise=False):
        super().__init__()
        self.bins = torch.Tensor(bins)
        self.sigma = sigma
        self.batchwise = batchwise
        if type(bins) is int:
            self.delta = (max - min) / bins * torch.ones(bins - 1)
            self.centers = float(min) + self.delta * (torch.arange(bins - 1).float() + 0.5)
        else:
            self.delta = torch.Tensor([np.diff(bins)]).float()
            self.centers = self.bins[:-1] + .5 * self.delta

    def forward(self, x):
        batches = len(x) if (len(x.shape) == 4 and self.batchwise) else 1
        x = x.view(batches, -1)
        x = x[:, None, :] - self.centers[..., None]
        half = self.delta[..., None] / 2
        return (torch.sigmoid(self.sigma * (x + half)) - torch.sigmoid(self.sigma * (x - half))).sum(2)

    def to(self, device):
        self.centers = self.centers.to(device)
        self.delta = self.delta.to(device)
        return self


def normal(x, sig=1, mu=0):
    return np.exp(-.5 * ((x - mu) / sig) ** 2) / np.sqrt(2 * np.pi * sig * sig)


def naive_upsample(img):
    """Baseline that splits every non-zero LR pixel at random over its 2x2 HR block (reference models.py:349-363);
    a host-side Python loop in the reference too, kept for evaluation/eval.py's NaiveGenerator."""
    up = torch.zeros(len(img), 1, *tuple(np.array(img.shape[-2:]) * 2))
    for batch in range(len(img)):
        for i, j in product(range(img.shape[-2]), range(img.shape[-1])):
            e = img[batch, 0, i, j]
            if e == 0:
                continue
            rand = e * normal(torch.rand(4))
            new_vals = torch.zeros(1)
            while new_vals.sum() <= 0:
                new_vals = torch.where(torch.randint(0, 2, (4,)) == 1, torch.zeros(4), rand)
            for k, (a, b) in enumerate(product(torch.arange(2) + 2 * i, torch.arange(2) + 2 * j)):
                up[batch, 0, a, b] = new_vals[k] / new_vals.sum() * e
    return up


class NaiveGenerator(nn.Module):
    def __init__(self, num_upsample):
        super().__init__()
        self.num_upsample = num_upsample

    def forward(self, x):
        for _ in range(self.num_upsample):
            x = naive_upsample(x)
        return x


def weight_reset(m):
    """reference models.py:375-378"""
    if isinstance(m, nn.Conv2d) or isinstance(m, nn.Linear):
        m.reset_parameters()


def uniform_reset(m):
    """reference models.py:380-384"""
    if isinstance(m, nn.Conv2d):
        nn.init.xavier_uniform_(m.weight)
        nn.init.constant_(m.bias, 0.)
