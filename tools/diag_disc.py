"""Per-layer diagnostics of the tensor-core discriminator engine against the fp32 autograd path."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import models
from oracle import oracle as O
from srb200 import disc_engine as DE, losses as LS, discriminator as Dm

dev = torch.device("cuda:0")
def cos(a, b):
    a, b = a.float().reshape(-1), b.float().reshape(-1)
    return (torch.dot(a, b) / (a.norm() * b.norm() + 1e-30)).item()
B = 6
d = models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64]); d.load_state_dict(O.make_discriminator_sd(71, gain=2.0)); d = d.to(dev)
eng = DE.DiscEngine(d.model, d.input_shape)
hr, _ = O.synthetic_jets(B, 80, 4, 72); fake, _ = O.synthetic_jets(B, 80, 4, 73); fake = fake * 0.7 + 0.01
hr, fake = hr.to(dev), fake.to(dev)
A = eng.forward([hr, fake], tag="rf")
both = torch.cat([hr, fake]).requires_grad_(True)
# reference activations layer by layer
h = both
acts = []
for conv, slope in eng.layers:
    h = Dm.conv2d(h, conv, False, slope)
    acts.append(h)
for l in range(eng.L - 1):
    a = A.a[l].float().permute(0, 3, 1, 2)
    print(f"act {l+1}: cos {cos(a, acts[l]):.6f} normratio {a.norm().item()/acts[l].norm().item():.4f} maxerr/range {(a-acts[l]).abs().max().item()/acts[l].abs().max().item():.4f}")
pred_ref = acts[-1]
print("pred maxerr/range", ((A.pred - pred_ref).abs().max() / pred_ref.abs().max()).item())
gp_ = torch.randn(2 * B, 1, 5, 5, generator=torch.Generator().manual_seed(5)).to(dev)
params = list(d.parameters())
ref_grads = torch.autograd.grad(pred_ref, params + [both] + acts[:-1], gp_)
grads = [(torch.zeros_like(params[2 * l]), torch.zeros_like(params[2 * l + 1])) for l in range(eng.L)]
eng.backward(A, gp_, grads, tag="bw")
ch = eng._chain_bufs(2 * B, dev, "bw")
na = len(params) + 1
for l in range(eng.L - 1):
    # ch[l] = gradient w.r.t. pre-activation of a_{l+1}; reference: grad wrt a_{l+1} times LReLU'
    gref = ref_grads[na + l] * torch.where(acts[l] > 0, 1.0, eng.slopes[l])
    g = ch[l].float().permute(0, 3, 1, 2)
    print(f"gz {l+1}: cos {cos(g, gref):.6f} normratio {g.norm().item()/gref.norm().item():.4f}")
for l in range(eng.L):
    print(f"dW {l}: cos {cos(grads[l][0], ref_grads[2*l]):.6f} normratio {grads[l][0].norm().item()/ref_grads[2*l].norm().item():.4f}   db cos {cos(grads[l][1], ref_grads[2*l+1]):.6f} ratio {grads[l][1].norm().item()/ref_grads[2*l+1].norm().item():.4f}")
# third opinion: plain torch (cuDNN) fp32
import torch.nn.functional as F
both2 = both.detach().clone().requires_grad_(True)
h = both2
for conv, slope in eng.layers:
    h = F.conv2d(h, conv.weight, conv.bias, conv.stride, 1)
    if slope is not None:
        h = F.leaky_relu(h, slope)
tg = torch.autograd.grad(h, params, gp_)
for l in range(eng.L):
    print(f"torch vs fp32-path dW {l}: normratio {ref_grads[2*l].norm().item()/tg[2*l].norm().item():.4f} cos {cos(ref_grads[2*l], tg[2*l]):.6f} | ours vs torch normratio {grads[l][0].norm().item()/tg[2*l].norm().item():.4f} cos {cos(grads[l][0], tg[2*l]):.6f} | db ours/torch {grads[l][1].norm().item()/tg[2*l+1].norm().item():.4f}")
for l in range(eng.L - 1):
    gref = ref_grads[na + l] * torch.where(acts[l] > 0, 1.0, eng.slopes[l])
    g = ch[l].float().permute(0, 3, 1, 2)
    s_ours, s_ref = g.sum((0, 2, 3)), gref.sum((0, 2, 3))
    print(f"sum gz{l+1}: torch-sum(ours)/ref {s_ours.norm().item()/s_ref.norm().item():.4f}  kernel db/torch-sum(ours) {grads[l][1].norm().item()/s_ours.norm().item():.4f}  flipfrac {((A.a[l].float().permute(0,3,1,2) > 0) != (acts[l] > 0)).float().mean().item():.5f}")
print("--- single-step check: torch dgrad from OUR gz_{l+1} and OUR mask vs kernel output")
for l in range(eng.L - 1, 0, -1):
    conv = eng.convs[l]
    s = conv.stride[0]
    gz_in = gp_ if l == eng.L - 1 else ch[l].float().permute(0, 3, 1, 2)
    a_in = A.a[l - 1].float().permute(0, 3, 1, 2)
    Hh, Ww = a_in.shape[2], a_in.shape[3]
    Ho, Wo = gz_in.shape[2], gz_in.shape[3]
    wq = conv.weight.detach().to(torch.bfloat16).float()
    gzq = gz_in.to(torch.bfloat16).float()
    ideal = F.conv_transpose2d(gzq, wq, None, s, 1, output_padding=(Hh - ((Ho - 1) * s + 1), Ww - ((Wo - 1) * s + 1)))
    ideal = ideal * torch.where(a_in > 0, 1.0, eng.slopes[l - 1])
    got = ch[l - 1].float().permute(0, 3, 1, 2)
    d = got - ideal
    print(f"layer {l} dgrad: maxerr/range {d.abs().max().item()/ideal.abs().max().item():.5f}  sum ratio {got.sum((0,2,3)).norm().item()/ideal.sum((0,2,3)).norm().item():.4f}  mean|d|/mean|ideal| {d.abs().mean().item()/ideal.abs().mean().item():.5f} bias {d.mean().item()/ideal.abs().mean().item():.6f}")
print("--- fp32 chain with OUR masks vs fp32 chain with fp32 masks vs ours (per-channel pixel sums)")
g_our_mask = gp_; g_ref_mask = gp_
for l in range(eng.L - 1, 0, -1):
    conv = eng.convs[l]; s = conv.stride[0]
    a_o = A.a[l - 1].float().permute(0, 3, 1, 2); a_r = acts[l - 1]
    Hh, Ww = a_r.shape[2], a_r.shape[3]; Ho, Wo = g_our_mask.shape[2], g_our_mask.shape[3]
    op = (Hh - ((Ho - 1) * s + 1), Ww - ((Wo - 1) * s + 1))
    g_our_mask = F.conv_transpose2d(g_our_mask, conv.weight.detach(), None, s, 1, output_padding=op) * torch.where(a_o > 0, 1.0, eng.slopes[l - 1])
    g_ref_mask = F.conv_transpose2d(g_ref_mask, conv.weight.detach(), None, s, 1, output_padding=op) * torch.where(a_r > 0, 1.0, eng.slopes[l - 1])
    got = ch[l - 1].float().permute(0, 3, 1, 2)
    sr = g_ref_mask.sum((0, 2, 3)).norm().item()
    print(f"gz{l}: sum(fp32 chain, our masks)/sum(ref) {g_our_mask.sum((0,2,3)).norm().item()/sr:.4f}   sum(ours)/sum(ref) {got.sum((0,2,3)).norm().item()/sr:.4f}  cancellation |sum|/sum|.| {g_ref_mask.sum().abs().item()/g_ref_mask.abs().sum().item():.5f}")
