"""Multi-GPU parity of the data-parallel GAN step (run under torchrun, one rank per GPU):
every rank steps on its shard with NCCL gradient all-reduce; rank 0 then replays all shards on private single-GPU
replicas, averages the flat gradients by hand and applies the same Adam - parameters must agree.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import models
from oracle import oracle as O
from srb200.train_step import GANTrainStep

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B, NB = 4, 2


def build(pg):
    G = models.GeneratorRRDB(1, filters=64, num_res_blocks=NB, num_upsample=2, res_scale=0.1)
    G.load_state_dict(O.make_generator_sd(5, num_res_blocks=NB, num_upsample=2)); G = G.to(dev).train()
    Ds = []
    for k in range(2):
        d = models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64]); d.load_state_dict(O.make_discriminator_sd(6 + k, gain=2.0))
        Ds.append(d.to(dev))
    return G, Ds, GANTrainStep(G, Ds, factor=4, process_group=pg)


def data(r):
    hr, lr = O.synthetic_jets(B, 80, 4, 50 + r)
    eps = [torch.rand(B, 1, 1, 1, generator=torch.Generator().manual_seed(60 + r + 10 * k)) for k in range(2)]
    return hr.to(dev), lr.to(dev), eps


G, Ds, tr = build(dist.group.WORLD)
hr, lr, eps = data(rank)
g_first = {}
for it in range(2):
    losses = tr.step(lr, hr, epsilons=eps)
    if it == 0:     # flat_g after a DP step = the all-reduced SUM of the ranks' gradients (Adam applies 1 / world)
        g_first = {k: o.flat_g.clone() for k, o in (("G", tr.opt_G), ("D0", tr.opt_D[0]), ("D1", tr.opt_D[1]))}
torch.cuda.synchronize()
# all ranks hold identical parameters
for name, flat in (("G", tr.opt_G.flat_p), ("D0", tr.opt_D[0].flat_p), ("D1", tr.opt_D[1].flat_p)):
    ref = flat.clone(); dist.broadcast(ref, 0)
    assert torch.equal(ref, flat), f"rank {rank}: {name} parameters diverged from rank 0"
if rank == 0:
    # replay: one replica per shard (no process group), gradients averaged by hand.  Each replica is stepped with a
    # private trainer whose Adam is intercepted: collect flat_g, then apply the averaged gradient to all replicas.
    reps = [build(None) for _ in range(world)]
    shards = [data(r) for r in range(world)]
    for it in range(2):
        saved = []
        for (Gr, Dr, t), (h, l, e) in zip(reps, shards):
            snap = {}
            for key, opt in (("G", t.opt_G), ("D0", t.opt_D[0]), ("D1", t.opt_D[1])):
                orig = opt.step
                def fake(gate=None, grad_scale=1.0, gathered=False, _o=opt, _k=key):
                    snap[_k] = (_o.flat_g.clone(), gate[0].clone() if gate is not None else None)
                opt.step = fake
            t.step(l, h, epsilons=e)
            saved.append(snap)
        from srb200.optim import FusedAdam
        if it == 0:
            # step 0 starts from identical parameters: NCCL's sum and the hand sum differ only by fp32 summation order
            # (at 2 ranks not at all; rings / trees / NVLS at 8 ranks add in another order)
            for key in ("G", "D0", "D1"):
                hand = sum(s[key][0] for s in saved)
                d, m = (g_first[key] - hand).abs().max().item(), hand.abs().max().item()
                print(f"{key}: step-0 gradient  max |allreduce - hand sum| = {d:.3e}  (|g| max {m:.3e})", flush=True)
                assert d <= 1e-5 * m, key
        for (Gr, Dr, t) in reps:
            for key, opt in (("G", t.opt_G), ("D0", t.opt_D[0]), ("D1", t.opt_D[1])):
                opt.step = FusedAdam.step.__get__(opt)
                opt.flat_g.copy_(sum(s[key][0] for s in saved))
                gate = None
                if saved[0][key][1] is not None:
                    gate = (sum(s[key][1] for s in saved) / world, t.d_threshold)
                opt.step(gate=gate, grad_scale=1.0 / world, gathered=True)
    torch.cuda.synchronize()
    # Adam's first steps move every element by ~lr * g / (|g| + eps): where |g| is at the fp32 noise of the sum, another
    # summation order can flip that element's whole step (2e-4).  So: all but 1e-4 of the elements within 2e-5, and no
    # element further than the two steps' bound 2 * 2 * lr.
    for key, a, b in (("G", tr.opt_G.flat_p, reps[0][2].opt_G.flat_p), ("D0", tr.opt_D[0].flat_p, reps[0][2].opt_D[0].flat_p),
                      ("D1", tr.opt_D[1].flat_p, reps[0][2].opt_D[1].flat_p)):
        diff = (a - b).abs()
        d, frac = diff.max().item(), (diff > 2e-5).float().mean().item()
        print(f"{key}: max |dp - replay| = {d:.3e}, fraction beyond 2e-5 = {frac:.2e}  (|p| max {a.abs().max().item():.3e})",
              flush=True)
        assert d <= 8.1e-4 and frac <= 1e-4, key
    print("dp_check ok", {k: float(v) for k, v in losses.items()}, flush=True)
dist.barrier()
dist.destroy_process_group()
