"""GPU probe: how far apart are (a) the fp64 oracle, (b) the fp32 oracle, (c) the bf16-autocast oracle, (d) our fp32
kernels, (e) our bf16 kernels — logits rel-err and worst gradient rel-err / cosine.  Sets the noise floor that the
parity gates in tests/ are judged against."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "vision-transformer_b200")); sys.path.insert(0, ROOT)
import torch
from oracle import ref_vit
import train_dp

torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def cos(a, b):
    a, b = a.double().flatten().cpu(), b.double().flatten().cpu()
    return float(a @ b / (a.norm() * b.norm()).clamp_min(1e-300))


def run(E, H, L, B, init):
    P, img, classes = 16, 224, 2
    cfg = ref_vit.Cfg(3, classes, (img // P) ** 2, E, P, H, L, B)
    tr32 = train_dp.DPTrainer(embedding_size=E, num_heads=H, num_blocks=L, batch_size=B, dtype=torch.float32, dropout_p=0.0)
    trbf = train_dp.DPTrainer(embedding_size=E, num_heads=H, num_blocks=L, batch_size=B, dtype=torch.bfloat16, dropout_p=0.0)
    if init == "kaiming":
        tr32.init_parameters(seed=5)
        sd = {k: v.cpu() for k, v in tr32.state_dict().items()}
    else:
        sd = ref_vit.random_state_dict(cfg, seed=5)
        tr32.load_state_dict(sd)
    trbf.load_state_dict(sd)
    gen = torch.Generator().manual_seed(6)
    x = torch.rand(B, 3, img, img, generator=gen); y = torch.randint(0, classes, (B,), generator=gen)
    res = {}
    for name, dt in (("f64", torch.float64), ("f32", torch.float32)):
        s = {k: v.cuda().to(dt) for k, v in sd.items()}
        res[name] = ref_vit.loss_and_grads(s, x.cuda().to(dt), y.cuda(), cfg)
    s = {k: v.cuda() for k, v in sd.items()}
    with torch.autocast("cuda", dtype=torch.bfloat16):
        leaf = {k: v.detach().clone().requires_grad_(True) for k, v in s.items()}
        logits = ref_vit.vit_forward(leaf, x.cuda(), cfg)
        loss = torch.nn.functional.cross_entropy(logits.float(), y.cuda())
    loss.backward()
    res["autocast"] = (logits.detach().float(), loss.detach(), {k: v.grad for k, v in leaf.items()})
    for name, tr in (("ours32", tr32), ("oursbf", trbf)):
        d = tr._dims(False); tr.forward(x.cuda(), d); tr.backward(y.cuda(), d)
        res[name] = (tr.logits.clone(), tr.loss.clone(), {k: v.clone() for k, v in tr.grads_by_ref_key().items()})
    print(f"== E={E} H={H} L={L} B={B} init={init}  logits(f64)={res['f64'][0][0].tolist()}")
    for a, b in (("f32", "f64"), ("ours32", "f64"), ("ours32", "f32"), ("autocast", "f64"), ("oursbf", "f64"), ("oursbf", "autocast")):
        ga, gb = res[a][2], res[b][2]
        worst_rel = max((rel(ga[k], gb[k]), k) for k in gb)
        worst_cos = min((cos(ga[k], gb[k]), k) for k in gb)
        print(f"{a:>9} vs {b:<9} logits rel {rel(res[a][0], res[b][0]):.3e} | grad worst rel {worst_rel[0]:.3e} ({worst_rel[1][-40:]}) "
              f"| worst cos {worst_cos[0]:.6f} ({worst_cos[1][-40:]})")


if __name__ == "__main__":
    run(384, 6, 12, 8, "oracle")
    run(384, 6, 12, 8, "kaiming")
    run(768, 12, 2, 4, "oracle")
    run(768, 12, 12, 4, "kaiming")
