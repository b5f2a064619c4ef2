"""GPU probe: benign (softened q/k) weights — gradient cosine of ours-bf16 and of the autocast oracle vs the fp32 oracle,
per tensor group, to find where bf16 error enters."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "vision-transformer_b200")); sys.path.insert(0, ROOT)
import torch
from oracle import ref_vit
import train_dp

torch.backends.cuda.matmul.allow_tf32 = False


def cos(a, b):
    a, b = a.double().flatten().cpu(), b.double().flatten().cpu()
    return float(a @ b / (a.norm() * b.norm()).clamp_min(1e-300))


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def run(E, H, L, B, soften, img=224, P=16):
    classes = 2
    cfg = ref_vit.Cfg(3, classes, (img // P) ** 2, E, P, H, L, B)
    sd = ref_vit.random_state_dict(cfg, seed=3)
    sd = {k: (v * soften if k.endswith(("key.weight", "query.weight")) else v) for k, v in sd.items()}
    trbf = train_dp.DPTrainer(embedding_size=E, num_heads=H, num_blocks=L, batch_size=B, dtype=torch.bfloat16,
                              dropout_p=0.0, image_size=img, patch_size=P)
    trbf.load_state_dict(sd)
    gen = torch.Generator().manual_seed(4)
    x = torch.rand(B, 3, img, img, generator=gen); y = torch.randint(0, classes, (B,), generator=gen)
    s = {k: v.cuda() for k, v in sd.items()}
    f32 = ref_vit.loss_and_grads(s, x.cuda(), y.cuda(), cfg)
    leaf = {k: v.detach().clone().requires_grad_(True) for k, v in s.items()}
    with torch.autocast("cuda", dtype=torch.bfloat16):
        logits = ref_vit.vit_forward(leaf, x.cuda(), cfg)
        loss = torch.nn.functional.cross_entropy(logits.float(), y.cuda())
    loss.backward()
    auto = (logits.detach().float(), None, {k: v.grad for k, v in leaf.items()})
    d = trbf._dims(False); trbf.forward(x.cuda(), d); trbf.backward(y.cuda(), d)
    ours = (trbf.logits.clone(), None, trbf.grads_by_ref_key())
    print(f"== E={E} H={H} L={L} B={B} soften={soften}: logits rel ours {rel(ours[0], f32[0]):.3e} autocast {rel(auto[0], f32[0]):.3e}")
    groups = {}
    for k in f32[2]:
        short = k.split(".")[-2] + "." + k.split(".")[-1] if "heads" in k else ".".join(k.split(".")[-3:])
        if "blocks" in k:
            short = "b" + k.split(".")[2] + "." + short
        groups.setdefault(short, []).append(k)
    for gname, keys in groups.items():
        co = min(cos(ours[2][k], f32[2][k]) for k in keys)
        ca = min(cos(auto[2][k], f32[2][k]) for k in keys)
        flag = " <<<" if co < 0.999 and co < ca - 2e-4 else ""
        print(f"  {gname:<34} ours {co:.5f}  autocast {ca:.5f}{flag}")


if __name__ == "__main__":
    run(768, 12, 1, 4, 0.25)
    run(768, 12, 2, 4, 0.25)
    run(768, 12, 1, 4, 0.02)


def flips(E=768, H=12, B=32, soften=0.02, img=224, P=16):
    """Block-0 post-ReLU activation `h` of our bf16 path vs the fp32 oracle: sign mismatches on token-0 rows and overall."""
    classes, L = 2, 1
    cfg = ref_vit.Cfg(3, classes, (img // P) ** 2, E, P, H, L, B)
    sd = ref_vit.random_state_dict(cfg, seed=3)
    sd = {k: (v * soften if k.endswith(("key.weight", "query.weight")) else v) for k, v in sd.items()}
    tr = train_dp.DPTrainer(embedding_size=E, num_heads=H, num_blocks=L, batch_size=B, dtype=torch.bfloat16, dropout_p=0.0)
    tr.load_state_dict(sd)
    gen = torch.Generator().manual_seed(4)
    x = torch.rand(B, 3, img, img, generator=gen)
    d = tr._dims(False); tr.forward(x.cuda(), d)
    M, T = tr.M, tr.T
    al = lambda n: (n + 255) // 256 * 256
    off = al(M * E * 2) + 2 * al(M * 4) + al(M * 3 * E * 2) + al(M * E * 2) + al(B * H * T * 4)
    x_mid = tr.blk_acts[0][off: off + M * E * 4].view(torch.float32).view(M, E)
    off += al(M * E * 4) + al(M * E * 2) + 2 * al(M * 4)
    h = tr.blk_acts[0][off: off + M * 4 * E * 2].view(torch.bfloat16).view(M, 4 * E)
    s = {k: v.cuda() for k, v in sd.items()}
    t = ref_vit.patch_embed(s, x.cuda(), P)
    pfx = "transformer_encoder.blocks.0."
    a, _ = ref_vit.multi_head(s, pfx + "multi_head.", ref_vit.layer_norm(s, pfx + "ln1.", t), H)
    xm = t + a
    pre = ref_vit.layer_norm(s, pfx + "ln2.", xm) @ s[pfx + "ffwd.mlp.0.weight"].t() + s[pfx + "ffwd.mlp.0.bias"]
    pre = pre.reshape(M, 4 * E)
    mism = (h.float() > 0) != (pre > 0)
    rows0 = torch.arange(B, device="cuda") * T
    print(f"flips B={B}: x_mid rel L2 {((x_mid - xm.reshape(M, E)).norm() / xm.norm()).item():.3e}  "
          f"flip frac all {mism.float().mean().item():.5f}  token0 rows {mism[rows0].float().mean().item():.5f}  "
          f"x_mid token0 rel {((x_mid[rows0] - xm.reshape(M, E)[rows0]).norm() / xm.reshape(M, E)[rows0].norm()).item():.3e}")


if __name__ == "__main__":
    flips()
    run(768, 12, 1, 32, 0.02)
