"""Generate tests/golden/*.pt by running the REAL reference modules (import from /root/reference, CPU, fp32).

    python tests/golden/make_golden.py          # only works in the build container; the GPU box has no /root/reference

Each fixture holds: the config, the reference state_dict (reference keys/shapes), a uint8 image batch (the model input
is image/255, what torchvision's ToTensor yields, reference train.py:151-155), labels, and the reference's logits,
mean cross-entropy loss, every parameter gradient, and block-0 attention probabilities (`attention_probs`,
reference transformer.py:48).  Both model sides run in eval() mode (dropout off) with grad enabled, SURVEY §8c.
"""
import importlib.util
import os
import sys

import torch

REF = "/root/reference/src/VisionTransformer"
HERE = os.path.dirname(os.path.abspath(__file__))

CASES = {
    # name: (C, classes, img, E, P, H, L, B, seed)
    "tiny": (3, 3, 32, 16, 8, 2, 2, 2, 11),          # T=17, hs=8
    "odd": (1, 5, 40, 24, 10, 3, 1, 3, 12),          # C=1, P=10 (K=100), hs=8, T=17
    "c1": (3, 2, 224, 4, 16, 4, 4, 5, 13),           # BASELINE configs[0]: repo default ViT, hs=1, T=197
}


def load_reference():
    spec = importlib.util.spec_from_file_location("ref_VisionTransformer", os.path.join(REF, "__init__.py"),
                                                  submodule_search_locations=[REF])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_VisionTransformer"] = mod
    spec.loader.exec_module(mod)
    from ref_VisionTransformer import config, vit
    return vit, config


def main():
    vit, config = load_reference()
    for name, (C, classes, img, E, P, H, L, B, seed) in CASES.items():
        torch.manual_seed(seed)
        N = (img // P) ** 2
        cfg = config.ViTConfig(input_channels=C, num_classes=classes, num_patches=N, embedding_size=E, patch_size=P,
                               num_heads=H, num_blocks=L, device="cpu", batch_size=B)
        model = vit.VisionTransformer(cfg).eval()
        # move LayerNorm/bias values off their trivial init so that every gradient path is exercised
        with torch.no_grad():
            for k, p in model.named_parameters():
                if k.endswith("bias") or "ln" in k or k.startswith("mlp.2"):
                    p.add_(0.1 * torch.randn_like(p))
        img_u8 = torch.randint(0, 256, (B, C, img, img), dtype=torch.uint8)
        labels = torch.randint(0, classes, (B,))
        x = img_u8.float() / 255.0
        logits = model(x)
        loss = torch.nn.CrossEntropyLoss()(logits, labels)
        loss.backward()
        probs0 = model.transformer_encoder.blocks[0].multi_head.attention_probs.detach()
        out = {
            "config": dict(input_channels=C, num_classes=classes, num_patches=N, embedding_size=E, patch_size=P,
                           num_heads=H, num_blocks=L, batch_size=B, image_size=img),
            "state_dict": {k: v.detach().clone() for k, v in model.state_dict().items()},
            "images_u8": img_u8, "labels": labels,
            "logits": logits.detach(), "loss": loss.detach(),
            "grads": {k: p.grad.detach().clone() for k, p in model.named_parameters()},
            "probs0": probs0 if probs0.numel() <= 200_000 else probs0[:1, :1].clone(),
        }
        path = os.path.join(HERE, f"{name}.pt")
        torch.save(out, path)
        print(name, "logits", logits.detach().flatten()[:4].tolist(), "loss", float(loss), os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
