"""ORACLE / CPU BASELINE — test infrastructure only (bench.py's cpu_baseline leg and `--impl reference` arm, tests).

One training step of the reference's path on the host CPU, through the oracle restatement (oracle/ref_vit.py):
forward (train mode, dropout p=0.2 with torch Bernoulli masks) + mean cross-entropy + backward + AdamW(lr=1e-4,
weight_decay=1e-4) — the step lines of reference src/train.py:89-96 — with as many host threads as torch uses.
/root/reference is not on the GPU box, so this port (pinned to the reference by tests/test_oracle.py) is what runs.
"""
import os
import time

import torch

from . import ref_vit


class CpuTrainer:
    def __init__(self, *, input_channels, num_classes, image_size, patch_size, embedding_size, num_heads, num_blocks,
                 batch_size, lr=1e-4, weight_decay=1e-4, seed=0):
        n = (image_size // patch_size) ** 2
        self.cfg = ref_vit.Cfg(input_channels, num_classes, n, embedding_size, patch_size, num_heads, num_blocks,
                               batch_size)
        self.img = image_size
        sd = ref_vit.random_state_dict(self.cfg, seed=seed)
        self.params = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        self.opt = torch.optim.AdamW(list(self.params.values()), lr=lr, weight_decay=weight_decay)
        self.gen = torch.Generator().manual_seed(seed + 1)

    def batch(self):
        c = self.cfg
        x = torch.rand(c.batch_size, c.input_channels, self.img, self.img, generator=self.gen)
        y = torch.randint(0, c.num_classes, (c.batch_size,), generator=self.gen)
        return x, y

    def step(self, x, y, dropout_p=0.2):
        c = self.cfg
        T, E = c.num_patches + 1, c.embedding_size
        masks = None
        if dropout_p > 0:
            masks = [tuple(torch.rand(c.batch_size, T, E, generator=self.gen) >= dropout_p for _ in range(2))
                     for _ in range(c.num_blocks)]
        logits = ref_vit.vit_forward(self.params, x, c, masks)
        loss = torch.nn.functional.cross_entropy(logits, y)
        self.opt.zero_grad(set_to_none=True)
        loss.backward()
        self.opt.step()
        return float(loss.detach())


def time_steps(trainer, steps, warmup):
    x, y = trainer.batch()
    for _ in range(warmup):
        trainer.step(x, y)
    t0 = time.perf_counter()
    for _ in range(steps):
        trainer.step(x, y)
    return (time.perf_counter() - t0) / max(steps, 1)


def host_description():
    model = "?"
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    model = line.split(":", 1)[1].strip()
                    break
    except OSError:
        pass
    return {"cpu_model": model, "os_cpu_count": os.cpu_count(), "torch_threads": torch.get_num_threads()}
