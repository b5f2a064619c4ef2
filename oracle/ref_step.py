"""ORACLE / BASELINE — test and measurement infrastructure only (bench.py's `--impl reference` arm, its
`cpu_baseline` leg and its `reference_gpu_eager` side number).  Never imported by the product path.

The reference's OWN training step: the unmodified modules staged under oracle/_ref/py/src (oracle/ref_loader.py)
driven by the step lines of its src/train.py:89-98 — `model(tensors)`, `CrossEntropyLoss`, `zero_grad(set_to_none)`,
`backward`, `torch.optim.AdamW(lr=1e-4, weight_decay=1e-4).step()` — in train() mode (dropout live), on the host CPU
with all its threads or, as the PyTorch-eager competitor SURVEY §0 names, on the same B200.
"""
import contextlib
import os
import time

import torch

from . import ref_loader


def use_all_host_threads():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to its children; the CPU arm must use every host core."""
    n = os.cpu_count() or 1
    try:
        n = min(n, len(os.sched_getaffinity(0)))
    except (AttributeError, OSError):
        pass
    torch.set_num_threads(max(n, 1))
    return torch.get_num_threads()


class RefTrainer:
    def __init__(self, *, input_channels, num_classes, image_size, patch_size, embedding_size, num_heads, num_blocks,
                 batch_size, device="cpu", autocast=False, lr=1e-4, weight_decay=1e-4, seed=0):
        vit, _, config = ref_loader.modules()
        self.device, self.autocast, self.B = device, autocast, batch_size
        self.img, self.C, self.classes = image_size, input_channels, num_classes
        torch.manual_seed(seed)
        cfg = config.ViTConfig(input_channels=input_channels, num_classes=num_classes,
                               num_patches=(image_size // patch_size) ** 2, embedding_size=embedding_size,
                               patch_size=patch_size, num_heads=num_heads, num_blocks=num_blocks, device=device,
                               batch_size=batch_size)
        self.model = vit.VisionTransformer(cfg)                                               # train.py:65
        self.opt = torch.optim.AdamW(self.model.parameters(), lr=lr, weight_decay=weight_decay)  # train.py:66
        self.model = self.model.to(device)                                                    # train.py:80
        self.crit = torch.nn.CrossEntropyLoss()                                               # train.py:81
        self.gen = torch.Generator().manual_seed(seed + 1)

    def batch(self):
        x = torch.rand(self.B, self.C, self.img, self.img, generator=self.gen)
        y = torch.randint(0, self.classes, (self.B,), generator=self.gen)
        return x.to(self.device), y.to(self.device)

    def step(self, x, y):
        ctx = torch.autocast(self.device.split(":")[0], dtype=torch.bfloat16) if self.autocast else \
            contextlib.nullcontext()
        with ctx:
            logits = self.model(x)                          # train.py:92
            loss = self.crit(logits.float(), y)             # train.py:93
        self.opt.zero_grad(set_to_none=True)                # train.py:94
        loss.backward()                                     # train.py:95
        self.opt.step()                                     # train.py:96
        return loss


def time_steps(trainer, steps, warmup):
    """Seconds per step.  CUDA: device-timed with events; CPU: wall clock."""
    x, y = trainer.batch()
    cuda = str(trainer.device).startswith("cuda")
    for _ in range(warmup):
        trainer.step(x, y)
    if cuda:
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            loss = trainer.step(x, y)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / max(steps, 1), float(loss.detach())
    t0 = time.perf_counter()
    for _ in range(steps):
        loss = trainer.step(x, y)
    return (time.perf_counter() - t0) / max(steps, 1), float(loss.detach())
