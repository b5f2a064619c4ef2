"""Data-parallel training step over the stage-level C ABI — the B200 replacement for the step lines of the
reference's src/train.py (:66 AdamW(lr, weight_decay=1e-4), :81/:93 CrossEntropyLoss, :94-96 zero_grad/backward/step).

One process per GPU.  All parameters live in ONE flat fp32 arena (forward order: embedding, block 0..L-1, head;
every tensor 256-byte aligned), gradients / Adam moments / bf16 shadow weights in arenas of the same layout, and
every activation buffer is allocated once: a steady-state step performs no allocation and no host sync.

Multi-GPU: the path shards by images (pure data parallel).  Gradients are exchanged on contiguous arena slices
("buckets", reverse layer order) issued on a side stream as soon as the backward stage that completes a bucket has
been enqueued, so the exchange overlaps the rest of backward; the fused AdamW kernel folds the 1/world mean into its
gradient read.  Two exchange schemes:
  * all-reduce (`zero1=False`): every rank sums every bucket and runs AdamW over the whole arena;
  * ZeRO-1 (`zero1=True`, default for world > 1): each bucket is REDUCE-SCATTERED (bf16 or fp32), every rank runs
    AdamW on its 1/world shard of every bucket only (Adam moments exist only for that shard), and the updated fp32
    parameters are ALL-GATHERED bucket by bucket in forward order on the side stream while the next forward already
    runs (each stage waits for the event of its own bucket).
`cls_tkn_embd` is a per-batch-slot parameter (reference vit.py:31-33) and is averaged across ranks like any other
parameter, exactly as DistributedDataParallel would treat it.
"""
import ctypes
import os
import sys

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
if _HERE not in sys.path:
    sys.path.insert(0, _HERE)

from VisionTransformer import _C  # noqa: E402
from VisionTransformer import _engine as eng  # noqa: E402

ALIGN = 64  # elements: 256 B in fp32, 128 B in bf16 — keeps every tensor TMA-legal (16 B) in both arenas
SHARD_ALIGN = ALIGN * 8  # every STAGE starts at a multiple of 512 elements, so any bucket (a union of whole stages)
#                          splits evenly over 1 / 2 / 4 / 8 ranks into ALIGN-aligned shards (ZeRO-1)
_SECTIONS = ("query", "key", "value")


# ---------------------------------------------------------------------------------------------------- layout
def param_layout(*, C, P, E, H, L, T, B, classes):
    """[(stage, field, reference-key or None, shape)] in arena order.  stage: 'embed' | int layer | 'head'."""
    F = 4 * E
    out = [("embed", "w", "emdeddings.sequence.0.weight", (E, C, P, P)),
           ("embed", "bias", "emdeddings.sequence.0.bias", (E,)),
           ("embed", "cls", "emdeddings.cls_tkn_embd", (B, 1, E)),
           ("embed", "pos", "emdeddings.pos_embd", (1, T, E))]
    for i in range(L):
        b = f"transformer_encoder.blocks.{i}."
        out += [(i, "ln1_w", b + "ln1.weight", (E,)), (i, "ln1_b", b + "ln1.bias", (E,)),
                (i, "ln2_w", b + "ln2.weight", (E,)), (i, "ln2_b", b + "ln2.bias", (E,)),
                (i, "w_qkv", None, (3 * E, E)),  # packed from heads.h.{query,key,value}.weight
                (i, "w_proj", b + "multi_head.proj.weight", (E, E)), (i, "b_proj", b + "multi_head.proj.bias", (E,)),
                (i, "w_fc1", b + "ffwd.mlp.0.weight", (F, E)), (i, "b_fc1", b + "ffwd.mlp.0.bias", (F,)),
                (i, "w_fc2", b + "ffwd.mlp.2.weight", (E, F)), (i, "b_fc2", b + "ffwd.mlp.2.bias", (E,))]
    out += [("head", "w1", "mlp.0.weight", (F, E)), ("head", "b1", "mlp.0.bias", (F,)),
            ("head", "ln_w", "mlp.2.weight", (F,)), ("head", "ln_b", "mlp.2.bias", (F,)),
            ("head", "w3", "mlp.3.weight", (classes, F)), ("head", "b3", "mlp.3.bias", (classes,))]
    return out


def arena_offsets(layout):
    offs, tot, prev = [], 0, None
    for stage, _, _, shape in layout:
        if stage != prev:
            tot = (tot + SHARD_ALIGN - 1) // SHARD_ALIGN * SHARD_ALIGN
            prev = stage
        n = 1
        for s in shape:
            n *= s
        offs.append((tot, n))
        tot += (n + ALIGN - 1) // ALIGN * ALIGN
    return offs, (tot + SHARD_ALIGN - 1) // SHARD_ALIGN * SHARD_ALIGN


def auto_bucket_mb(total_elems):
    """Bucket size (MB of fp32 gradients) of the exchange when the caller does not choose one: TWO buckets — 80 % of the
    arena (the head and all but the first few blocks), whose exchange overlaps the rest of the backward pass, and the
    remainder, the only part left exposed.  A collective that co-runs with the persistent GEMM kernels takes SMs from
    their static tile walk, so FEW launches beat many small ones.  Measured on 4 x B200 (profiles/r02_bucket_policy_n4.txt):
    ViT-B/16 (337 MB) one bucket 33.53-33.64 ms, [279 | 58] MB 33.27, seven 48 MB buckets 33.46, one GPU 32.55;
    ViT-L/16 (1.17 GB) [1025 | 149] MB 105.7 ms, three buckets 106.0, twenty-five 48 MB buckets 107.6-107.9, one GPU 104.6."""
    return 0.8 * total_elems * 4 / 2 ** 20


def make_buckets(layout, offs, total, bucket_elems):
    """Contiguous arena slices in the order their gradients complete (head first, embedding last).
    -> [(start, end, stage_that_completes_it)]"""
    stages = []  # (stage, start, end) in forward order
    for (stage, _, _, _), (o, n) in zip(layout, offs):
        end = (o + n + ALIGN - 1) // ALIGN * ALIGN
        if stages and stages[-1][0] == stage:
            stages[-1] = (stage, stages[-1][1], end)
        else:
            stages.append((stage, o, end))
    buckets, cur_end = [], total
    for stage, start, _ in reversed(stages):
        if cur_end - start >= bucket_elems or start == 0:   # the embedding stage (offset 0) always closes the last one
            buckets.append((start, cur_end, stage))
            cur_end = start
    return buckets


class BucketReducer:
    """The one exchange step of the data-parallel path: sum-all-reduce of contiguous gradient-arena buckets, each
    launched as soon as the backward stage that completes it has been enqueued.  On CUDA the collective runs on a
    side stream behind an event recorded on the compute stream (so it overlaps the remaining backward); on CPU
    tensors (gloo, used by the world_size-2 tests) the same calls run without streams."""

    def __init__(self, flat_grad, buckets, process_group=None, world_size=1):
        self.g, self.buckets, self.pg, self.world = flat_grad, buckets, process_group, int(world_size)
        self.cuda = flat_grad.is_cuda
        if self.cuda and self.world > 1:
            self.stream = torch.cuda.Stream(device=flat_grad.device)
            self.events = [torch.cuda.Event() for _ in buckets]

    def ready(self, stage, handles):
        """Launch the all-reduce of every bucket completed by `stage`'s backward (already enqueued)."""
        if self.world <= 1:
            return
        for k, (a, b, st) in enumerate(self.buckets):
            if st != stage:
                continue
            if self.cuda:
                self.events[k].record(torch.cuda.current_stream(self.g.device))
                with torch.cuda.stream(self.stream):
                    self.stream.wait_event(self.events[k])
                    handles.append(torch.distributed.all_reduce(self.g[a:b], group=self.pg, async_op=True))
            else:
                handles.append(torch.distributed.all_reduce(self.g[a:b], group=self.pg, async_op=True))

    @staticmethod
    def finish(handles):
        for h in handles:
            h.wait()  # CUDA: the compute stream waits for the reduced buckets; the host does not block


class ShardedReducer:
    """ZeRO-1 exchange of the data-parallel path.  Backward order: each completed bucket [a, b) is staged
    (`pack(g[a:b], gb[a:b])`: bf16 copy + the fp32 arena cleared, one pass) and reduce-scattered, so rank r receives
    the SUM over ranks of its shard [a + r*(b-a)/world, a + (r+1)*(b-a)/world) into `gs` (all shards back to back,
    bucket order).  After the sharded AdamW, `gather(p)` all-gathers every bucket of the parameter arena in place, in
    FORWARD order, and records one event per bucket for the next forward to wait on.  CUDA: everything runs on a side
    stream; CPU tensors (gloo, the world_size-2 tests): same calls, no streams.

    pack(g_slice, out_slice_or_None): stage one bucket — out = cast(g) when the exchange dtype is bf16, g = 0."""

    def __init__(self, flat_grad, buckets, process_group, world_size, rank, comm_dtype, pack):
        self.g, self.buckets, self.pg, self.world, self.rank = flat_grad, buckets, process_group, int(world_size), int(rank)
        self.comm_dtype, self.pack = comm_dtype, pack
        self.cuda = flat_grad.is_cuda
        dev = flat_grad.device
        for a, b, _ in buckets:
            if (b - a) % (self.world * ALIGN) != 0:
                raise ValueError(f"bucket [{a},{b}) does not split into {self.world} {ALIGN}-element-aligned shards")
        self.shard_len = [(b - a) // self.world for a, b, _ in buckets]
        self.shard_off = [sum(self.shard_len[:k]) for k in range(len(buckets))]
        self.shard_total = sum(self.shard_len)
        self.gs = torch.zeros(self.shard_total, dtype=comm_dtype, device=dev)         # my reduced gradient shards
        self.gb = torch.zeros(flat_grad.numel(), dtype=comm_dtype, device=dev) if comm_dtype != flat_grad.dtype else None
        if self.cuda:
            self.stream = torch.cuda.Stream(device=dev)
            self.events = [torch.cuda.Event() for _ in buckets]
            self.gather_events = [torch.cuda.Event() for _ in buckets]
        self.gather_pending = False

    def shard_range(self, k):
        """(arena start, arena end) of THIS rank's shard of bucket k."""
        a, _, _ = self.buckets[k]
        return a + self.rank * self.shard_len[k], a + (self.rank + 1) * self.shard_len[k]

    def _scatter(self, k):
        a, b, _ = self.buckets[k]
        out = self.gs[self.shard_off[k]: self.shard_off[k] + self.shard_len[k]]
        if self.gb is not None:
            self.pack(self.g[a:b], self.gb[a:b])
            torch.distributed.reduce_scatter_tensor(out, self.gb[a:b], group=self.pg)
        else:
            torch.distributed.reduce_scatter_tensor(out, self.g[a:b], group=self.pg)
            self.pack(self.g[a:b], None)

    def ready(self, stage, handles):
        for k, (a, b, st) in enumerate(self.buckets):
            if st != stage:
                continue
            if self.cuda:
                self.events[k].record(torch.cuda.current_stream(self.g.device))
                with torch.cuda.stream(self.stream):
                    self.stream.wait_event(self.events[k])
                    self._scatter(k)   # synchronous collective w.r.t. the SIDE stream; the compute stream runs on
            else:
                self._scatter(k)
            handles.append(k)

    def finish(self, handles):
        if self.cuda and handles:
            torch.cuda.current_stream(self.g.device).wait_stream(self.stream)

    def gather(self, arena, after=None):
        """In-place all-gather of every bucket of `arena` from the ranks' shards, forward order (embedding first).
        after(k, a, b): optional per-bucket follow-up on the same stream (the bf16 shadow cast)."""
        order = range(len(self.buckets) - 1, -1, -1)
        if self.cuda:
            self.stream.wait_stream(torch.cuda.current_stream(arena.device))
        for k in order:
            a, b, _ = self.buckets[k]
            s0, s1 = self.shard_range(k)
            if self.cuda:
                with torch.cuda.stream(self.stream):
                    torch.distributed.all_gather_into_tensor(arena[a:b], arena[s0:s1], group=self.pg)
                    if after is not None:
                        after(k, a, b)
                    self.gather_events[k].record(self.stream)
            else:
                torch.distributed.all_gather_into_tensor(arena[a:b], arena[s0:s1].clone(), group=self.pg)
                if after is not None:
                    after(k, a, b)
        self.gather_pending = self.cuda

    def wait_bucket(self, k):
        if self.cuda and self.gather_pending:
            torch.cuda.current_stream(self.g.device).wait_event(self.gather_events[k])

    def wait_all(self):
        if self.cuda and self.gather_pending:
            torch.cuda.current_stream(self.g.device).wait_stream(self.stream)
            self.gather_pending = False


# ---------------------------------------------------------------------------------------------------- trainer
class DPTrainer:
    def __init__(self, *, input_channels=3, num_classes=2, image_size=224, patch_size=16, embedding_size=768,
                 num_heads=12, num_blocks=12, batch_size=256, dtype=torch.bfloat16, device=None, lr=1e-4,
                 weight_decay=1e-4, betas=(0.9, 0.999), eps=1e-8, dropout_p=eng.DROPOUT_P, process_group=None,
                 world_size=1, bucket_mb=None, seed=0, last_block_row0=False, zero1=None, grad_comm=None):
        dev = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        eng.require_cuda(torch.empty(0, device=dev), "trainer device")
        if _C.lib().vitb200_device_ok() != 1:
            raise RuntimeError("libvitb200 runs on sm_100 (B200) only")
        self.dev, self.dtype, self.code = dev, dtype, _C.dtype_code(dtype)
        gh = image_size // patch_size
        self.B, self.C, self.P, self.E, self.H, self.L = batch_size, input_channels, patch_size, embedding_size, \
            num_heads, num_blocks
        self.T, self.classes, self.img = gh * gh + 1, num_classes, image_size
        self.M = self.B * self.T
        self.lr, self.wd, self.betas, self.eps = lr, weight_decay, betas, eps
        self.dropout_p, self.seed, self.step_no = float(dropout_p), int(seed), 0
        self.pg, self.world = process_group, int(world_size)
        # every data-parallel rank draws its OWN dropout masks (independent regularisation over the global batch)
        self.rank = torch.distributed.get_rank(process_group) if (self.world > 1 and
                                                                  torch.distributed.is_initialized()) else 0
        # opt-in dead-row elimination: only token 0 of the LAST block feeds the classifier (reference vit.py:80), so
        # that block's proj / LayerNorm-2 / MLP and their backward run on B rows instead of B*T.  Same logits, same
        # gradients; fewer executed FLOPs than the algorithmic count, hence off unless asked for.
        self.row0 = bool(last_block_row0) and num_blocks >= 1

        self.layout = param_layout(C=self.C, P=self.P, E=self.E, H=self.H, L=self.L, T=self.T, B=self.B,
                                   classes=self.classes)
        self.offs, self.total = arena_offsets(self.layout)
        z = lambda dt, n=None: torch.zeros(self.total if n is None else n, dtype=dt, device=dev)  # noqa: E731
        self.p, self.g = z(torch.float32), z(torch.float32)
        self.shadow = z(torch.bfloat16) if dtype == torch.bfloat16 else None
        if bucket_mb is None or bucket_mb <= 0:
            bucket_mb = auto_bucket_mb(self.total)
        self.buckets = make_buckets(self.layout, self.offs, self.total, int(bucket_mb * (1 << 20) / 4))
        # exchange scheme (module docstring): ZeRO-1 by default when there is more than one rank; gradients travel
        # in bf16 in bf16 mode and in fp32 in fp32 mode (the DP == single-GPU 1e-5 gate)
        self.zero1 = (self.world > 1) if zero1 is None else (bool(zero1) and self.world > 1)
        gc = grad_comm if grad_comm is not None else ("bf16" if dtype == torch.bfloat16 else "fp32")
        if gc not in ("bf16", "fp32"):
            raise ValueError(f"grad_comm must be 'bf16' or 'fp32', got {grad_comm!r}")
        self.comm_dtype = torch.bfloat16 if gc == "bf16" else torch.float32
        if self.zero1:
            self.reducer = ShardedReducer(self.g, self.buckets, self.pg, self.world, self.rank, self.comm_dtype,
                                          self._pack_grads)
            # Adam moments exist only for this rank's shards (bucket order, back to back)
            self.m, self.v = z(torch.float32, self.reducer.shard_total), z(torch.float32, self.reducer.shard_total)
            stage_bucket = {}
            for k, (a, b, _) in enumerate(self.buckets):
                for (stage, _, _, _), (o, _) in zip(self.layout, self.offs):
                    if a <= o < b:
                        stage_bucket[stage] = k
            self._stage_bucket = stage_bucket
        else:
            self.m, self.v = z(torch.float32), z(torch.float32)
            self.reducer = None   # BucketReducer, built below once the side streams exist

        # pointer tables per stage
        self._views = {}
        for (stage, field, _, shape), (o, n) in zip(self.layout, self.offs):
            self._views[(stage, field)] = (o, n, shape)
        lp = {"w_qkv", "w_proj", "w_fc1", "w_fc2"}

        def ptrs(stage, fields, grads=False, lowp=()):
            d = {}
            for f in fields:
                o, n, _ = self._views[(stage, f)]
                src = self.g if grads else (self.shadow if (f in lowp and self.shadow is not None) else self.p)
                d[f] = src[o:o + n]
            return d

        self.embed_p = ptrs("embed", ("bias", "cls", "pos"))
        o, n, _ = self._views[("embed", "w")]
        self.embed_p["w_f32"] = self.p[o:o + n]
        self.embed_p["w_lp"] = self.shadow[o:o + n] if self.shadow is not None else None
        self.embed_g = ptrs("embed", _C.EMBED_FIELDS_G, grads=True)
        self.block_p = [ptrs(i, _C.BLOCK_FIELDS_P, lowp=lp) for i in range(self.L)]
        self.block_g = eng.chain_block_grads([ptrs(i, _C.BLOCK_FIELDS_G, grads=True) for i in range(self.L)])
        self.head_p = ptrs("head", _C.HEAD_FIELDS)
        o, n, _ = self._views[("head", "w1")]
        self.head_p["w1_lp"] = self.shadow[o:o + n] if self.shadow is not None else None
        self.head_g = ptrs("head", _C.HEAD_FIELDS, grads=True)

        # activations / scratch, allocated once
        self.plan = eng.Plan(B=self.B, T=self.T, E=self.E, H=self.H, C=self.C, P=self.P, Hi=self.img, Wi=self.img,
                             classes=self.classes, L=self.L, dtype=dtype)
        pl = self.plan
        self.emb_acts = pl.new_bytes(pl.embed_act_bytes, dev)
        self.xs = [pl.new_stream_tensor(dev) for _ in range(self.L + 1)]
        self.blk_acts = [pl.new_bytes(pl.block_act_bytes, dev) for _ in range(self.L)]
        self.head_acts = pl.new_bytes(pl.head_act_bytes, dev)
        self.scratch = pl.new_bytes(max(pl.block_scratch_bytes, pl.head_act_bytes), dev)
        self.gstream = pl.new_stream_tensor(dev)
        self.dy = [pl.new_lp(dev), pl.new_lp(dev)]
        self.logits = torch.empty(self.B, self.classes, dtype=torch.float32, device=dev)
        self.dlogits = torch.empty_like(self.logits)
        self.loss = torch.zeros(1, dtype=torch.float32, device=dev)

        # host <-> device staging for step_host()
        self.copy_stream = torch.cuda.Stream(device=dev)
        if self.reducer is None:
            self.reducer = BucketReducer(self.g, self.buckets, self.pg, self.world)
        self._in = [(torch.empty(self.B, self.C, self.img, self.img, dtype=torch.float32, device=dev),
                     torch.empty(self.B, dtype=torch.int64, device=dev)) for _ in range(2)]
        self._in_u8 = None  # uint8 staging buffers, allocated on the first step_host_u8
        self._in_ready = [torch.cuda.Event() for _ in range(2)]
        self._in_free = [torch.cuda.Event() for _ in range(2)]
        self._slot = 0
        self.loss_host = torch.zeros(64, dtype=torch.float32).pin_memory()
        self._loss_ev = [torch.cuda.Event() for _ in range(64)]
        self.h2d_bytes = self._in[0][0].numel() * 4 + self.B * 8
        self.d2h_bytes = 4

    # ------------------------------------------------------------------------------------------ parameters
    def view(self, stage, field, arena=None):
        if self.zero1:
            self.reducer.wait_all()
        o, n, shape = self._views[(stage, field)]
        return (self.p if arena is None else arena)[o:o + n].view(shape)

    def load_state_dict(self, sd):
        """Fill the arena from a state_dict with the REFERENCE's keys (per-head q/k/v packed on the fly)."""
        E, H = self.E, self.H
        hs = E // H
        with torch.no_grad():
            for stage, field, key, shape in self.layout:
                dst = self.view(stage, field)
                if key is not None:
                    src = sd[key]
                    if tuple(src.shape) != tuple(shape):
                        raise RuntimeError(f"size mismatch for {key}: {tuple(src.shape)} vs {tuple(shape)}")
                    dst.copy_(src.to(self.dev, torch.float32))
                else:
                    for s, sec in enumerate(_SECTIONS):
                        for h in range(H):
                            k = f"transformer_encoder.blocks.{stage}.multi_head.heads.{h}.{sec}.weight"
                            dst[s * E + h * hs: s * E + (h + 1) * hs].copy_(sd[k].to(self.dev, torch.float32))
        self.sync_shadow()

    def init_parameters(self, seed=0):
        """Random init with the reference's distributions (vit.py:31-36 randn for cls/pos; nn.Linear / nn.Conv2d
        default Kaiming-uniform(a=sqrt(5)) = U(+-1/sqrt(fan_in)) for weights and biases; LayerNorm 1 / 0)."""
        gen = torch.Generator(device=self.dev).manual_seed(int(seed))
        fan_in = {"w": self.C * self.P * self.P, "bias": self.C * self.P * self.P, "w_qkv": self.E, "w_proj": self.E,
                  "b_proj": self.E, "w_fc1": self.E, "b_fc1": self.E, "w_fc2": 4 * self.E, "b_fc2": 4 * self.E,
                  "w1": self.E, "b1": self.E, "w3": 4 * self.E, "b3": 4 * self.E}
        with torch.no_grad():
            for stage, field, _, shape in self.layout:
                dst = self.view(stage, field)
                if field in ("cls", "pos"):
                    dst.copy_(torch.randn(shape, generator=gen, device=self.dev))
                elif field in ("ln1_w", "ln2_w", "ln_w"):
                    dst.fill_(1.0)
                elif field in ("ln1_b", "ln2_b", "ln_b"):
                    dst.zero_()
                else:
                    bound = fan_in[field] ** -0.5
                    dst.copy_((torch.rand(shape, generator=gen, device=self.dev) * 2 - 1) * bound)
        self.sync_shadow()

    def state_dict(self):
        E, H = self.E, self.H
        hs = E // H
        out = {}
        for stage, field, key, _ in self.layout:
            t = self.view(stage, field).detach().clone()
            if key is not None:
                out[key] = t
            else:
                for h in range(H):
                    for sec in ("key", "query", "value"):
                        s = _SECTIONS.index(sec)
                        out[f"transformer_encoder.blocks.{stage}.multi_head.heads.{h}.{sec}.weight"] = \
                            t[s * E + h * hs: s * E + (h + 1) * hs].clone()
        return out

    def grads_by_ref_key(self):
        E, H = self.E, self.H
        hs = E // H
        out = {}
        for stage, field, key, _ in self.layout:
            t = self.view(stage, field, self.g)
            if key is not None:
                out[key] = t
            else:
                for h in range(H):
                    for s, sec in enumerate(_SECTIONS):
                        out[f"transformer_encoder.blocks.{stage}.multi_head.heads.{h}.{sec}.weight"] = \
                            t[s * E + h * hs: s * E + (h + 1) * hs]
        return out

    # ---- optimizer state in the reference's torch.optim.AdamW format (train.py:66,73,110) ----------
    def _moment_views(self, arena):
        """reference key -> view of `arena` (per-head slices of the packed q/k/v rows), same mapping as grads."""
        E, H = self.E, self.H
        hs = E // H
        out = {}
        for stage, field, key, _ in self.layout:
            t = self.view(stage, field, arena)
            if key is not None:
                out[key] = t
            else:
                for h in range(H):
                    for s_, sec in enumerate(_SECTIONS):
                        out[f"transformer_encoder.blocks.{stage}.multi_head.heads.{h}.{sec}.weight"] = \
                            t[s_ * E + h * hs: s_ * E + (h + 1) * hs]
        return out

    def _full_moments(self):
        """(m, v) as full arenas.  ZeRO-1 keeps only this rank's shards: gather them (collective: every rank calls)."""
        if not self.zero1:
            return self.m, self.v
        red, out = self.reducer, []
        for sh in (self.m, self.v):
            full = torch.zeros(self.total, dtype=torch.float32, device=self.dev)
            for k, (a, b, _) in enumerate(self.buckets):
                torch.distributed.all_gather_into_tensor(full[a:b], sh[red.shard_off[k]: red.shard_off[k] + red.shard_len[k]],
                                                         group=self.pg)
            out.append(full)
        return out

    def _store_moments(self, m_full, v_full):
        if not self.zero1:
            self.m.copy_(m_full)
            self.v.copy_(v_full)
            return
        red = self.reducer
        for k in range(len(self.buckets)):
            s0, s1 = red.shard_range(k)
            so, n = red.shard_off[k], red.shard_len[k]
            self.m[so:so + n].copy_(m_full[s0:s1])
            self.v[so:so + n].copy_(v_full[s0:s1])

    def optimizer_state_dict(self):
        """What `torch.optim.AdamW(reference_model.parameters()).state_dict()` would hold after the same steps:
        state indexed by position in the REFERENCE's parameter list, so a checkpoint written here resumes in the
        unmodified reference (src/train.py:73) and vice versa."""
        from VisionTransformer.checkpoint import reference_param_keys
        keys = reference_param_keys(self.L, self.H)
        m_full, v_full = self._full_moments()
        m, v = self._moment_views(m_full), self._moment_views(v_full)
        state = {}
        if self.step_no > 0:
            for i, k in enumerate(keys):
                state[i] = {"step": torch.tensor(float(self.step_no)), "exp_avg": m[k].detach().clone(),
                            "exp_avg_sq": v[k].detach().clone()}
        group = {"lr": self.lr, "betas": tuple(self.betas), "eps": self.eps, "weight_decay": self.wd, "amsgrad": False,
                 "foreach": None, "maximize": False, "capturable": False, "differentiable": False, "fused": None,
                 "decoupled_weight_decay": True, "params": list(range(len(keys)))}
        return {"state": state, "param_groups": [group]}

    def load_optimizer_state_dict(self, sd):
        """Inverse of optimizer_state_dict(); accepts the reference's own checkpoint['optimizer_state_dict']."""
        from VisionTransformer.checkpoint import reference_param_keys
        keys = reference_param_keys(self.L, self.H)
        if len(sd["param_groups"]) != 1 or len(sd["param_groups"][0]["params"]) != len(keys):
            raise ValueError(f"loaded state dict contains a parameter group that doesn't match the size of "
                             f"optimizer's group ({len(keys)} parameters in the reference list)")
        g = sd["param_groups"][0]
        self.lr, self.betas, self.eps, self.wd = g["lr"], tuple(g["betas"]), g["eps"], g["weight_decay"]
        m_full = torch.zeros(self.total, dtype=torch.float32, device=self.dev)
        v_full = torch.zeros_like(m_full)
        m, v = self._moment_views(m_full), self._moment_views(v_full)
        steps = set()
        with torch.no_grad():
            for i, k in enumerate(keys):
                st = sd["state"].get(i)
                if st is None:
                    continue
                if tuple(st["exp_avg"].shape) != tuple(m[k].shape):
                    raise ValueError(f"size mismatch for optimizer state of {k}: {tuple(st['exp_avg'].shape)} vs "
                                     f"{tuple(m[k].shape)}")
                m[k].copy_(st["exp_avg"].to(self.dev, torch.float32))
                v[k].copy_(st["exp_avg_sq"].to(self.dev, torch.float32))
                steps.add(int(float(st["step"])))
            self._store_moments(m_full, v_full)
        if len(steps) > 1:
            raise ValueError(f"per-parameter step counters differ ({sorted(steps)}); the flat AdamW kernel keeps one")
        self.step_no = steps.pop() if steps else 0

    def sync_shadow(self):
        if self.shadow is not None:
            eng.cast_bf16(self.p, self.shadow)

    def broadcast_parameters(self, src=0):
        if self.world > 1:
            torch.distributed.broadcast(self.p, src=src, group=self.pg)
            self.sync_shadow()

    # ------------------------------------------------------------------------------------------ one step
    def _dims(self, train):
        seed = (self.seed * 0x9E3779B97F4A7C15 + self.step_no * 0xD1B54A32D192ED03 +
                self.rank * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        return self.plan.dims(self.dropout_p if train else 0.0, seed)

    def _params_ready(self, stage):
        """ZeRO-1: the all-gather of the previous step's updated parameters runs on the side stream, bucket by bucket
        in forward order; a stage's kernels wait only for the bucket that holds its parameters."""
        if self.zero1:
            self.reducer.wait_bucket(self._stage_bucket[stage])

    def forward(self, images, d, u8_layout=0):
        pl = self.plan
        self._params_ready("embed")
        pl.embed_fwd(d, self.embed_p, images, self.xs[0], self.emb_acts, u8_layout)
        for i in range(self.L):
            self._params_ready(i)
            pl.block_fwd(d, i, self.block_p[i], self.xs[i], self.xs[i + 1], self.blk_acts[i],
                         row0=self.row0 and i == self.L - 1)
        self._params_ready("head")
        pl.head_fwd(d, self.head_p, self.xs[self.L], self.logits, self.head_acts, row0=self.row0)
        return self.logits

    def _pack_grads(self, g_slice, out_slice):
        """Exchange staging of one bucket (ShardedReducer): bf16 copy (if any) + fp32 gradient range cleared."""
        _C.check(_C.lib().vitb200_grad_pack(g_slice.data_ptr(), None if out_slice is None else out_slice.data_ptr(),
                                            g_slice.numel(), _C.stream_ptr(self.dev)))

    def _bucket_ready(self, stage, handles):
        self.reducer.ready(stage, handles)

    def _check_labels(self, labels):
        if labels.dtype != torch.int64 or labels.device != self.dev or not labels.is_contiguous() or \
                labels.numel() != self.B:
            raise RuntimeError(f"labels must be a contiguous int64 tensor of {self.B} class indices on {self.dev}, got "
                               f"{labels.dtype} {tuple(labels.shape)} on {labels.device}")

    def backward(self, labels, d):
        pl, L = self.plan, _C.lib()
        self._check_labels(labels)
        _C.check(L.vitb200_cross_entropy(self.logits.data_ptr(), labels.data_ptr(), self.loss.data_ptr(),
                                         self.dlogits.data_ptr(), self.B, self.classes, 1.0, _C.stream_ptr()))
        handles = []
        dy_a, dy_b = self.dy
        pl.head_bwd(d, self.head_p, self.head_g, self.head_acts, self.dlogits, self.gstream, dy_a, self.scratch,
                    row0=self.row0)
        self._bucket_ready("head", handles)
        for i in range(self.L - 1, -1, -1):
            pl.block_bwd(d, i, self.block_p[i], self.block_g[i], self.xs[i], self.blk_acts[i], self.gstream, dy_a,
                         dy_b, self.scratch, row0=self.row0 and i == self.L - 1)
            dy_a, dy_b = dy_b, dy_a
            self._bucket_ready(i, handles)
        pl.embed_bwd(d, self.embed_g, self.gstream, self.emb_acts)
        self._bucket_ready("embed", handles)
        self.reducer.finish(handles)

    def _optimizer_step_sharded(self, h):
        """ZeRO-1: AdamW on this rank's shard of every bucket (gradients = the reduce-scattered sums, moments exist for
        the shard only), then the all-gather of the updated parameters (+ bf16 shadow cast) on the side stream."""
        red, L = self.reducer, _C.lib()
        red.wait_all()   # a previous step's gathers write the non-owned ranges of p / shadow: finish them first
        bf = self.comm_dtype == torch.bfloat16
        for k in range(len(self.buckets)):
            s0, s1 = red.shard_range(k)
            so, n = red.shard_off[k], red.shard_len[k]
            gsk = red.gs[so:so + n]
            _C.check(L.vitb200_adamw_range(self.p[s0:s1].data_ptr(), None if bf else gsk.data_ptr(),
                                           gsk.data_ptr() if bf else None, self.m[so:so + n].data_ptr(),
                                           self.v[so:so + n].data_ptr(),
                                           None if self.shadow is None else self.shadow[s0:s1].data_ptr(), n, 0, h,
                                           _C.stream_ptr(self.dev)))

        def shadow_cast(k, a, b):
            if self.shadow is not None:
                _C.check(L.vitb200_cast_f32_bf16(self.p[a:b].data_ptr(), self.shadow[a:b].data_ptr(), b - a,
                                                 _C.stream_ptr(self.dev)))
        red.gather(self.p, after=shadow_cast)

    def optimizer_step(self):
        self.step_no += 1
        h = _C.AdamWHyper(self.lr, self.betas[0], self.betas[1], self.eps, self.wd, self.step_no, 1.0 / self.world)
        if self.zero1:
            return self._optimizer_step_sharded(h)
        # one launch: AdamW over the arena, bf16 shadow written, gradient arena cleared for the next backward
        _C.check(_C.lib().vitb200_adamw_range(self.p.data_ptr(), self.g.data_ptr(), None, self.m.data_ptr(),
                                              self.v.data_ptr(),
                                              None if self.shadow is None else self.shadow.data_ptr(), self.total, 1,
                                              h, _C.stream_ptr(self.dev)))

    def _on_device(fn):
        """Run a public entry point with the trainer's device current, so that torch's 'current stream' (where the
        C ABI enqueues) is the stream of THIS trainer's device even when the caller's current device differs."""
        import functools

        @functools.wraps(fn)
        def wrapped(self, *a, **k):
            with torch.cuda.device(self.dev):
                return fn(self, *a, **k)
        return wrapped

    @_on_device
    def step(self, images, labels, train=True):
        """fwd + CE + bwd + grad all-reduce + AdamW on device-resident inputs; returns the (device) loss tensor."""
        if images.shape != self._in[0][0].shape:
            raise RuntimeError(f"Sizes of tensors must match: expected image batch {tuple(self._in[0][0].shape)}, got "
                               f"{tuple(images.shape)} (cls_tkn_embd is per batch slot, reference vit.py:31-33,41)")
        d = self._dims(train)
        self.forward(images, d)
        self.backward(labels, d)
        self.optimizer_step()
        return self.loss

    @_on_device
    def step_host(self, images_host, labels_host):
        """Same step from PINNED host tensors: H2D of this step's inputs on a copy stream (double-buffered, so it
        overlaps the previous step's compute) and an async D2H of the loss.  -> index into self.loss_host."""
        k = self._slot
        self._slot ^= 1
        x, y = self._in[k]
        cur = torch.cuda.current_stream(self.dev)
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self._in_free[k])
            x.copy_(images_host, non_blocking=True)
            y.copy_(labels_host, non_blocking=True)
            self._in_ready[k].record(self.copy_stream)
        cur.wait_event(self._in_ready[k])
        d = self._dims(True)
        self.forward(x, d)
        self.backward(y, d)
        self._in_free[k].record(cur)  # the labels are read by the loss kernel at the top of backward
        self.optimizer_step()
        li = (self.step_no - 1) % self.loss_host.numel()
        self.loss_host[li:li + 1].copy_(self.loss, non_blocking=True)
        self._loss_ev[li].record(cur)
        return li

    @_on_device
    def step_host_u8(self, images_host_u8, labels_host, layout=0):
        """step_host fed with raw uint8 pixels: pinned (B,C,H,W) [layout 0] or (B,H,W,C) [layout 1]
        uint8 -> one H2D of 1 byte per pixel, the ToTensor /255 (train.py:151-155) fused into the patch gather."""
        if self._in_u8 is None:
            self._in_u8 = [torch.empty(self.B * self.C * self.img * self.img, dtype=torch.uint8, device=self.dev)
                           for _ in range(2)]
        if images_host_u8.dtype != torch.uint8 or images_host_u8.numel() != self._in_u8[0].numel():
            raise RuntimeError(f"expected {self._in_u8[0].numel()} uint8 pixels, got {images_host_u8.dtype} x "
                               f"{images_host_u8.numel()}")
        k = self._slot
        self._slot ^= 1
        x, y = self._in_u8[k], self._in[k][1]
        cur = torch.cuda.current_stream(self.dev)
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self._in_free[k])
            x.copy_(images_host_u8.reshape(-1), non_blocking=True)
            y.copy_(labels_host, non_blocking=True)
            self._in_ready[k].record(self.copy_stream)
        cur.wait_event(self._in_ready[k])
        d = self._dims(True)
        self.forward(x, d, u8_layout=layout)
        self.backward(y, d)
        self._in_free[k].record(cur)
        self.optimizer_step()
        li = (self.step_no - 1) % self.loss_host.numel()
        self.loss_host[li:li + 1].copy_(self.loss, non_blocking=True)
        self._loss_ev[li].record(cur)
        return li

    # ------------------------------------------------------------------------------------------ CUDA graph
    _SEED_STRIDE = 0xD1B54A32D192ED03       # per-step term of the dropout seed (see _dims)

    def capture_graph(self):
        """Capture one whole training step (fwd + CE + bwd + AdamW with fused zero_grad) into a CUDA graph — SURVEY
        8f-3 / reference train.py:92-96 as ONE graph launch.  Kernel arguments are frozen at capture, so everything
        that changes from step to step lives in a 32-byte device "step state" that the graph itself advances with its
        first node: the step counter, the per-step term of the dropout seed (kernels add it to their seed when they
        draw a mask) and AdamW's bias corrections.  A replay draws exactly the masks, and applies exactly the update,
        that the eager `step()` with the same step number would.  Single GPU only (the multi-GPU exchange lives on
        side streams); call it after at least one eager step, so that no first-use set-up lands inside the capture."""
        if self.world != 1:
            raise RuntimeError("capture_graph: single-GPU only (the data-parallel exchange runs on side streams)")
        if self.step_no < 1:
            raise RuntimeError("capture_graph: run at least one eager step() first")
        with torch.cuda.device(self.dev):
            self._gstate = torch.zeros(_C.STEP_STATE_BYTES, dtype=torch.uint8, device=self.dev)
            self._gx, self._gy = torch.empty_like(self._in[0][0]), torch.empty_like(self._in[0][1])   # static inputs
            base_seed = (self.seed * 0x9E3779B97F4A7C15 + self.rank * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
            d = self.plan.dims(self.dropout_p, base_seed, seed_dev=self._gstate.data_ptr())
            side = torch.cuda.Stream(device=self.dev)
            side.wait_stream(torch.cuda.current_stream(self.dev))
            graph = torch.cuda.CUDAGraph()
            # "relaxed": the C ABI calls cudaFuncSetAttribute / cuTensorMapEncodeTiled (host-only, no stream work) while
            # it enqueues; nothing in it allocates or synchronises
            with torch.cuda.graph(graph, stream=side, capture_error_mode="relaxed"):
                L = _C.lib()
                _C.check(L.vitb200_step_state_advance(self._gstate.data_ptr(), self.betas[0], self.betas[1],
                                                      self._SEED_STRIDE, _C.stream_ptr(self.dev)))
                self.forward(self._gx, d)
                self.backward(self._gy, d)
                h = _C.AdamWHyper(self.lr, self.betas[0], self.betas[1], self.eps, self.wd, 1, 1.0)
                _C.check(L.vitb200_adamw_range_dev(self.p.data_ptr(), self.g.data_ptr(), None, self.m.data_ptr(),
                                                   self.v.data_ptr(),
                                                   None if self.shadow is None else self.shadow.data_ptr(), self.total,
                                                   1, h, self._gstate.data_ptr(), _C.stream_ptr(self.dev)))
            torch.cuda.current_stream(self.dev).wait_stream(side)
            self._graph, self._graph_step = graph, -1
        return graph

    @_on_device
    def step_graph(self, images, labels):
        """Replay the captured step on device-resident inputs (copied into the graph's static buffers).  Interleaves
        freely with step(): the device step counter is re-synchronised whenever the host's moved in between."""
        if getattr(self, "_graph", None) is None:
            self.capture_graph()
        self._check_labels(labels)
        if self._graph_step != self.step_no:
            self._gstate[8:12].view(torch.int32).fill_(self.step_no)
        self._gx.copy_(images, non_blocking=True)
        self._gy.copy_(labels, non_blocking=True)
        self._graph.replay()
        self.step_no += 1
        self._graph_step = self.step_no
        return self.loss

    def read_loss(self, li):
        self._loss_ev[li].synchronize()
        return float(self.loss_host[li])

    # ------------------------------------------------------------------------------------------ accounting
    def flops_per_image(self):
        """Algorithmic training FLOPs per image = 3 x forward."""
        E, T, L, N = self.E, self.T, self.L, self.T - 1
        fwd = L * T * (24 * E * E + 4 * T * E) + N * 2 * self.C * self.P * self.P * E + 8 * E * E + 8 * E * self.classes
        return 3.0 * fwd

    def activation_bytes(self):
        pl = self.plan
        return (pl.embed_act_bytes + self.L * pl.block_act_bytes + (self.L + 2) * self.M * self.E * 4 +
                pl.head_act_bytes + max(pl.block_scratch_bytes, pl.head_act_bytes))
