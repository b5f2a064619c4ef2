"""Checkpoint compatibility with the reference.

The reference saves `{'epoch', 'model_state_dict', 'optimizer_state_dict', 'loss', 'step'}` (src/train.py:107-113) and
resumes with `model.load_state_dict` + `optimizer.load_state_dict` (src/train.py:68-75).  Model state_dicts already
interchange key for key (the packed q/k/v parameter is split into the reference's per-head keys on save and
re-packed on load).  `torch.optim.AdamW.state_dict()` however is indexed by POSITION in `model.parameters()`:
the reference has 3H per-head matrices per block where this package has ONE packed `[3E, E]` parameter.  This
module translates the optimizer state between the two parameter lists, losslessly:

    reference list : `reference_param_keys(L, H)`   (4 + L*(3H+10) + 6 entries, pinned against the real reference's
                                                     state_dict order by tests/test_host.py)
    packed list    : `packed_param_keys(L)`         (4 + L*11 + 6 entries = named_parameters() of this package)

Pure host-side tensor bookkeeping: no kernels, works on CPU tensors (checkpoints are loaded on the host).

    python -m VisionTransformer.checkpoint to-packed    ref.pt  ours.pt  --heads 12
    python -m VisionTransformer.checkpoint to-reference ours.pt ref.pt   --heads 12
"""
import argparse

import torch

_SECTIONS = ("query", "key", "value")      # row blocks of the packed [3E, E] parameter
_REF_HEAD_ORDER = ("key", "query", "value")  # registration order inside the reference's Head (transformer.py:12-17)


def reference_param_keys(num_blocks, num_heads):
    """Names in the order of the reference's `model.parameters()` (vit.py:47-80, transformer.py:10-90)."""
    keys = ["emdeddings.cls_tkn_embd", "emdeddings.pos_embd", "emdeddings.sequence.0.weight",
            "emdeddings.sequence.0.bias"]
    for i in range(num_blocks):
        b = f"transformer_encoder.blocks.{i}."
        for h in range(num_heads):
            keys += [f"{b}multi_head.heads.{h}.{n}.weight" for n in _REF_HEAD_ORDER]
        keys += [b + "multi_head.proj.weight", b + "multi_head.proj.bias", b + "ffwd.mlp.0.weight", b + "ffwd.mlp.0.bias",
                 b + "ffwd.mlp.2.weight", b + "ffwd.mlp.2.bias", b + "ln1.weight", b + "ln1.bias", b + "ln2.weight",
                 b + "ln2.bias"]
    keys += ["mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias", "mlp.3.weight", "mlp.3.bias"]
    return keys


def packed_param_keys(num_blocks):
    """Names in the order of this package's `model.named_parameters()` (qkv_weight registered before proj)."""
    keys = ["emdeddings.cls_tkn_embd", "emdeddings.pos_embd", "emdeddings.sequence.0.weight",
            "emdeddings.sequence.0.bias"]
    for i in range(num_blocks):
        b = f"transformer_encoder.blocks.{i}."
        keys += [b + "multi_head.qkv_weight", b + "multi_head.proj.weight", b + "multi_head.proj.bias",
                 b + "ffwd.mlp.0.weight", b + "ffwd.mlp.0.bias", b + "ffwd.mlp.2.weight", b + "ffwd.mlp.2.bias",
                 b + "ln1.weight", b + "ln1.bias", b + "ln2.weight", b + "ln2.bias"]
    keys += ["mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias", "mlp.3.weight", "mlp.3.bias"]
    return keys


def _head_rows(key, num_heads, head_size):
    """'...multi_head.heads.<h>.<name>.weight' -> (packed key, first row) of its slice in the packed parameter."""
    pre, rest = key.split("multi_head.heads.")
    h, name, _ = rest.split(".")
    return pre + "multi_head.qkv_weight", _SECTIONS.index(name) * num_heads * head_size + int(h) * head_size


def _infer_counts(n_params, num_heads, packed):
    per_block = 11 if packed else 3 * num_heads + 10
    if (n_params - 10) % per_block != 0 or n_params < 10:
        raise ValueError(f"optimizer state holds {n_params} parameters: not a "
                         f"{'packed' if packed else 'reference'} ViT parameter list with {num_heads} heads")
    return (n_params - 10) // per_block


def _group_template(opt_sd, n_params):
    groups = opt_sd.get("param_groups", [])
    if len(groups) != 1:
        raise ValueError("expected the single AdamW parameter group of src/train.py:66")
    g = dict(groups[0])
    g["params"] = list(range(n_params))
    return g


def optimizer_state_to_packed(ref_opt_sd, num_heads):
    """AdamW state_dict over the REFERENCE parameter list -> over this package's packed list.  Per-head moments are
    stacked into the packed rows; their step counters must agree (they always do: all parameters step together)."""
    n_ref = len(ref_opt_sd["param_groups"][0]["params"])
    L = _infer_counts(n_ref, num_heads, packed=False)
    ref_keys, pk_keys = reference_param_keys(L, num_heads), packed_param_keys(L)
    ref_index = {k: i for i, k in enumerate(ref_keys)}
    state = ref_opt_sd.get("state", {})
    out_state = {}
    for j, k in enumerate(pk_keys):
        if not k.endswith("multi_head.qkv_weight"):
            if ref_index[k] in state:
                out_state[j] = {n: (v.clone() if torch.is_tensor(v) else v) for n, v in state[ref_index[k]].items()}
            continue
        pre = k[:-len("qkv_weight")]
        parts = {(h, name): state.get(ref_index[f"{pre}heads.{h}.{name}.weight"])
                 for h in range(num_heads) for name in _SECTIONS}
        present = [p for p in parts.values() if p is not None]
        if not present:
            continue
        if len(present) != len(parts):
            raise ValueError(f"{pre}: optimizer state present for only some heads")
        steps = {float(p["step"]) for p in present}
        if len(steps) != 1:
            raise ValueError(f"{pre}: per-head step counters differ ({sorted(steps)})")
        entry = {"step": present[0]["step"].clone() if torch.is_tensor(present[0]["step"]) else present[0]["step"]}
        for mom in (n for n in present[0] if n != "step"):
            entry[mom] = torch.cat([parts[(h, name)][mom] for name in _SECTIONS for h in range(num_heads)], dim=0)
        out_state[j] = entry
    return {"state": out_state, "param_groups": [_group_template(ref_opt_sd, len(pk_keys))]}


def optimizer_state_to_reference(packed_opt_sd, num_heads):
    """AdamW state_dict over this package's packed parameter list -> over the reference's list (what the reference's
    `optimizer.load_state_dict` at src/train.py:73 expects)."""
    n_pk = len(packed_opt_sd["param_groups"][0]["params"])
    L = _infer_counts(n_pk, num_heads, packed=True)
    ref_keys, pk_keys = reference_param_keys(L, num_heads), packed_param_keys(L)
    pk_index = {k: i for i, k in enumerate(pk_keys)}
    state = packed_opt_sd.get("state", {})
    out_state = {}
    for i, k in enumerate(ref_keys):
        if "multi_head.heads." not in k:
            if pk_index[k] in state:
                out_state[i] = {n: (v.clone() if torch.is_tensor(v) else v) for n, v in state[pk_index[k]].items()}
            continue
        st = state.get(pk_index[_head_rows(k, num_heads, 1)[0]])
        if st is None:
            continue
        rows = next(v for n, v in st.items() if n != "step").shape[0]
        hs = rows // (3 * num_heads)
        _, r0 = _head_rows(k, num_heads, hs)
        entry = {"step": st["step"].clone() if torch.is_tensor(st["step"]) else st["step"]}
        for mom, v in st.items():
            if mom != "step":
                entry[mom] = v[r0:r0 + hs].clone()
        out_state[i] = entry
    return {"state": out_state, "param_groups": [_group_template(packed_opt_sd, len(ref_keys))]}


def convert_checkpoint(ckpt, num_heads, to):
    """A whole `torch.save`d dict of src/train.py:107-113.  The model state_dict already uses the reference's keys in
    both directions; only 'optimizer_state_dict' changes."""
    out = dict(ckpt)
    if "optimizer_state_dict" in ckpt:
        fn = optimizer_state_to_packed if to == "packed" else optimizer_state_to_reference
        out["optimizer_state_dict"] = fn(ckpt["optimizer_state_dict"], num_heads)
    return out


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("direction", choices=["to-packed", "to-reference"])
    ap.add_argument("src")
    ap.add_argument("dst")
    ap.add_argument("--heads", type=int, required=True)
    a = ap.parse_args(argv)
    ckpt = torch.load(a.src, map_location="cpu", weights_only=False)
    torch.save(convert_checkpoint(ckpt, a.heads, "packed" if a.direction == "to-packed" else "reference"), a.dst)


if __name__ == "__main__":
    main()
