"""Format gpurun_out/r02_parity_table.jsonl (written by tests/test_reference_gpu.py on the GPU box) into the committed
profiles/r02_parity_table.txt.  usage: python tools/parity_table.py [in.jsonl] > profiles/r02_parity_table.txt"""
import json
import sys

path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/r02_parity_table.jsonl"
rows = [json.loads(l) for l in open(path) if l.strip()]
print("# Parity of the CUDA path against the REAL reference (oracle/_ref/py, imported on the same B200), written by")
print("# tests/test_reference_gpu.py.  'ours' = drop-in package through the C ABI; 'autocast ref' = the reference's own")
print("# modules under torch.autocast(bf16); both measured against the reference in fp32 (fp64 for the fp32 row).")
print("# rel = max-norm relative error of the logits; cos = gradient cosine per reference state_dict key (min / mean).")
print()
f = lambda v: f"{v:.3e}" if isinstance(v, float) and abs(v) < 1e-2 else (f"{v:.5f}" if isinstance(v, float) else str(v))
for r in rows:
    c = r.get("config", "?")
    if "ours_worst_grad_l2" in r:
        print(f"{c}\n    logits rel: ours {f(r['ours_logits_rel'])}  | reference fp32 {f(r['ref_fp32_logits_rel'])}   (vs reference float64)")
        print(f"    worst gradient rel-L2: ours {f(r['ours_worst_grad_l2'])} ({r['ours_worst_grad_l2_key']}) | reference fp32 {f(r['ref_fp32_worst_grad_l2'])}")
        print(f"    worst gradient max-norm: ours {f(r['ours_worst_grad_maxnorm'])} ({r['ours_worst_grad_maxnorm_key']}) | reference fp32 {f(r['ref_fp32_worst_grad_maxnorm'])}")
        print(f"    gate: {r['gate']}")
    elif "ours_min_grad_cos" in r:
        print(f"{c}   [{r['weights']}]")
        print(f"    logits rel: ours {f(r['ours_logits_rel'])} | autocast ref {f(r['autocast_ref_logits_rel'])}")
        print(f"    min cos: ours {f(r['ours_min_grad_cos'])} ({r['ours_min_grad_cos_key']}) | autocast ref {f(r['autocast_ref_min_grad_cos'])}")
        print(f"    mean cos: ours {f(r['ours_mean_grad_cos'])} | autocast ref {f(r['autocast_ref_mean_grad_cos'])}")
        print(f"    gate: {r['gate']}")
    elif "per_tensor_outliers_ours" in r:
        print(f"{c}   per-tensor outliers (deficit > max(1e-3, 3x the other's)): ours {r['per_tensor_outliers_ours']} | autocast ref "
              f"{r['per_tensor_outliers_autocast_ref']} of {r['tensors']} tensors; gross (ours) {r['gross_outliers_ours']}")
    else:
        print(c + "\n    " + "\n    ".join(f"{k}: {v}" for k, v in r.items() if k != "config"))
    print()
