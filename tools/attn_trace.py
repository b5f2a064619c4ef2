"""Event trace of CTA 0 of the tcgen05 attention-backward kernels (clock64 stamps written by the kernel when
VITB200_ATTN_TRACE holds a device address).  Prints, per role, the events of the first items with deltas."""
import math, os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, os.path.join(ROOT, "vision-transformer_b200"))
import torch
from VisionTransformer import _C

B, T, H, hs = 256, 197, 12, 64
E = H * hs
L = _C.lib()
scale = math.sqrt(hs)
qkv = (torch.randn(B * T, 3 * E, device="cuda") * 0.05).bfloat16()
out = torch.empty(B * T, E, device="cuda", dtype=torch.bfloat16)
lse = torch.empty(B, H, T, device="cuda")
d_o = torch.randn(B * T, E, device="cuda").bfloat16()
dqkv = torch.empty_like(qkv)
delta = torch.empty(B, H, T, device="cuda")
_C.check(L.vitb200_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), _C.BF16, B, T, H, hs, scale, _C.stream_ptr()))
bwd = lambda: _C.check(L.vitb200_attn_bwd(qkv.data_ptr(), out.data_ptr(), d_o.data_ptr(), lse.data_ptr(), delta.data_ptr(), dqkv.data_ptr(), _C.BF16, B, T, H, hs, scale, _C.stream_ptr()))
for _ in range(3):
    bwd()
trace = torch.zeros(4096, dtype=torch.int64, device="cuda")
os.environ["VITB200_ATTN_TRACE"] = hex(trace.data_ptr())
bwd()
torch.cuda.synchronize()
del os.environ["VITB200_ATTN_TRACE"]
tr = trace.cpu().tolist()
names = {1: "mma: wait slot_free", 2: "mma: slot free", 3: "mma: sd issued", 4: "mma: wait ew_done", 5: "mma: ew done",
         6: "mma: out issued", 7: "mma: wait kv_full", 8: "mma: kv full", 10: "ew: wait sd_full", 11: "ew: sd full",
         12: "ew: chunk done", 14: "ew: wait acc_full", 15: "ew: acc full", 16: "ew: end"}
for mode in (1, 0):
    base = mode * 2048
    evs = []
    for role in range(4):
        for v in tr[base + role * 512: base + (role + 1) * 512]:
            if v:
                evs.append((v >> 8, role, v & 255))
    evs.sort()
    t0 = evs[0][0]
    print(f"==== MODE {mode}: {len(evs)} events; first 120 (clk since start, role, event)")
    last = {0: t0, 1: t0, 2: t0, 3: t0}
    for clk, role, tag in evs[:int(os.environ.get('NEV', 120))]:
        print(f"{clk - t0:8d}  +{clk - last[role]:6d}  r{role}  {names.get(tag, tag)}")
        last[role] = clk
    # per-item period for the MMA thread: time between consecutive 'kv full'
    kv = [c for c, r, t in evs if r == 0 and t == 8]
    print("item periods (clk):", [b - a for a, b in zip(kv, kv[1:])][:12])
