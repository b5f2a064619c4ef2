"""Event trace of CTA 0 of the tcgen05 attention-FORWARD kernel (clock64 stamps written by the kernel when
VITB200_ATTN_TRACE holds a device address).  Prints per role the time spent between events, and the first events."""
import math, os, sys, collections
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
fwd = lambda: _C.check(L.vitb200_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), _C.BF16, B, T, H, hs, scale, _C.stream_ptr()))
for _ in range(3):
    fwd()
trace = torch.zeros(4096, dtype=torch.int64, device="cuda")
os.environ["VITB200_ATTN_TRACE"] = hex(trace.data_ptr())
fwd()
torch.cuda.synchronize()
tr = trace.cpu().tolist()
names = {1: "mma: wait t_free", 2: "mma: t_free ok", 3: "mma: S issued", 4: "mma: wait p_ready", 5: "mma: p_ready ok",
         6: "mma: PV issued", 7: "mma: wait kv_full", 8: "mma: kv_full ok", 10: "ew: wait s_full", 11: "ew: s_full ok",
         12: "ew: pass1 done", 13: "ew: P written", 14: "ew: o_full ok", 15: "ew: epilogue done"}
evs = []
for role in range(5):
    for v in tr[role * 512:(role + 1) * 512]:
        if v:
            evs.append((v >> 8, role, v & 255))
evs.sort()
t0 = evs[0][0]
rn = {0: "mma", 1: "ew t0 h0", 2: "ew t0 h1", 3: "ew t1 h0", 4: "ew t1 h1"}
seg = collections.defaultdict(list)
last = {}
for clk, role, tag in evs:
    if role in last:
        seg[(role, last[role][1], tag)].append(clk - last[role][0])
    last[role] = (clk, tag)
span = evs[-1][0] - t0
print(f"{len(evs)} events, span {span} clk")
for (role, a, b), v in sorted(seg.items()):
    v2 = v[2:] if len(v) > 4 else v      # skip the cold start
    print(f"  {rn[role]:9s} {names.get(a,a):18s} -> {names.get(b,b):18s} n={len(v):3d} median {sorted(v2)[len(v2)//2]:6d}  sum {sum(v):7d} ({100*sum(v)/span:4.1f}% of span)")
kv = [c for c, r, t in evs if r == 1 and t == 11]
print("unit periods of ew t0 h0 (clk):", [b - a for a, b in zip(kv, kv[1:])][:14])
for clk, role, tag in evs[:int(os.environ.get("NEV", 90))]:
    print(f"{clk - t0:8d}  {rn[role]:9s} {names.get(tag, tag)}")
