"""Per-tile timeline of CTA 0 from a -DTLB200_TRACE build of the library (clock64 stamps, cycles):
  TLB200_LIB=tools/_ab/libtlb200_trace.so python tools/trace_tiles.py [m n k] [fp8]
Build: nvcc ... -DTLB200_TRACE -o tools/_ab/libtlb200_trace.so tilelang_mesh_b200/csrc/tlb200.cu"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import _abi, gemm as G
args = [a for a in sys.argv[1:] if a.isdigit()]
m, n, k = (int(x) for x in args[:3]) if len(args) >= 3 else (8192, 8192, 8192)
dt = torch.float8_e4m3fn if "fp8" in sys.argv else torch.bfloat16
cfg = {}
for a in sys.argv[1:]:
    if "=" in a:
        kk, v = a.split("="); cfg[kk] = int(v)
debug_mode = cfg.pop("debug", 0)  # development knobs of the traced build (GemmKernelParams::debug_mode); the product is garbage
lib = _abi.load()
lib.tlb200_debug_set_trace.argtypes = [ctypes.c_void_p]
a = torch.randn(m, k, device="cuda").to(dt); b = torch.randn(n, k, device="cuda").to(dt)
c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
tr = torch.zeros(256 * 16, dtype=torch.int64, device="cuda")
for _ in range(3): G.gemm(a, b, c, **cfg)
torch.cuda.synchronize()
if "cold" in sys.argv:  # read 512 MB first: the traced call finds nothing of its operands in L2
    torch.empty(512 << 20, dtype=torch.uint8, device="cuda").view(torch.int64).sum().item()
lib.tlb200_debug_set_trace(tr.data_ptr())
lib.tlb200_debug_set_mode(debug_mode)
G.gemm(a, b, c, **cfg)
torch.cuda.synchronize()
lib.tlb200_debug_set_trace(None)
lib.tlb200_debug_set_mode(0)
t = tr.cpu().view(-1, 16)
if "small" in sys.argv:
    r = t[0].tolist(); z = r[13]
    names = {13: "kernel entry", 14: "set-up done (barriers, TMEM, cluster sync)", 11: "producer issues first K block", 0: "MMA warp holds accumulator",
             12: "producer warp 0 has issued its 3rd K block (blocks 0, 3, 6; no stage was waited for)", 1: "first K block landed", 2: "last commit issued", 3: "epilogue sees accumulator full", 4: "epilogue: accumulator handed back", 15: "epilogue warp done"}
    for k in sorted(names, key=lambda k: r[k]):
        print(f"{r[k]-z:8d} cycles  {names[k]}")
    if r[9]:  # -DTLB200_TRACE_WAIT builds only (two clock reads per K block perturb the loop they measure)
        print(f"MMA warp waited for operand data {r[9]} cycles of the {r[2]-r[1]} between the first K block and the last commit")
    else:
        print(f"main loop: {r[2]-r[1]} cycles between the first K block landing and the last commit")
    q = t[200].tolist()
    if q[0]:
        print(f"producer warp 0 after set-up: +{q[0]-r[14]} role dispatch, +{q[1]-q[0]} L2 policy, +{q[2]-q[1]} griddepcontrol.wait, +{q[3]-q[2]} tile coordinates, +{r[11]-q[3]} K range -> first issue")
    print("K block | producer: starts waiting for the stage, +wait, +expect_tx, +A issue, +B issue | MMA thread sees the operands (after the B issue)")
    for kb in range(64):
        q = t[64 + kb].tolist()
        if not q[0]:
            continue
        print(f"    {kb:2d} | {q[4]-z:7d} {q[0]-q[4]:6d} {q[1]-q[0]:5d} {q[2]-q[1]:5d} {q[3]-q[2]:5d} | {q[8]-z:7d} ({q[8]-q[3]:6d})")
    print("plan", G.plan_for(a, b, c, **cfg)); sys.exit(0)
print("plan", G.plan_for(a, b, c, **cfg))
rows = [r for r in t.tolist() if r[2] != 0]
t0 = rows[0][0]
print("tile | acc free->1st kb | mainloop (1st kb->last commit) | commit->epi0 sees full | epi0 acc0 drain | epi0 acc1 wait | acc1 drain | epi4: acc0 drain, acc1 drain | last drain -> next tile's MMA holds acc | producer lead (1st issue before acc free)")
for i, r in enumerate(rows):
    nxt = rows[i + 1][0] if i + 1 < len(rows) else 0
    last_done = max(r[6], r[10])
    print(f"{i:3d} | {r[1]-r[0]:6d} | {r[2]-r[1]:7d} | {r[3]-r[2]:6d} | {r[4]-r[3]:6d} | {r[5]-r[4]:6d} | {r[6]-r[5]:6d} | {r[8]-r[7]:6d} {r[10]-r[9]:6d} | {(nxt-last_done) if nxt else 0:6d} | {r[0]-r[11]:7d}   exposed(commit->next acc free) {(nxt - r[2]) if nxt else 0}"
          f" | epi0 acc0, 2nd slab: full->ld done {r[12]-r[3]}, ->store-read wait done {r[13]-r[12]}, ->st.shared+fence {r[14]-r[13]}, ->store issued {r[15]-r[14]}")
