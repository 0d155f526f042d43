"""Per-kernel count of the Blackwell mnemonics in the built library's SASS (`cuobjdump -sass`):
UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA bulk-tensor load / store, UTCBAR = tcgen05.commit.
Usage: python tools/sass_summary.py > profiles/sass_rN.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "tilelang_mesh_b200", "lib", "libtlb200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
pat = re.compile(r"\b(UTC[A-Z]*MMA(?:\.[A-Z0-9_]+)*|LDTM(?:\.[A-Za-z0-9_]+)*|STTM(?:\.[A-Za-z0-9_]+)*|UTMALDG(?:\.[A-Z0-9_]+)*|UTMASTG(?:\.[A-Z0-9_]+)*|UTCBAR(?:\.[A-Z0-9_]+)*|UTMAPF(?:\.[A-Z0-9_]+)*|SYNCS(?:\.[A-Z0-9_]+)*|ELECT|UTCATOMSWS(?:\.[A-Z0-9_]+)*)")
print("# cuobjdump -sass tilelang_mesh_b200/lib/libtlb200.so  (sm_100a) - mnemonic counts per kernel")
for i, chunk in enumerate(sass.split("Function : ")[1:]):
    head = chunk.split("\n", 1)[0].strip()
    cnt = collections.Counter(pat.findall(chunk))
    n_instr = len(re.findall(r"^\s+/\*[0-9a-f]{4,}\*/", chunk, flags=re.M))
    nm = names[i] if i < len(names) else head
    nm = re.sub(r"\((?:tlb200::)?\w+\)$|\([^()]*\)$", "", nm.strip())
    if not any(k.startswith(("UTC", "LDTM", "UTMA")) for k in cnt):
        print(f"{nm}: {n_instr} instructions, no tensor-core / TMA instructions")
        continue
    print(f"{nm}: {n_instr} instructions")
    for k in sorted(cnt):
        if k.startswith("SYNCS") or k == "ELECT":
            continue
        print(f"    {k:40s} {cnt[k]}")
