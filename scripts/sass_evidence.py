"""SASS evidence for the Blackwell-native kernels: mnemonic histogram per kernel + the tcgen05 / TMEM lines.
python scripts/sass_evidence.py > profiles/rN_sass_evidence.md   (needs only cuobjdump, no GPU)"""
import collections, os, re, subprocess, sys
LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "nerf_cuda_b200", "libnerf_b200.so")
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
kern, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.match(r"\s+Function : (\S+)", line)
    if m:
        cur = m.group(1); kern[cur] = []
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
    if m and cur:
        kern[cur].append((m.group(1), m.group(2).strip()))
print("# SASS evidence (cuobjdump -sass nerf_cuda_b200/libnerf_b200.so, sm_100a)\n")
print("`UTCHMMA` = tcgen05.mma kind::f16, `LDTM`/`STTM` = tcgen05.ld/st, `UTCBAR` = tcgen05.commit, `UTCATOMSWS` = tcgen05.alloc/dealloc "
      "(B200_PROFILING.md, 'What proves a Blackwell-native kernel').\n")
KEY = ("UTC", "LDTM", "STTM", "SYNCS", "FMUL2", "HMMA")
for k, ins in kern.items():
    if not ins:
        continue
    hist = collections.Counter()
    for _, t in ins:
        t = re.sub(r"^@!?U?P\d+\s+", "", t)
        hist[t.split()[0].split(".")[0]] += 1
    top = ", ".join(f"{m} {c}" for m, c in hist.most_common(14))
    print(f"## `{k}` — {len(ins)} instructions\n\n{top}\n")
    key = [(a, t) for a, t in ins if any(x in t for x in KEY)]
    if any("UTC" in t or "LDTM" in t for _, t in key):
        cnt = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0] for _, t in key)
        print("tensor-core / TMEM instructions: " + ", ".join(f"`{m}` x{c}" for m, c in sorted(cnt.items())) + "\n")
        print("```")
        shown = 0
        for a, t in key:
            if ("UTC" in t or "LDTM" in t or "STTM" in t) and shown < 40:
                print(f"/*{a}*/ {t}")
                shown += 1
        print("```\n")
