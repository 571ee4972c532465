#!/bin/bash
# Which Blackwell-native instructions each kernel of libb200parallm.so contains (cuobjdump -sass, counted per function):
# UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM / STTM = tcgen05.ld / .st, UTMALDG / UTMASTG / UTMAREDG = TMA tensor
# load / store / reduce-add, SYNCS = mbarrier ops, HMMA = legacy mma.sync, LDGSTS = cp.async, REDUX / MATCH etc. not listed.
LIB=${1:-mlx_parallm_b200/libb200parallm.so}
python3 - "$LIB" <<'PY'
import re, subprocess, sys, collections
out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
names = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "SYNCS", "HMMA", "LDGSTS"]
cnt = collections.OrderedDict()
fn = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        cnt.setdefault(fn, collections.Counter())
        continue
    if fn is None:
        continue
    for n in names:
        if re.search(r"\s" + n + r"[.\s]", line):
            cnt[fn][n] += 1
dem = subprocess.run(["cu++filt"] + list(cnt), capture_output=True, text=True).stdout.splitlines()
print(f"{'kernel':95s} " + " ".join(f"{n:>8s}" for n in names))
tot = collections.Counter()
for (fn, c), d in zip(cnt.items(), dem):
    d = d.split(">(")[0].replace("b200p::", "") + (">" if ">(" in d else "")
    if not any(c.values()):
        continue
    tot.update(c)
    print(f"{d[:95]:95s} " + " ".join(f"{c[n]:8d}" for n in names))
print(f"{'TOTAL':95s} " + " ".join(f"{tot[n]:8d}" for n in names))
PY
