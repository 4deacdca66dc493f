"""Per-kernel census of the Blackwell-specific SASS in libuuomocap.so -> profiles/<tag>_sass_census.txt
    python tools/sass_census.py r2      (cuobjdump only: no GPU needed)"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "rX"
lib = os.path.join(ROOT, "uuo-mocap_b200", "lib", "libuuomocap.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
WANT = [("UTCHMMA", "tcgen05.mma kind::tf32"), ("LDTM", "tcgen05.ld"), ("UBLKCP", "cp.async.bulk (1-D bulk copy)"),
        ("UTMALDG", "cp.async.bulk.tensor (TMA tile load)"), ("SYNCS", "mbarrier"), ("CREDUX", "redux.sync.{min,max}.f32"),
        ("REDUX", "redux.sync (int)"), ("LDGSTS", "cp.async"), ("FFMA2", "fma.rn.f32x2"), ("FMUL2", "mul.rn.f32x2"),
        ("FADD2", "add/sub.rn.f32x2"), ("MUFU.RSQ", "rsqrt.approx"), ("HMMA", "legacy mma.sync (should be 0)"),
        ("HGMMA", "wgmma (should be 0)")]
per = collections.OrderedDict()
cur = None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        cur = per.setdefault(name, collections.Counter())
        continue
    if cur is None:
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        cur["_n"] += 1
        for key, _ in WANT:
            if op.startswith(key) and not (key == "REDUX" and op.startswith("CREDUX")):
                cur[key] += 1
arch = re.findall(r"arch = (sm_\w+)", txt)
out = ["# SASS census of uuo-mocap_b200/lib/libuuomocap.so (%s): cuobjdump -sass, instruction counts per kernel" % tag,
       "# arch: %s; %d kernels" % (", ".join(sorted(set(arch))), len(per)),
       "# " + "; ".join("%s = %s" % kv for kv in WANT), ""]
cols = [k for k, _ in WANT]
out.append("%-44s %6s " % ("kernel", "instr") + " ".join("%8s" % c for c in cols))
tot = collections.Counter()
for name, c in per.items():
    short = re.sub(r"^void |\(anonymous namespace\)::|<unnamed>::", "", name)[:44]
    out.append("%-44s %6d " % (short, c["_n"]) + " ".join("%8d" % c[k] for k in cols))
    tot.update(c)
out.append("%-44s %6d " % ("TOTAL", tot["_n"]) + " ".join("%8d" % tot[k] for k in cols))
os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
open(os.path.join(ROOT, "profiles", "%s_sass_census.txt" % tag), "w").write("\n".join(out) + "\n")
print("\n".join(out[:6] + out[-3:]))
