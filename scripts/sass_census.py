"""SASS census of libfi_b200.so: per kernel symbol, how many of the instructions that prove the
Blackwell data path are really in the binary.  Runs without a GPU (cuobjdump only).

    python scripts/sass_census.py            # writes profiles/sass_census.txt

Mnemonics (B200_PROFILING.md): UTMALDG = TMA tensor load (cp.async.bulk.tensor), UTCQMMA / UTCHMMA =
tcgen05.mma (kind::f8f6f4 / f16), LDTM / STTM = tcgen05.ld / st (TMEM), UTCBAR = tcgen05.commit,
HMMA / QMMA = mma.sync, LDSM = ldmatrix, MOVM = movmatrix, SYNCS = mbarrier ops, LDGSTS = cp.async,
ACQBULK / UBLKCP = bulk copies, REDG / ATOMG = global reductions / atomics, MULTIMEM via "MULTIMEM" /
".MMA" spellings are not used: the NVLS store shows up as ST.E with the .MULTIMEM? suffix (listed
under `multimem`)."""
import re
import subprocess
import sys
from collections import OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "flashinfer_b200" / "libfi_b200.so"
PATTERNS = OrderedDict([
    ("UTMALDG", r"\bUTMALDG"), ("UTMAPF", r"\bUTMAPF"), ("UTCQMMA", r"\bUTCQMMA"), ("UTCHMMA", r"\bUTCHMMA"),
    ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UTCBAR", r"\bUTCBAR"), ("HMMA", r"\bHMMA"),
    ("QMMA", r"\bQMMA"), ("LDSM", r"\bLDSM"), ("MOVM", r"\bMOVM"), ("SYNCS", r"\bSYNCS"),
    ("LDGSTS", r"\bLDGSTS"), ("ATOMG", r"\bATOMG"), ("multimem", r"MULTIMEM|\.MMEM"),
    ("ACQBULK", r"\bACQBULK"), ("MEMBAR", r"\bMEMBAR"),
])


def demangle(names):
    r = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True)
    return r.stdout.splitlines() if r.returncode == 0 else names


def main():
    lib = Path(sys.argv[1]) if len(sys.argv) > 1 else LIB
    sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
    kernels, cur = OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = {"_insts": 0, **{k: 0 for k in PATTERNS}}
            continue
        if cur is None or "/*" not in line:
            continue
        if re.search(r"/\*[0-9a-f]{4,}\*/", line):
            kernels[cur]["_insts"] += 1
            for k, pat in PATTERNS.items():
                if re.search(pat, line):
                    kernels[cur][k] += 1
    names = demangle(list(kernels))
    cols = [k for k in PATTERNS if any(v[k] for v in kernels.values())]
    out = [f"SASS census of {lib.name} ({len(kernels)} kernels, sm_100a; cuobjdump -sass | per-symbol mnemonic counts)",
           "", "insts " + " ".join(f"{c:>8s}" for c in cols) + "  kernel"]
    tot = {c: 0 for c in cols}
    for (sym, c), name in sorted(zip(kernels.items(), names), key=lambda t: t[1]):
        short = re.sub(r"\(fib200::\w+(, CUtensorMap_st, CUtensorMap_st)?\)|\(.*\)$", "", name).replace("void ", "")
        out.append(f"{c['_insts']:5d} " + " ".join(f"{c[k]:8d}" for k in cols) + f"  {short}")
        for k in cols:
            tot[k] += c[k]
    out.append("total " + " ".join(f"{tot[k]:8d}" for k in cols))
    text = "\n".join(out) + "\n"
    (ROOT / "profiles" / "sass_census.txt").write_text(text)
    print(text)


if __name__ == "__main__":
    main()
