"""Aggregate an `ncu --page source --csv --print-source sass` dump by device function.
usage: python tools/ncu_funcs.py sass.csv kernels.o 'stepKernelILi8E'
Function boundaries come from the local FUNC symbols of the kernel's .text section (cuobjdump -elf)."""
import csv, sys, subprocess, re, collections
sass_csv, obj, kpat = sys.argv[1:4]
elf = subprocess.run(["cuobjdump", "-elf", obj], capture_output=True, text=True).stdout
# find section index of the kernel text
secidx = None
for l in elf.split("\n"):
    a = l.split()
    if len(a) >= 9 and a[-1].startswith(".text.") and kpat in a[-1] and "PROGBITS" in l:
        secidx = int(a[0], 16)
syms = []
for l in elf.split("\n"):
    a = l.split()
    # symbol table rows: idx value size info other shndx name
    if len(a) == 7 and a[1].startswith("0x"):
        try:
            val, size, shndx = int(a[1], 16), int(a[2], 16), int(a[5], 16)
        except ValueError:
            continue
        if shndx == secidx and size > 0:
            syms.append((val, size, a[6]))
syms.sort()
rows = list(csv.reader(open(sass_csv)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hdr]
isamp, iinst = H.index("# Samples"), H.index("Instructions Executed")
ithr = H.index("Thread Instructions Executed")
stall = {h: i for i, h in enumerate(H) if h.startswith("stall_") and "Not Issued" not in h}
base = int(rows[hdr + 1][0], 16)
agg = collections.defaultdict(lambda: [0, 0, 0, collections.Counter(), 0])
def fn(off):
    name = "kernel"
    for v, s, n in syms:
        if v <= off < v + s: name = n
    return name
for r in rows[hdr + 1:]:
    try: off = int(r[0], 16) - base
    except ValueError: continue
    f = fn(off)
    a = agg[f]
    a[0] += int(r[isamp]); a[1] += int(r[iinst]); a[2] += 1; a[4] += int(r[ithr])
    for h, i in stall.items():
        a[3][h] += int(r[i])
ts = sum(a[0] for a in agg.values()); ti = sum(a[1] for a in agg.values())
print(f"total samples {ts}  warp-instructions {ti}  sass {sum(a[2] for a in agg.values())}")
for f, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    top = ", ".join(f"{k[6:]} {100*v/max(1,a[0]):.0f}%" for k, v in a[3].most_common(4))
    nm = re.sub(r"^.*_cu_[0-9a-f]{8}", "", f)[:40]
    print(f"{100*a[0]/ts:5.1f}%s {100*a[1]/ti:5.1f}%i thr={a[4]/max(1,a[1]):4.1f} sass={a[2]:5d}  {nm:42s} {top}")
