"""Instruction mix of one device function inside an object: python tools/sass_mix.py obj.o funcname"""
import re, subprocess, sys, collections
obj, fn = sys.argv[1], sys.argv[2]
elf = subprocess.run(["cuobjdump", "-elf", obj], capture_output=True, text=True).stdout
start = size = None
for l in elf.split("\n"):
    a = l.split()
    if len(a) >= 7 and a[2].startswith("0x") and fn in a[-1] and "stepKernel" not in a[-1][:20]:
        start, size = int(a[1], 16), int(a[2], 16)
sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
mix = collections.Counter(); n = 0
for l in sass.split("\n"):
    m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(.*?);", l)
    if not m: continue
    adr = int(m.group(1), 16)
    if start <= adr < start + size:
        ins = m.group(2).split()
        op = ins[1] if ins[0].startswith("@") else ins[0]
        mix[op.split(".")[0]] += 1; n += 1
print(fn, "start", hex(start), "bytes", size, "instructions", n)
print(mix.most_common(25))
