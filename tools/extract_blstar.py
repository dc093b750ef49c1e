"""Build-container tool: reads the BL* (Andronescu et al. 2010) 37 C energy
parameters out of the reference's src/boltzmann_param.c and writes them as a
flat little-endian int32 blob, dafs_b200/data/blstar37.i32, in the order the
reference's copy_* functions consume them (boltzmann_param.c:5908-6026).  The
blob is parameter data (the published parameters_BLstar_Vienna.txt set), not
source; libdafs_cuda embeds it and oracle/mccaskill_oracle.c reads it.
    python tools/extract_blstar.py [/root/reference/src/boltzmann_param.c [OUT]]
Layout (int32 each): magic 0x424C5331, then
  stack[7][7]  mismatchH[7][5][5]  mismatchI[7][5][5]  dangle5[8][5]  dangle3[8][5]
  int11[7][7][5][5]  int21[7][7][5][5][5]  int22[7][7][4][4][4][4]
  hairpin[31]  bulge[31]  internal[31]  ML{base, closing, intern, TerminalAU}
  ninio{m, max}  ntetra  then ntetra x {6 chars as ints, energy}"""
import os
import re
import struct
import sys

src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/src/boltzmann_param.c"
out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dafs_b200", "data", "blstar37.i32")
txt = open(src).read()
txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
CONST = {"INF": 10000000, "DEF": -50, "NST": 0}


def arr(name, n):
    m = re.search(r"static int %s\[\]\s*=\s*\{(.*?)\};" % name, txt, flags=re.S)
    vals = [CONST[t] if t in CONST else int(t) for t in re.findall(r"-?\d+|INF|DEF|NST", m.group(1))]
    assert len(vals) == n, (name, len(vals), n)
    return vals


data = [0x424C5331]
data += arr("stack37a", 49) + arr("mismatchH37a", 175) + arr("mismatchI37a", 175)
data += arr("dangle5_37a", 40) + arr("dangle3_37a", 40)
data += arr("int11_37a", 1225) + arr("int21_37a", 6125) + arr("int22_37a", 12544)
data += arr("hairpin37a", 31) + arr("bulge37a", 31) + arr("internal_loop37a", 31)
data += arr("MLparams_a", 4) + arr("ninio_a", 2)
tl = re.findall(r'\{"([ACGU]{6})",\s*(-?\d+)\}', txt)
data.append(len(tl))
for s, e in tl:
    data += [ord(c) for c in s] + [int(e)]
open(out, "wb").write(struct.pack("<%di" % len(data), *data))
print(out, len(data), "ints,", len(tl), "tetraloops")
