"""bin2inc.py IN OUT -- little-endian int32 blob -> comma-separated C initialiser list."""
import struct
import sys

d = open(sys.argv[1], "rb").read()
v = struct.unpack("<%di" % (len(d) // 4), d)
with open(sys.argv[2], "w") as f:
    for k in range(0, len(v), 16):
        f.write(",".join(str(x) for x in v[k:k + 16]) + ",\n")
