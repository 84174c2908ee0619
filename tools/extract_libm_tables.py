#!/usr/bin/env python
"""Print glibc's __exp2f_data.tab and __powf_log2_data.tab as found in this machine's libm.so.6 (located by their first entries),
in the form csrc/device/libm_exact.cuh holds them (VPT_EXP2F_TABLE_INIT / VPT_POWF_LOG2_TABLE_INIT). Development aid: the tables
are constants of glibc >= 2.28's float exp/pow; the exhaustive sweep in oracle/libm_check.cpp is what proves they are right."""
import struct
import sys

path = sys.argv[1] if len(sys.argv) > 1 else "/lib/x86_64-linux-gnu/libm.so.6"
b = open(path, "rb").read()
i = b.find(struct.pack("<QQ", 0x3FF0000000000000, 0x3FEFD9B0D3158574))
print("exp2f tab @", hex(i))
print(", ".join("0x%016xULL" % x for x in struct.unpack("<32Q", b[i:i + 256])))
first = struct.pack("<dd", float.fromhex("0x1.661ec79f8f3bep+0"), float.fromhex("-0x1.efec65b963019p-2"))
i = -1
while True:
    i = b.find(first, i + 1)
    if i < 0:
        break
    tail = struct.unpack("<5d", b[i + 256:i + 296])
    if tail[4] != 0.0 and abs(tail[4] - 1.4427) < 1e-3:  # the 5-term polynomial ends with ~1/ln2: __powf_log2_data (log2f's has 4 terms)
        print("powf log2 tab @", hex(i))
        print(", ".join(x.hex() for x in struct.unpack("<32d", b[i:i + 256])))
