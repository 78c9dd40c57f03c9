#!/usr/bin/env python
"""Prints the constant tables of glibc's powf (sysdeps/ieee754/flt-32/e_powf.c: __powf_log2_data, __exp2f_data) as
found in THIS image's libm.so.6, as C initialisers for path-tracing_b200/csrc/pt_libm.cuh.

The reference's gamma resolve calls glm::pow -> std::pow(float, float) = glibc powf (main.cpp:147,180); the device
restates that algorithm (double-precision table + polynomial, one rounding) so that the 8-bit image is identical.
The tables are located by the first entry of each (values published in glibc's e_powf_log2_data.c / e_exp2f_data.c);
tests/test_host.py checks the restated algorithm against the host's powf bit for bit."""
import struct
import sys

data = open(sys.argv[1] if len(sys.argv) > 1 else "/lib/x86_64-linux-gnu/libm.so.6", "rb").read()
first = struct.pack("<dd", float.fromhex("0x1.661ec79f8f3bep+0"), float.fromhex("-0x1.efec65b963019p-2"))
hits = []
i = data.find(first)
while i >= 0:
    hits.append(i)
    i = data.find(first, i + 1)
# two log2 tables start with this entry: log2f's (4 polynomial coefficients) and powf's (5); take the one followed by 5
for h in hits:
    poly = struct.unpack("<5d", data[h + 256:h + 296])
    if abs(poly[4] - 1.4426950408889634) < 1e-9:
        base = h
tab = struct.unpack("<32d", data[base:base + 256])
poly = struct.unpack("<5d", data[base + 256:base + 296])
print("// __powf_log2_data: 16 x (invc, logc), then the degree-5 polynomial")
print("__device__ const double kPowLog2Tab[32] = {")
for k in range(16):
    print(f"    {tab[2 * k].hex()}, {tab[2 * k + 1].hex()},")
print("};")
print("__device__ const double kPowLog2Poly[5] = {" + ", ".join(p.hex() for p in poly) + "};")
e = data.find(struct.pack("<QQ", 0x3ff0000000000000, 0x3fefd9b0d3158574))
t = struct.unpack("<32Q", data[e:e + 256])
rest = struct.unpack("<4d", data[e + 256:e + 288])
print("// __exp2f_data: tab[32] (bits of 2^(i/32) minus i << 47), shift_scaled, poly[3]")
print("__device__ const unsigned long long kExp2fTab[32] = {")
for k in range(0, 32, 4):
    print("    " + ", ".join(f"0x{v:016x}ull" for v in t[k:k + 4]) + ",")
print("};")
print(f"// shift_scaled = {rest[0].hex()}, poly = " + ", ".join(v.hex() for v in rest[1:]))
