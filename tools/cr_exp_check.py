"""cr_exp (csrc/cuda/cr_exp.h) and the host libm's exp against 80-digit arithmetic: how often is each NOT the
correctly rounded double? Usage: /tmp/cr_exp_check 200000 | python tools/cr_exp_check.py"""
import struct
import sys
from decimal import Decimal, getcontext
getcontext().prec = 80
n = bad_cr = bad_libm = differ = 0
for line in sys.stdin:
    xa, ya, yb = (int(v, 16) for v in line.split())
    x = struct.unpack("<d", struct.pack("<Q", xa))[0]
    want = float(Decimal(x).exp())  # Decimal -> float conversion rounds to nearest even
    w = struct.unpack("<Q", struct.pack("<d", want))[0]
    n += 1
    bad_cr += ya != w
    bad_libm += yb != w
    differ += ya != yb
print(f"{n} arguments: cr_exp not correctly rounded {bad_cr}, libm exp not correctly rounded {bad_libm}, cr_exp != libm {differ}")
