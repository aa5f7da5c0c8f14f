"""Generates the constants of include/gz_detmath.h with mpmath (60 digits)."""
import mpmath as mp
mp.mp.dps = 60

def split(x):
    hi = float(x)
    lo = float(x - mp.mpf(hi))
    return hi, lo

print("// atan(k/8), k = 0..8, as hi + lo")
for k in range(9):
    hi, lo = split(mp.atan(mp.mpf(k) / 8))
    print("  {%s, %s}," % (hi.hex(), lo.hex()))
for name, v in (("pi", mp.pi), ("pio2", mp.pi / 2)):
    hi, lo = split(v)
    print(name, hi.hex(), lo.hex())
# Cody-Waite pi/2 in three pieces, the first two with 33 significant bits
p = mp.pi / 2
import math
def trunc_bits(x, bits):
    m, e = math.frexp(float(x))
    return math.ldexp(math.floor(m * 2 ** bits) / 2 ** bits, e)
p1 = trunc_bits(p, 33)
p2 = trunc_bits(p - mp.mpf(p1), 33)
p3 = float(p - mp.mpf(p1) - mp.mpf(p2))
print("P1", p1.hex(), "P2", p2.hex(), "P3", p3.hex())
print("two_over_pi", float(2 / mp.pi).hex())
