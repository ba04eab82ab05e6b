"""CPU: the two fixed-point formats behind the order-independent sums (melissa_b200/csrc/fx128.cuh), restated
operation by operation with exact Python integers.  Pins what the kernels rely on: conversion error <= 2^-52 of one
term, exact (hence order-independent) addition, the carry between the two words, the treatment of tiny negative
terms (the fraction must not round up to 1), and that four 20-bit limbs absorb 4096 additions without a carry."""
import math
import random
from fractions import Fraction

import numpy as np

MASK64 = (1 << 64) - 1


def fx_from_double(v):
    """fx_from_double: hi = floor(v) (DADD.RM against 1.5 * 2^52), lo = the fraction rounded toward zero to 52 bits,
    scaled to 2^-64.  Returns (lo, hi) as the two unsigned 64-bit words the device adds atomically."""
    hi = math.floor(v)
    frac = Fraction(v) - hi                                   # exact, in [0, 1)
    f = float(frac)                                           # __dadd_rz(v, -floor): round toward zero
    if Fraction(f) > frac:
        f = np.nextafter(f, 0.0)
    assert 0.0 <= f < 1.0
    mant = int(Fraction(f) * (1 << 52))                       # bits of (1 + f) below the leading one
    return (mant << 12) & MASK64, hi & MASK64


def fx_add(acc, term):
    lo = acc[0] + term[0]
    carry = lo >> 64
    return lo & MASK64, (acc[1] + term[1] + carry) & MASK64


def fx_value(acc):
    hi = acc[1] - (1 << 64) if acc[1] >> 63 else acc[1]
    return Fraction(hi) + Fraction(acc[0], 1 << 64)


def test_one_term_is_represented_to_two_to_the_minus_52():
    rng = random.Random(1)
    for _ in range(4000):
        v = rng.choice([1e-300, 1e-17, 1e-3, 1.0, 37.5, 1e6, 2.0 ** 49]) * rng.uniform(-1, 1)
        got = fx_value(fx_add((0, 0), fx_from_double(v)))
        assert abs(got - Fraction(v)) <= Fraction(1, 1 << 52), v
        assert got <= Fraction(v)                             # always rounded down: no term can push a sum up


def test_a_tiny_negative_term_does_not_become_zero_plus_one():
    # floor(-1e-20) = -1, fraction = 1 - 1e-20, which a round-to-nearest subtraction turns into 1.0 (the bound was off
    # by exactly +1 once because of it): the fraction is rounded toward zero instead
    for v in (-1e-20, -2.0 ** -60, -5e-324):
        got = fx_value(fx_add((0, 0), fx_from_double(v)))
        assert Fraction(-1, 1 << 51) <= got <= 0, (v, float(got))


def test_sums_do_not_depend_on_the_order_of_the_additions():
    rng = random.Random(2)
    terms = [fx_from_double(rng.gauss(0, 1) * 10 ** rng.uniform(-8, 4)) for _ in range(3000)]
    ref = (0, 0)
    for t in terms:
        ref = fx_add(ref, t)
    for seed in range(5):
        random.Random(seed).shuffle(terms)
        acc = (0, 0)
        for t in terms:
            acc = fx_add(acc, t)
        assert acc == ref
    # and the sum is the exact sum of the converted terms (carries included)
    assert fx_value(ref) == sum((fx_value(t) for t in terms), Fraction(0))


def limbs_from_double(v):
    """limb_add_shared: value * 2^44 floored, cut into l0, l1, l2 (20 bits each) and a signed top limb."""
    v16 = v * 16.0
    hi = math.floor(v16)
    rem = Fraction(v16) - hi
    r = float(rem)
    if Fraction(r) > rem:
        r = np.nextafter(r, 0.0)
    lo = math.floor(r * 2.0 ** 40)
    top = hi >> 20                                            # arithmetic shift: sign lives here
    return [lo & 0xfffff, (lo >> 20) & 0xfffff, hi & 0xfffff, top & 0xffffffff]


def limbs_value(acc):
    top = acc[3] - (1 << 32) if acc[3] >> 31 else acc[3]
    return Fraction(acc[0] + (acc[1] << 20) + (acc[2] << 40) + (top << 60), 1 << 44)


def test_four_limbs_take_4096_additions_without_a_carry_between_words():
    rng = random.Random(3)
    acc = [0, 0, 0, 0]
    exact = Fraction(0)
    for i in range(4096):
        v = rng.uniform(-1, 1) * 10 ** rng.uniform(-6, 9) if i % 7 else math.copysign(2.0 ** 35 - 1, rng.uniform(-1, 1))
        limbs = limbs_from_double(v)
        assert all(0 <= x < (1 << 20) for x in limbs[:3])
        assert abs(limbs_value(limbs) - Fraction(v)) <= Fraction(1, 1 << 44)
        for j in range(3):
            acc[j] += limbs[j]
            assert acc[j] < (1 << 32)                         # the 12 spare bits of the 32-bit word hold
        acc[3] = (acc[3] + limbs[3]) & 0xffffffff
        exact += limbs_value(limbs)
    assert limbs_value(acc) == exact
