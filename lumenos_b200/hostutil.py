"""Host-side parameter material that the Go host obtains from Lattigo, restated in Python so that the benchmark
and the examples can run without the oracle: NTT-friendly prime chains (rlwe.GenModuli as used by
fhe/bfv.go:160-185), the primitive 2N-th roots of each ring, and the plaintext-field twiddle table
core.PrimeField.RootsForward (core/field.go:138-197).  Pure integer arithmetic on Python ints; nothing here
touches ciphertext data."""
import math
import random

import numpy as np

T_DEFAULT = 144115188075593729  # 2^57 - 2^18 + 1 (cmd/server/main.go:22)

_SMALL = (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37)


def is_prime(n):
    if n < 2:
        return False
    for p in _SMALL:
        if n % p == 0:
            return n == p
    d, s = n - 1, 0
    while d % 2 == 0:
        d //= 2
        s += 1
    for a in _SMALL:
        x = pow(a, d, n)
        if x in (1, n - 1):
            continue
        for _ in range(s - 1):
            x = x * x % n
            if x == n - 1:
                break
        else:
            return False
    return True


def _rho(n):
    if n % 2 == 0:
        return 2
    rng = random.Random(n)
    while True:
        c = rng.randrange(1, n)
        f = lambda v: (v * v + c) % n  # noqa: E731
        x = y = 2
        d = 1
        while d == 1:
            x, y = f(x), f(f(y))
            d = math.gcd(abs(x - y), n)
        if d != n:
            return d


def prime_factors(n):
    out, stack = set(), [n]
    while stack:
        m = stack.pop()
        if m == 1:
            continue
        if is_prime(m):
            out.add(m)
            continue
        d = _rho(m)
        stack += [d, m // d]
    return sorted(out)


def primitive_root(q, start=3):
    """lattigo ring.PrimitiveRoot: first generator found counting up from `start`"""
    fs = prime_factors(q - 1)
    g = start
    while any(pow(g, (q - 1) // f, q) == 1 for f in fs):
        g += 1
    return g


class _PrimeGen:
    """lattigo ring.NTTFriendlyPrimesGenerator.NextAlternatingPrime"""

    def __init__(self, bits, nth):
        self.bits, self.nth = bits, nth
        self.next = (1 << bits) + 1
        self.prev = (1 << bits) + 1 - nth
        self.cn, self.cp = True, self.prev > 0

    def __call__(self):
        while True:
            if not (self.cn or self.cp):
                raise RuntimeError("prime generator exhausted")
            if self.cn:
                if self.next.bit_length() > self.bits + 1:
                    self.cn = False
                else:
                    c = self.next
                    self.next += self.nth
                    if is_prime(c):
                        return c
            if self.cp:
                if self.prev.bit_length() < self.bits or self.prev < self.nth:
                    self.cp = False
                else:
                    c = self.prev
                    self.prev -= self.nth
                    if is_prime(c):
                        return c


def generate_moduli(logN, k, logq=(58, 56), logp=(55, 55)):
    """Q = [logq[0]-bit, logq[1]-bit x (k-1)], P = logp (fhe/bfv.go:160-178), Lattigo's alternating search"""
    nth = 2 << logN
    gens = {}
    def take(bits):
        if bits not in gens:
            gens[bits] = _PrimeGen(bits, nth)
        return gens[bits]()
    q = [take(logq[0])] + [take(logq[1]) for _ in range(k - 1)]
    p = [take(b) for b in logp]
    return q, p


def psi_of(q, logN, primroot_start=3):
    return pow(primitive_root(q, primroot_start), (q - 1) // (2 << logN), q)


def num_q_primes_for_ntt(ntt_size, t=T_DEFAULT):
    """fhe/bfv.go:141-158"""
    return (ntt_size & -ntt_size).bit_length() - 1 + (0 if t.bit_length() > 45 else -2)


def field_roots_forward(n_field, t=T_DEFAULT, primroot_start=3):
    """core.NewPrimeField(t, n_field).RootsForward: RootsForward[bitrev(j)] = psi^j * 2^64 mod t (core/field.go:172-194)"""
    if n_field & (n_field - 1) or (t - 1) % (2 * n_field):
        raise ValueError("invalid field parameters")
    logn = n_field.bit_length() - 1
    psi = pow(primitive_root(t, primroot_start), (t - 1) // (2 * n_field), t)
    out = np.zeros(n_field, np.uint64)
    cur = (1 << 64) % t
    for j in range(n_field):
        out[int(format(j, "0%db" % logn)[::-1], 2) if logn else 0] = cur
        cur = cur * psi % t
    return out
