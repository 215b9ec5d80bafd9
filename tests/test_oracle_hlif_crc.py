"""The arithmetic of csrc/crc32.cuh (per-chunk CRC-32 of the HLIF container) restated in Python and held
against zlib: slice-by-4 tables, the warp's slicing of a chunk (lane 0 the head, lanes 1..31 equal slices), and
the combination tree crc(A || B) = x^(8|B|) * crc(A) + crc(B) with one operator per chunk and one squaring per level."""
import random
import zlib

import pytest

POLY = 0xEDB88320
M32 = 0xFFFFFFFF

# csrc/crc32.cuh kX2n
X2N = [0x40000000, 0x20000000, 0x08000000, 0x00800000, 0x00008000, 0xedb88320, 0xb1e6b092, 0xa06a2517,
       0xed627dae, 0x88d14467, 0xd7bbfe6a, 0xec447f11, 0x8e7ea170, 0x6427800e, 0x4d47bae0, 0x09fe548f,
       0x83852d0f, 0x30362f1a, 0x7b5a9cc3, 0x31fec169, 0x9fec022a, 0x6c8dedc4, 0x15d6874d, 0x5fde7a4e,
       0xbad90e37, 0x2e4e5eef, 0x4eaba214, 0xa8a472c0, 0x429a969e, 0x148d302a, 0xc40ba6d0, 0xc4e22c3c]


def make_tables():
    t0 = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ (POLY if c & 1 else 0)
        t0.append(c)
    t = [t0]
    for k in range(1, 4):
        t.append([(t[k - 1][i] >> 8) ^ t0[t[k - 1][i] & 0xFF] for i in range(256)])
    return t


T = make_tables()


def crc_bytes(data, misalign=0):
    """finished CRC-32: bytes up to a 4-byte boundary (the slice starts `misalign` bytes past one), words, tail"""
    c, i, n = M32, 0, len(data)
    while i < n and (misalign + i) & 3:
        c = (c >> 8) ^ T[0][(c ^ data[i]) & 0xFF]
        i += 1
    while i + 4 <= n:
        c ^= int.from_bytes(data[i:i + 4], "little")
        c = T[3][c & 0xFF] ^ T[2][(c >> 8) & 0xFF] ^ T[1][(c >> 16) & 0xFF] ^ T[0][c >> 24]
        i += 4
    while i < n:
        c = (c >> 8) ^ T[0][(c ^ data[i]) & 0xFF]
        i += 1
    return c ^ M32


def multmodp(a, b):
    p = 0
    for i in range(31, -1, -1):
        if (a >> i) & 1:
            p ^= b
        b = (b >> 1) ^ (POLY if b & 1 else 0)
    return p


def x2nmodp(n, k):
    p = 0x80000000
    while n:
        if n & 1:
            p = multmodp(X2N[k & 31], p)
        n >>= 1
        k += 1
    return p


def warp_crc(data, misalign=0):
    n = len(data)
    L = (n // 32) & ~3
    head = n - 31 * L
    c = []
    for lane in range(32):
        a = 0 if lane == 0 else head + (lane - 1) * L
        b = head if lane == 0 else a + L
        c.append(crc_bytes(data[a:b], misalign + a))
    op = x2nmodp(L, 3)
    s = 1
    while s < 32:
        c = [multmodp(op, c[i]) ^ (c[i + s] if i + s < 32 else c[i]) for i in range(32)]  # every lane, as the shuffle does
        op = multmodp(op, op)
        s *= 2
    return c[0]


def test_table_of_squares():
    assert X2N[0] == 1 << 30
    for i in range(1, 32):
        assert X2N[i] == multmodp(X2N[i - 1], X2N[i - 1])


@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 5, 31, 32, 33, 127, 128, 129, 131, 1000, 4095, 4096, 4100, 16384, 65535, 65536])
def test_warp_crc_equals_zlib(n):
    rng = random.Random(n)
    data = bytes(rng.getrandbits(8) for _ in range(n))
    for misalign in (0, 1, 3):
        assert crc_bytes(data, misalign) == zlib.crc32(data)
        assert warp_crc(data, misalign) == zlib.crc32(data)
