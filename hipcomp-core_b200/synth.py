"""Seeded synthetic workloads of the BASELINE configs, generated with torch ops so
that multi-GiB inputs are built on the GPU in seconds (same code runs on CPU).

C1 zipf_text   : Zipf(1.2) words from a 4096-word vocabulary; 10 % of 64 KiB chunks
                 uniform random, 5 % one repeated 37-byte line.
C2/C4 log_lines: "<ISO timestamp, monotone +U(0,50) ms> <LEVEL in 5> <component in 64>
                 msg=<template in 256> a=<5 digits> b=<3 digits>\\n"
C3 sorted_ints : cumsum(step), step = 0 w.p. 0.7 else Geometric(0.05), as int32.
"""
import numpy as np
import torch

_BLOCK = 64 << 20


def _gather_pieces(src, src_off, lens):
    """Concatenate src[src_off[i] : src_off[i]+lens[i]] for all i (flattened, in order)."""
    lens = lens.reshape(-1)
    src_off = src_off.reshape(-1)
    out_start = torch.cumsum(lens, 0) - lens
    total = int(lens.sum().item())
    idx = torch.repeat_interleave(src_off - out_start, lens, output_size=total)
    idx += torch.arange(total, device=src.device)
    return src[idx]


def _pool(pieces, device):
    buf = np.frombuffer(b"".join(pieces), dtype=np.uint8).copy()
    lens = np.array([len(p) for p in pieces], dtype=np.int64)
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]])
    return (torch.from_numpy(buf).to(device), torch.from_numpy(offs).to(device), torch.from_numpy(lens).to(device))


def _digits(val, width):
    """val: int64 tensor [N] -> uint8 [N, width] of zero-padded decimal digits."""
    cols = []
    for k in range(width - 1, -1, -1):
        cols.append(((val // (10 ** k)) % 10 + 48).to(torch.uint8))
    return torch.stack(cols, 1)


def log_lines(nbytes, seed, device="cuda"):
    rs = np.random.default_rng(seed)
    letters = list(b"abcdefghijklmnopqrstuvwxyz")
    levels = [b"DEBUG", b"INFO", b"WARN", b"ERROR", b"TRACE"]
    comps = [b"comp-%02d." % i + bytes(rs.choice(letters, 6).tolist()) for i in range(64)]
    tmpls = [b" ".join(bytes(rs.choice(letters, int(rs.integers(3, 9))).tolist()) for _ in range(int(rs.integers(3, 8))))
             for _ in range(256)]
    consts = [b" ", b" msg=", b" a=", b" b=", b"\n"]
    pool, poff, plen = _pool(levels + comps + tmpls + consts, device)
    L0, C0, T0, K0 = 0, 5, 5 + 64, 5 + 64 + 256
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = torch.empty(nbytes, dtype=torch.uint8, device=device)
    done, t0 = 0, 1_700_000_000_000
    while done < nbytes:
        want = min(_BLOCK, nbytes - done)
        n = want // 56 + 16  # shortest possible line is > 56 bytes
        dt = torch.randint(0, 50, (n,), generator=g, device=device)
        t = t0 + torch.cumsum(dt, 0)
        t0 = int(t[-1].item())
        ts = torch.empty((n, 24), dtype=torch.uint8, device=device)
        ts[:, 0:11] = torch.tensor(list(b"2026-11-04T"), dtype=torch.uint8, device=device)
        ts[:, 11:13] = _digits((t // 3600000) % 24, 2)
        ts[:, 13] = 58
        ts[:, 14:16] = _digits((t // 60000) % 60, 2)
        ts[:, 16] = 58
        ts[:, 17:19] = _digits((t // 1000) % 60, 2)
        ts[:, 19] = 46
        ts[:, 20:23] = _digits(t % 1000, 3)
        ts[:, 23] = 90
        a = _digits(torch.randint(0, 100000, (n,), generator=g, device=device), 5)
        b = _digits(torch.randint(0, 1000, (n,), generator=g, device=device), 3)
        dyn = torch.cat([ts, a, b], 1).reshape(-1)  # 32 bytes per line
        src = torch.cat([pool, dyn])
        base = pool.numel() + torch.arange(n, device=device) * 32
        lvl = torch.randint(0, 5, (n,), generator=g, device=device) + L0
        cmp_ = torch.randint(0, 64, (n,), generator=g, device=device) + C0
        tpl = torch.randint(0, 256, (n,), generator=g, device=device) + T0

        def const(k):
            return torch.full((n,), K0 + k, device=device)

        ids = torch.stack([lvl, cmp_, tpl, const(0), const(1), const(2), const(3), const(4)], 1)
        o_, l_ = poff[ids], plen[ids]
        full = lambda v: torch.full((n,), v, device=device, dtype=torch.int64)
        # ts ' ' level ' ' comp ' msg=' tmpl ' a=' a ' b=' b '\n'
        offs = torch.stack([base, o_[:, 3], o_[:, 0], o_[:, 3], o_[:, 1], o_[:, 4], o_[:, 2], o_[:, 5], base + 24,
                            o_[:, 6], base + 29, o_[:, 7]], 1)
        lens = torch.stack([full(24), l_[:, 3], l_[:, 0], l_[:, 3], l_[:, 1], l_[:, 4], l_[:, 2], l_[:, 5], full(5),
                            l_[:, 6], full(3), l_[:, 7]], 1)
        blk = _gather_pieces(src, offs, lens)
        out[done : done + want] = blk[:want]
        done += want
    return out


def zipf_text(nbytes, seed, device="cuda", chunk=65536, mixed=True):
    rs = np.random.default_rng(seed)
    letters = list(b"abcdefghijklmnopqrstuvwxyz")
    vocab = [bytes(rs.choice(letters, int(rs.integers(2, 13))).tolist()) for _ in range(4096)]
    pool, poff, plen = _pool(vocab + [b" ", b"\n"], device)
    probs = torch.tensor(1.0 / np.arange(1, 4097) ** 1.2, dtype=torch.float32, device=device)
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = torch.empty(nbytes, dtype=torch.uint8, device=device)
    done = 0
    while done < nbytes:
        want = min(_BLOCK, nbytes - done)
        n = want // 3 + 16
        w = torch.multinomial(probs, n, replacement=True, generator=g)
        sep = 4096 + (torch.rand(n, generator=g, device=device) < 0.08).long()
        ids = torch.stack([w, sep], 1)
        blk = _gather_pieces(pool, poff[ids], plen[ids])
        out[done : done + want] = blk[:want]
        done += want
    if mixed:
        nchunks = nbytes // chunk
        kind = torch.rand(nchunks, generator=g, device=device)
        view = out[: nchunks * chunk].view(nchunks, chunk)
        rnd = torch.nonzero(kind < 0.10).reshape(-1)
        if rnd.numel():
            view[rnd] = torch.randint(0, 256, (rnd.numel(), chunk), generator=g, device=device, dtype=torch.uint8)
        rep = torch.nonzero((kind >= 0.10) & (kind < 0.15)).reshape(-1)
        if rep.numel():
            line = torch.tensor(list(b"the quick brown fox jumps over a dog\n"), dtype=torch.uint8, device=device)
            view[rep] = line.repeat(chunk // 37 + 1)[:chunk]
    return out


def sorted_ints(nelems, seed, device="cuda", dtype=torch.int32):
    """dtype: torch.int8 / int16 / int32 / int64 - the same column, wrapped to the element width."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = torch.empty(nelems, dtype=dtype, device=device)
    done, carry = 0, 0
    blk = _BLOCK // 4
    while done < nelems:
        n = min(blk, nelems - done)
        geo = torch.empty(n, device=device).geometric_(0.05, generator=g).long()
        zero = torch.rand(n, generator=g, device=device) < 0.7
        steps = torch.where(zero, torch.zeros_like(geo), geo)
        vals = carry + torch.cumsum(steps, 0)
        carry = int(vals[-1].item())
        out[done : done + n] = vals.to(dtype)  # wraps modulo 2^width (harmless for the codec)
        done += n
    return out
