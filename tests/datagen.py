"""Seeded synthetic inputs shared by the tests (small versions of the BASELINE configs)."""
import numpy as np

_LEVELS = [b"DEBUG", b"INFO", b"WARN", b"ERROR", b"TRACE"]


def log_lines(nbytes, seed):
    """C2-style log text: monotone timestamps, 5 levels, 64 components, 256 templates."""
    rng = np.random.default_rng(seed)
    comps = [b"comp-%02d.%s" % (i, bytes(rng.choice(list(b"abcdefghij"), 6).tolist())) for i in range(64)]
    tmpls = [b" ".join(bytes(rng.choice(list(b"abcdefghijklmnopqrstuvwxyz"), int(rng.integers(3, 9))).tolist())
                       for _ in range(int(rng.integers(3, 8)))) for _ in range(256)]
    out = bytearray()
    t = 1_700_000_000_000
    while len(out) < nbytes:
        t += int(rng.integers(0, 50))
        out += b"2026-11-04T%02d:%02d:%02d.%03dZ %s %s msg=%s a=%d b=%d\n" % (
            (t // 3600000) % 24, (t // 60000) % 60, (t // 1000) % 60, t % 1000,
            _LEVELS[int(rng.integers(0, 5))], comps[int(rng.integers(0, 64))],
            tmpls[int(rng.integers(0, 256))], int(rng.integers(0, 100000)), int(rng.integers(0, 1000)))
    return np.frombuffer(bytes(out[:nbytes]), dtype=np.uint8).copy()


def zipf_text(nbytes, seed):
    """C1-style text: Zipf(1.2) words from a 4096-word vocabulary."""
    rng = np.random.default_rng(seed)
    vocab = [bytes(rng.choice(list(b"abcdefghijklmnopqrstuvwxyz"), int(rng.integers(2, 13))).tolist())
             for _ in range(4096)]
    out = bytearray()
    while len(out) < nbytes:
        idx = np.minimum(rng.zipf(1.2, 2048) - 1, 4095)
        for i in idx:
            out += vocab[int(i)]
            out += b"\n" if rng.random() < 0.08 else b" "
    return np.frombuffer(bytes(out[:nbytes]), dtype=np.uint8).copy()


def sorted_ints(n, seed, dtype=np.int32):
    """C3-style sorted column: cumsum of steps, 0 w.p. 0.7 else Geometric(0.05)."""
    rng = np.random.default_rng(seed)
    steps = np.where(rng.random(n) < 0.7, 0, rng.geometric(0.05, n))
    return np.cumsum(steps).astype(dtype)


def mixed_chunks(chunk_bytes, n_chunks, seed):
    """List of chunks cycling text / log / random / zeros / repeated line / ints."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n_chunks):
        k = i % 6
        if k == 0:
            c = zipf_text(chunk_bytes, seed + i)
        elif k == 1:
            c = log_lines(chunk_bytes, seed + i)
        elif k == 2:
            c = rng.integers(0, 256, chunk_bytes, dtype=np.uint8)
        elif k == 3:
            c = np.zeros(chunk_bytes, np.uint8)
        elif k == 4:
            line = b"the quick brown fox jumps over the dog\n"[:37]
            c = np.frombuffer((line * (chunk_bytes // 37 + 1))[:chunk_bytes], np.uint8).copy()
        else:
            c = sorted_ints((chunk_bytes + 3) // 4, seed + i).view(np.uint8)[:chunk_bytes].copy()
        out.append(c)
    return out
