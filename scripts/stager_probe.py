"""Times the staging pipeline's directions separately (b200compStagerCompress / Decompress / RoundTrip) for several
slice counts on C2-like data: where the e2e milliseconds go."""
import importlib, sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
S = importlib.import_module("hipcomp-core_b200.staging")
SY = importlib.import_module("hipcomp-core_b200.synth")
nbytes, chunk = 1 << 30, 65536
raw = SY.log_lines(nbytes, 0xB202, "cuda")
n = nbytes // chunk
h_raw, h_back = S.PinnedArray(nbytes), S.PinnedArray(nbytes)
h_blob, h_sizes, h_stat = S.PinnedArray(n * 65808), S.PinnedArray(n, np.uint64), S.PinnedArray(n, np.int32)
torch.from_numpy(h_raw.array).copy_(raw)
for slices in (4, 8, 16, 32):
    st = S.Stager("lz4", chunk, nbytes, slices=slices)
    def t(fn, reps=4):
        fn(); t0 = time.perf_counter()
        for _ in range(reps): fn()
        return (time.perf_counter() - t0) / reps * 1e3
    tc = t(lambda: st.compress(h_raw, h_blob, h_sizes))
    td = t(lambda: st.decompress(h_blob, h_sizes, h_back, h_stat))
    tr = t(lambda: st.roundtrip(h_raw, h_blob, h_sizes, h_back, h_stat))
    print(f"slices {slices:2d}: compress {tc:6.2f} ms  decompress {td:6.2f} ms  roundtrip {tr:6.2f} ms", flush=True)
    st.close()
