#!/usr/bin/env python
"""HLIF (manager) throughput: one buffer in, one self-describing container out, and back.
usage: bench_hlif.py [size_mib] [--checksums]   - CUDA events on the manager's stream, 3 warm-up + 5 timed calls each way."""
import importlib, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
M = importlib.import_module("hipcomp-core_b200.managers")
L = importlib.import_module("hipcomp-core_b200._lib")
S = importlib.import_module("hipcomp-core_b200.synth")
mib = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
n = mib << 20
dev = torch.device("cuda", 0)
cases = [("lz4", lambda: M.LZ4Manager(65536), S.log_lines(n, 0xB202, dev)),
         ("snappy", lambda: M.SnappyManager(65536), S.zipf_text(n, 0xB201, dev, chunk=65536)),
         ("cascaded(1,1,1)", lambda: M.CascadedManager(L.CascadedOpts(4096, L.TYPE_INT, 1, 1, 1)),
          S.sorted_ints(n // 4, 0xB203, dev).view(torch.uint8)),
         ("cascaded(default 2,1,1)", lambda: M.CascadedManager(), S.sorted_ints(n // 4, 0xB203, dev).view(torch.uint8))]
if "--checksums" in sys.argv:  # the same managers with per-chunk CRC-32 filled on compress and verified on decompress
    def with_crc(make):
        def f():
            m = make()
            m.set_checksum_mode(1)
            return m
        return f
    cases = [(name + " + crc32", with_crc(make), raw) for name, make, raw in cases[:3]]
ev = lambda: torch.cuda.Event(enable_timing=True)
for name, make, raw in cases:
    mgr = make()
    cc = mgr.configure_compression(n)
    comp = torch.empty(cc.max_compressed_buffer_size, dtype=torch.uint8, device=dev)
    back = torch.empty_like(raw)
    tc = td = 0.0
    for it in range(8):
        a, b, c = ev(), ev(), ev()
        a.record()
        mgr.compress(raw, comp, cc)
        b.record()
        dc = mgr.configure_decompression(comp)
        c0 = ev(); c0.record()
        mgr.decompress(back, comp, dc)
        c.record()
        torch.cuda.synchronize()
        if it >= 3:
            tc += a.elapsed_time(b); td += c0.elapsed_time(c)
    size = mgr.get_compressed_output_size(comp)
    assert torch.equal(back, raw), name
    print(json.dumps({"manager": name, "MiB": mib, "compress_GBps": round(n * 5 / tc / 1e6, 1),
                      "decompress_GBps": round(n * 5 / td / 1e6, 1), "ratio": round(n / size, 3)}), flush=True)
    mgr.close()
