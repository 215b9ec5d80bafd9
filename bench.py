#!/usr/bin/env python
"""bench.py — headline benchmark of the batched codec hot path (see DESIGN.md "Measurement").

Default run (`python bench.py [--gpus N]`): the three codecs of the metric, each on its BASELINE.json config,
  C2  LZ4      64 KiB chunks of 1 GiB synthetic log text per GPU      -> the line's `value`
  C1  Snappy   64 KiB chunks of 256 MiB mixed-entropy text            -> `codecs.snappy`
  C3  Cascaded RLE=1 delta=1 bitpack on 4 GiB of sorted int32         -> `codecs.cascaded`
One step = one CompressAsync over the whole batch followed by one DecompressAsync of its output.
metric = uncompressed GB (1e9 B) per second through that round trip, whole job (all ranks).  Chunks are
independent: every rank owns its own buffer ("scaling": "weak"), there is no data-path collective.
`e2e` is the same round trip from pinned host buffers through the library's staging pipeline
(b200compStagerRoundTrip), all PCIe traffic inside the timed region.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference|refgpu] [--codec lz4|snappy|cascaded]
  python bench.py --workload c4 [--gpus N]     C4: LZ4 decompress-only, 4 Mi x 4 KiB chunks, strong scaling
  python bench.py --workload c5 [--gpus N]     C5: input-class sweep (zeros/random/text/ints) x all codecs, one line each
"""
import argparse
import ctypes as C
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

GB = 1e9


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "refgpu"],
                    help="b200: this library; reference: the CPU codecs on the host cores; refgpu: the reference's "
                         "own GPU build (oracle/_ref/libhipcomp_ref.so) through the same ABI, for context")
    ap.add_argument("--codec", default=None, choices=["lz4", "snappy", "cascaded"],
                    help="one codec only (default: LZ4 as the line's value, Snappy and Cascaded as codecs.*)")
    ap.add_argument("--workload", default="default", choices=["default", "c4", "c5"])
    ap.add_argument("--c5-gib", type=int, default=16, help="C5: GiB per input class (whole job), cut into waves of <= 4 GiB per GPU")
    ap.add_argument("--size-mib", type=int, default=None, help="uncompressed MiB per GPU (default: the BASELINE config)")
    ap.add_argument("--chunk", type=int, default=None)
    ap.add_argument("--cpu-sample-mib", type=int, default=None, help="bound the CPU arm's buffer (default: the full config)")
    ap.add_argument("--cpu-threads", type=int, default=None)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--data", default="default", choices=["default", "zeros", "random", "text", "logs", "ints"],
                    help="input class (C5 sweep); default = the codec's BASELINE config generator")
    ap.add_argument("--casc-opts", default="1,1,1", help="cascaded num_RLEs,num_deltas,use_bp (C3 = 1,1,1; library default 2,1,1)")
    ap.add_argument("--casc-type", default="int", choices=["char", "short", "int", "longlong"])
    ap.add_argument("--e2e-slices", type=int, default=16, help="slices the staging pipeline cuts the buffer into")
    ap.add_argument("--e2e-mib", type=int, default=1024, help="upper bound on the e2e buffer per GPU (pinned host memory)")
    return ap.parse_args()


# workload table: BASELINE.json configs C2 (lz4, the metric's config), C1 (snappy), C3 (cascaded)
WORKLOADS = {
    "lz4": dict(name="C2: LZ4 compress+decompress, 64 KiB chunks of 1 GiB synthetic log text", size_mib=1024,
                chunk=65536, seed=0xB202, gen="log_lines"),
    "snappy": dict(name="C1: Snappy compress+decompress, 64 KiB chunks of 256 MiB mixed-entropy text", size_mib=256,
                   chunk=65536, seed=0xB201, gen="zipf_text"),
    "cascaded": dict(name="C3: Cascaded RLE=1 delta=1 bitpack, 4 GiB int32 sorted column, 64 KiB partitions",
                     size_mib=4096, chunk=65536, seed=0xB203, gen="sorted_ints"),
}
CASC_TYPES = {"char": (0, 1), "short": (2, 2), "int": (4, 4), "longlong": (6, 8)}  # hipcompType_t id, element bytes


def workload_for(args, codec):
    """The BASELINE config of `codec` with the command line's overrides; the name says what was overridden."""
    wl = dict(WORKLOADS[codec])
    base = dict(wl)
    if args.size_mib:
        wl["size_mib"] = args.size_mib
    if args.chunk:
        wl["chunk"] = args.chunk
    notes = []
    if wl["size_mib"] != base["size_mib"]:
        notes.append(f"{wl['size_mib']} MiB per GPU")
    if wl["chunk"] != base["chunk"]:
        notes.append(f"{wl['chunk']} B chunks")
    if args.data != "default":
        notes.append(f"input class {args.data}")
    if codec == "cascaded" and args.casc_opts != "1,1,1":
        notes.append(f"opts {args.casc_opts}")
    if codec == "cascaded" and args.casc_type != "int":
        notes.append(f"type {args.casc_type}")
    if notes:
        wl["name"] = f"{codec} (derived from {base['name'].split(':')[0]}): " + ", ".join(notes)
    return wl


def make_config(wl, world, nbytes, nchunks):
    return {"workload": wl["name"], "bytes_per_gpu": nbytes, "chunk_bytes": wl["chunk"], "chunks_per_gpu": nchunks,
            "sharding": f"{world} rank(s), independent chunk ranges, no collective",
            "l2": "inputs (>= 256 MiB per step) larger than the 126 MB L2; no explicit flush", "seed": wl["seed"]}


# ----------------------------------------------------------------------------- CPU codecs
class CpuCodec:
    """The CPU implementation timed as the reference arm / cpu_baseline.

    lz4     : liblz4.so.1 LZ4_compress_default / LZ4_decompress_safe (the CPU LZ4 block codec the north_star
              names; the reference ships none of its own), driven by oracle/cpu_bench.c (pthreads, clock_gettime)
    cascaded: oracle/cascaded_oracle.c (CPU restatement; no other CPU implementation exists), same C harness
    snappy  : the snappy bundled with pyarrow (stand-in for share/snappy's python-snappy dependency; hidden
              symbols, so it is driven from Python threads - pyarrow releases the GIL inside the codec)
    """

    def __init__(self, codec, chunk, casc=None, threads=None):
        self.codec, self.chunk = codec, chunk
        self.threads = threads or os.cpu_count() or 1
        self.casc = casc  # (type id, rle, delta, bp, width)
        if codec == "snappy":
            import pyarrow as pa

            self.pa = pa.Codec("snappy")
            self.kind, self.desc = "reference", "stock snappy (pyarrow bundled), Python threads"
        else:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import oracle_lib

            self.lib = oracle_lib.lib()
            self.lib.cpu_bench_run.restype = C.c_int
            self.lib.cpu_bench_run.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int,
                                               C.POINTER(C.c_int), C.POINTER(C.c_double)]
            if codec == "lz4":
                self.kind, self.desc = "reference", "liblz4.so.1 LZ4_compress_default+LZ4_decompress_safe, C pthreads harness"
            else:
                self.kind, self.desc = "port", "oracle/cascaded_oracle.c compress+decompress, C pthreads harness"

    def roundtrip_seconds(self, raw, threads=None, passes=1):
        """raw: np.uint8 array.  Returns (t_compress, t_decompress, compressed_bytes of one pass)."""
        T = threads or self.threads
        if self.codec != "snappy":
            out = (C.c_double * 3)()
            casc = (C.c_int * 5)(*(self.casc or (4, 1, 1, 1, 4)))
            rc = self.lib.cpu_bench_run(0 if self.codec == "lz4" else 2, raw.ctypes.data, raw.size, self.chunk, T, passes,
                                        casc, out)
            if rc != 0:
                raise RuntimeError(f"cpu_bench_run -> {rc}")
            return out[0], out[1], int(out[2])
        n, ch = (len(raw) + self.chunk - 1) // self.chunk, self.chunk
        comp = [None] * n
        mv = memoryview(raw)

        def run(fn):
            errs = []

            def work(k):
                try:
                    for i in range(k, n, T):
                        fn(i)
                except Exception as e:  # pragma: no cover
                    errs.append(e)

            ts = [threading.Thread(target=work, args=(k,)) for k in range(min(T, n))]
            t0 = time.perf_counter()
            for _ in range(passes):
                for t in ts:
                    t.start()
                for t in ts:
                    t.join()
                ts = [threading.Thread(target=work, args=(k,)) for k in range(min(T, n))]
            dt = time.perf_counter() - t0
            if errs:
                raise errs[0]
            return dt

        def enc(i):
            comp[i] = self.pa.compress(mv[i * ch:(i + 1) * ch])

        def dec(i):
            self.pa.decompress(comp[i], decompressed_size=min(ch, len(raw) - i * ch))

        tc = run(enc)
        td = run(dec)
        return tc, td, sum(c.size for c in comp)

    def baseline(self, raw, passes=1, steps=1, with_single=True):
        """cpu_baseline record: all threads over `raw` (x passes) per step, plus a 1-thread row on a bounded sample."""
        tcs, tds, cb = [], [], 0
        for _ in range(steps):
            tc, td, cb = self.roundtrip_seconds(raw, passes=passes)
            tcs.append(tc)
            tds.append(td)
        tot = len(raw) * passes * steps
        rec = {"value": tot / (sum(tcs) + sum(tds)) / GB, "unit": "GB/s", "cores": self.threads, "kind": self.kind,
               "sample": f"{len(raw) >> 20} MiB of the same workload x {passes} pass(es) per step, {self.desc}, {self.threads} threads",
               "compress_GBps": tot / sum(tcs) / GB, "decompress_GBps": tot / sum(tds) / GB,
               "compression_ratio": len(raw) / max(cb, 1), "seconds_per_step": (sum(tcs) + sum(tds)) / steps}
        if with_single:
            small = raw[: min(len(raw), 64 << 20)]
            tc, td, _ = self.roundtrip_seconds(small, threads=1)
            rec["single_thread"] = {"value": len(small) / (tc + td) / GB, "compress_GBps": len(small) / tc / GB,
                                    "decompress_GBps": len(small) / td / GB, "sample": f"first {len(small) >> 20} MiB, 1 thread"}
        return rec


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    # The sampler is started BEFORE the warm-up steps (nvidia-smi needs up to a second to produce its first line
    # on an 8-GPU box, longer than a short timed region) and stop() keeps the samples whose timestamp falls
    # inside the timed window [mark_begin, mark_end]; if the window is shorter than the sampling period it
    # falls back to the samples taken under the same load during warm-up and says so in "window".
    FIELDS = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc = gpu_index, None
        self.t0 = self.t1 = None

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    @staticmethod
    def _stamp(text):
        import datetime
        try:
            return datetime.datetime.strptime(text.strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()
        except ValueError:
            return None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out = self.proc.communicate(timeout=5)[0]
        except Exception:
            self.proc.kill()
            out = ""
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = []
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                rows.append((self._stamp(f[0]), float(f[1]), float(f[2]),
                             {nme for nme, v in zip(names, f[4:8]) if v.lower().startswith("active")}))
            except ValueError:
                continue
        window = "timed region"
        inside = [r for r in rows if r[0] is not None and self.t0 is not None and self.t1 is not None
                  and self.t0 - 0.05 <= r[0] <= self.t1 + 0.05]
        if not inside:
            inside, window = rows, "warm-up + timed region (timed region shorter than the sampling period)"
        sm = [r[1] for r in inside]
        mx = [r[2] for r in inside]
        reasons = set().union(*[r[3] for r in inside]) if inside else set()
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "window": window}


# ----------------------------------------------------------------------------- GPU arm
def load_traffic(codec, direction, nbytes):
    """DRAM bytes per launch of the dominant kernel, from the committed `ncu --set full` capture
    (profiles/traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of one launch and the uncompressed
    bytes that launch processed), scaled to this run's bytes per launch.  None when no capture is recorded."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[codec][direction]
        return {"traffic": t["dram_bytes"] / t["uncompressed_bytes"] * nbytes / GB, "traffic_unit": "GB per launch",
                "traffic_source": t["source"]}
    except Exception:
        return {"traffic": None}


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def dist_setup(n_gpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local) if torch.cuda.is_available() else None)
    return world, rank, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist

        dist.barrier()


def max_over_ranks(value, world, device):
    if world == 1:
        return value
    import torch.distributed as dist

    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, world, device):
    if world == 1:
        return value
    import torch.distributed as dist

    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def make_data(args, wl, rank, device):
    S = importlib.import_module("hipcomp-core_b200.synth")
    nbytes = wl["size_mib"] << 20
    seed = wl["seed"] + 1000 * rank
    data = getattr(args, "data", "default")
    if data == "zeros":
        return torch.zeros(nbytes, dtype=torch.uint8, device=device)
    if data == "random":
        g = torch.Generator(device=device)
        g.manual_seed(seed)
        return torch.randint(0, 256, (nbytes,), dtype=torch.uint8, device=device, generator=g)
    gen = {"text": "zipf_text", "logs": "log_lines", "ints": "sorted_ints"}.get(data, wl["gen"])
    wl = dict(wl, gen=gen)
    if wl["gen"] == "log_lines":
        return S.log_lines(nbytes, seed, device)
    if wl["gen"] == "zipf_text":
        return S.zipf_text(nbytes, seed, device, chunk=wl["chunk"])
    width = CASC_TYPES[getattr(args, "casc_type", "int")][1]
    tdt = {1: torch.int8, 2: torch.int16, 4: torch.int32, 8: torch.int64}[width]
    return S.sorted_ints(nbytes // width, seed, device, dtype=tdt).view(torch.uint8)


class Plan:
    """Pre-allocated buffers + argument lists so the timed loop is just the two C-ABI calls."""

    def __init__(self, codec, raw, chunk, casc_opts=(1, 1, 1), casc_type=4):
        B = importlib.import_module("hipcomp-core_b200.batched")
        L = importlib.import_module("hipcomp-core_b200._lib")
        self.B, self.L, self.codec, self.chunk = B, L, codec, chunk
        dev = raw.device
        self.n = raw.numel() // chunk
        self.raw = raw
        self.opts = L.CascadedOpts(4096, casc_type, *casc_opts) if codec == "cascaded" else B.default_opts(codec)
        off = np.arange(self.n, dtype=np.int64) * chunk
        self.src = B.DeviceBatch(raw, off, np.full(self.n, chunk, np.int64))
        self.stride = (B.max_compressed_chunk_bytes(codec, chunk, self.opts) + 15) // 16 * 16
        self.comp = B.DeviceBatch.empty(self.n, self.stride, dev)
        self.comp_sizes = torch.zeros(self.n, dtype=torch.int64, device=dev)
        self.temp_bytes = B.compress_temp_bytes(codec, self.n, chunk, self.opts)
        self.temp = torch.empty(max(self.temp_bytes, 16), dtype=torch.uint8, device=dev)
        self.dtemp_bytes = B.decompress_temp_bytes(codec, self.n, chunk)
        self.dtemp = torch.empty(max(self.dtemp_bytes, 16), dtype=torch.uint8, device=dev)
        self.back = torch.empty_like(raw)
        self.dst = B.DeviceBatch(self.back, off, np.full(self.n, chunk, np.int64))
        self.actual = torch.zeros(self.n, dtype=torch.int64, device=dev)
        self.status = torch.zeros(self.n, dtype=torch.int32, device=dev)
        name = {"lz4": "LZ4", "snappy": "Snappy", "cascaded": "Cascaded"}[codec]
        lib = L.lib()
        self.f_comp = getattr(lib, f"hipcompBatched{name}CompressAsync")
        self.f_dec = getattr(lib, f"hipcompBatched{name}DecompressAsync")
        if hasattr(lib, "b200compGatherChunksAsync"):
            self.f_gather = lib.b200compGatherChunksAsync
            self.f_gather.restype = C.c_int
            self.f_gather.argtypes = [C.c_void_p] * 4 + [C.c_size_t, C.c_void_p]

    def compress(self):
        """One CompressAsync over the whole batch."""
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        st = self.f_comp(self.src.ptrs.data_ptr(), self.src.sizes.data_ptr(), self.chunk, self.n, self.temp.data_ptr(),
                         self.temp_bytes, self.comp.ptrs.data_ptr(), self.comp_sizes.data_ptr(), self.opts, s)
        assert st == 0, f"CompressAsync -> {st}"

    def decompress(self):
        """One DecompressAsync of the compress output."""
        s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        st = self.f_dec(self.comp.ptrs.data_ptr(), self.comp_sizes.data_ptr(), self.dst.sizes.data_ptr(),
                        self.actual.data_ptr(), self.n, self.dtemp.data_ptr(), self.dtemp_bytes, self.dst.ptrs.data_ptr(),
                        self.status.data_ptr(), s)
        assert st == 0, f"DecompressAsync -> {st}"

    def verify(self):
        torch.cuda.synchronize()
        assert int(self.status.abs().sum().item()) == 0, "a chunk failed to decompress"
        assert bool((self.actual == self.chunk).all().item())
        assert torch.equal(self.back, self.raw), "round trip is not bit-exact"




def copy_ceiling(nbytes, device):
    """Plain pinned-memory copies of the same size on two streams: what PCIe gives this rank (H2D, D2H, both at once)."""
    n = min(nbytes, 1 << 30)
    h_a = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    h_b = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    d_a = torch.empty(n, dtype=torch.uint8, device=device)
    d_b = torch.empty(n, dtype=torch.uint8, device=device)
    s1, s2 = torch.cuda.Stream(device), torch.cuda.Stream(device)

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps

    def h2d():
        with torch.cuda.stream(s1):
            d_a.copy_(h_a, non_blocking=True)

    def d2h():
        with torch.cuda.stream(s2):
            h_b.copy_(d_b, non_blocking=True)

    def both():
        h2d()
        d2h()

    t1, t2, t3 = timed(h2d), timed(d2h), timed(both)
    return {"h2d_GBps": n / t1 / GB, "d2h_GBps": n / t2 / GB, "duplex_GBps_each_way": n / t3 / GB, "bytes": n}


def run_e2e(codec, plan, raw, args, world, device, comp_bytes):
    """Host-resident input -> compressed blob in host memory -> host-resident output, through the library's staging
    pipeline (b200compStagerRoundTrip, csrc/stager.cu): pinned raw -H2D-> CompressAsync -> chunks packed into the
    pinned blob by the device + sizes D2H -> blob slice H2D -> DecompressAsync -> D2H raw + statuses.  Every byte
    makes the whole trip inside the timed region; slices overlap on three streams."""
    S = importlib.import_module("hipcomp-core_b200.staging")
    nbytes = min(raw.numel(), args.e2e_mib << 20) // plan.chunk * plan.chunk
    n = nbytes // plan.chunk
    st = S.Stager(codec, plan.chunk, nbytes, opts=plan.opts if codec == "cascaded" else None, slices=args.e2e_slices)
    h_raw, h_back = S.PinnedArray(nbytes), S.PinnedArray(nbytes)
    h_blob, h_sizes, h_status = S.PinnedArray(n * plan.stride), S.PinnedArray(n, np.uint64), S.PinnedArray(n, np.int32)
    torch.from_numpy(h_raw.array).copy_(raw[:nbytes])
    K = max(2, min(args.steps, 5))
    total = st.roundtrip(h_raw, h_blob, h_sizes, h_back, h_status)  # warm-up + check
    assert (h_status.array == 0).all() and np.array_equal(h_back.array, h_raw.array), "e2e round trip failed"
    h_back.array[:] = 0
    barrier(world)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(K):
        total = st.roundtrip(h_raw, h_blob, h_sizes, h_back, h_status)  # blocks until the host buffers are complete
    t = time.perf_counter() - t0
    barrier(world)
    assert (h_status.array == 0).all() and np.array_equal(h_back.array, h_raw.array), "e2e round trip failed"
    t = max_over_ranks(t, world, device)
    tot_bytes = sum_over_ranks(float(nbytes), world, device)
    rec = {"value": tot_bytes * K / t / GB, "unit": "GB/s", "h2d_bytes_per_step": nbytes + total + 16 * n,
           "d2h_bytes_per_step": total + 8 * n + nbytes + 4 * n, "steps": K, "ms_per_step": t / K * 1e3,
           "bytes_per_gpu": nbytes, "slices": args.e2e_slices, "streams": 3,
           "timing": "host wall clock around K blocking library calls (max over ranks)",
           "path": "b200compStagerRoundTrip: pinned host raw -> H2D -> CompressAsync -> chunks packed into the pinned host "
                   "blob by the device + sizes D2H -> blob H2D -> DecompressAsync -> D2H raw+statuses; slices pipelined "
                   "over 3 streams, all copies inside the timed region"}
    st.close()
    for a in (h_raw, h_back, h_blob, h_sizes, h_status):
        a.free()
    # what plain pinned-memory copies give with every rank copying at once (the host side is shared by the ranks)
    barrier(world)
    cc = copy_ceiling(nbytes, device)
    rec["copy_ceiling"] = {"h2d_GBps": sum_over_ranks(cc["h2d_GBps"], world, device),
                           "d2h_GBps": sum_over_ranks(cc["d2h_GBps"], world, device),
                           "duplex_GBps_each_way": sum_over_ranks(cc["duplex_GBps_each_way"], world, device),
                           "bytes_per_rank": cc["bytes"], "ranks_copying_at_once": world,
                           "note": "sum over ranks of each rank's own rate, all ranks copying concurrently"}
    return rec


def measure_codec(args, codec, world, rank, local, device, with_cpu):
    """Times K steps of `codec` on its workload; returns the record (rank 0) or None."""
    wl = workload_for(args, codec)
    raw = make_data(args, wl, rank, device)
    nbytes = raw.numel()
    casc_opts = tuple(int(x) for x in args.casc_opts.split(","))
    ctype, cwidth = CASC_TYPES[args.casc_type]
    plan = Plan(codec, raw, wl["chunk"], casc_opts, ctype)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    clocks = ClockSampler(local)
    clocks.start()
    for _ in range(max(args.warmup, 3)):
        plan.compress()
        plan.decompress()
    plan.verify()
    comp_bytes = int(plan.comp_sizes.sum().item())

    # ---- timed region: K steps, inputs resident in HBM.  Every step touches >= 256 MiB of input plus its
    # outputs, far more than the 126 MB L2, so nothing survives in L2 from one step to the next.
    barrier(world)
    torch.cuda.synchronize()
    clocks.mark_begin()
    marks = []
    for _ in range(args.steps):
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        plan.compress()
        e1.record()
        plan.decompress()
        e2.record()
        marks.append((e0, e1, e2))
    torch.cuda.synchronize()
    clocks.mark_end()
    barrier(world)
    clk = clocks.stop()
    K = args.steps
    t_comp = sum(a.elapsed_time(b) for a, b, _ in marks) / 1e3
    t_dec = sum(b.elapsed_time(c) for _, b, c in marks) / 1e3
    t_total = marks[0][0].elapsed_time(marks[-1][2]) / 1e3
    t_total_max = max_over_ranks(t_total, world, device)
    t_comp_max = max_over_ranks(t_comp, world, device)
    t_dec_max = max_over_ranks(t_dec, world, device)
    total_bytes = sum_over_ranks(float(nbytes), world, device)
    total_comp = sum_over_ranks(float(comp_bytes), world, device)
    plan.verify()
    e2e = None
    if not args.no_e2e and args.impl == "b200":
        e2e = run_e2e(codec, plan, raw, args, world, device, comp_bytes)
    if rank != 0:
        return None
    peak, peak_src = load_peaks()
    alg_comp = nbytes + comp_bytes  # compress: raw read + compressed written (per launch, this rank)
    alg_dec = comp_bytes + nbytes
    comp_ms, dec_ms = t_comp / K * 1e3, t_dec / K * 1e3
    dominant = "compress" if t_comp >= t_dec else "decompress"

    def roof(direction):
        alg, ms = (alg_comp, comp_ms) if direction == "compress" else (alg_dec, dec_ms)
        ach = alg / (ms / 1e3) / GB
        return {"bound": "hbm", "kernel": f"{codec}::{direction}", "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, **load_traffic(codec, direction, nbytes), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg, "ms_per_launch": ms}

    launches = {"lz4": 2, "snappy": 2, "cascaded": CASCADED_LAUNCHES_PER_STEP}[codec]
    rec = {
        "metric": f"{codec}_roundtrip_throughput",
        "value": total_bytes * K / t_total_max / GB,
        "unit": "GB/s",
        "ms_per_step": t_total_max / K * 1e3,
        "dtype": "u8" if codec != "cascaded" else {1: "int8", 2: "int16", 4: "int32", 8: "int64"}[cwidth],
        "config": make_config(wl, world, nbytes, plan.n),
        "compress_GBps": total_bytes * K / t_comp_max / GB,
        "decompress_GBps": total_bytes * K / t_dec_max / GB,
        "compression_ratio": total_bytes / total_comp,
        "roofline": roof(dominant),
        "roofline_compress": roof("compress"),
        "roofline_decompress": roof("decompress"),
        "gpu_launches": launches * K,
        "clocks": clk,
    }
    if e2e:
        rec["e2e"] = e2e
    if with_cpu:
        cpu = CpuCodec(codec, wl["chunk"], casc=(ctype, *casc_opts, cwidth), threads=args.cpu_threads)
        mib = min(args.cpu_sample_mib or wl["size_mib"], wl["size_mib"])
        rec["cpu_baseline"] = cpu.baseline(raw[: mib << 20].cpu().numpy())
    del plan, raw
    torch.cuda.empty_cache()
    return rec


CASCADED_LAUNCHES_PER_STEP = 6  # 1 compress + 5 decompress instantiations (4 exit after reading partition 0's header)


def cascaded_default_options(args, world, rank, local, device):
    """The C3 column once more (1 GiB) with the library's default Cascaded options (num_RLEs 2, num_deltas 1,
    bit-packed) - the same call `--codec cascaded --casc-opts 2,1,1 --size-mib 1024 --no-e2e --no-cpu` makes, so that
    the default run's record carries it.  Device-timed only, N = 1 only; never allowed to break the headline line."""
    if world != 1 or args.impl != "b200" or args.casc_opts != "1,1,1" or args.casc_type != "int" or args.data != "default":
        return None
    try:
        sub = argparse.Namespace(**vars(args))
        sub.casc_opts, sub.size_mib, sub.no_e2e = "2,1,1", 1024, True
        rec = measure_codec(sub, "cascaded", world, rank, local, device, False)
        keep = ("config", "ms_per_step", "compress_GBps", "decompress_GBps", "compression_ratio", "roofline_compress",
                "roofline_decompress", "gpu_launches")
        return {"value": rec["value"], **{k: rec[k] for k in keep}}
    except Exception as e:  # noqa: BLE001 - an auxiliary record
        return {"error": f"{type(e).__name__}: {e}"[:300]}


def run_b200(args):
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    world, rank, local = dist_setup(args.gpus)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    codecs = [args.codec] if args.codec else ["lz4", "snappy", "cascaded"]
    with_cpu = world == 1 and not args.no_cpu and args.impl == "b200"
    recs = {}
    for c in codecs:
        recs[c] = measure_codec(args, c, world, rank, local, device, with_cpu)
    if rank != 0:
        return
    head = recs[codecs[0]]
    line = {
        **({"impl": "refgpu"} if args.impl == "refgpu" else {}),
        "metric": head["metric"], "value": head["value"], "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": head["dtype"], "data": "synthetic",
        **{k: head[k] for k in ("config", "compress_GBps", "decompress_GBps", "compression_ratio", "roofline",
                                "roofline_compress", "roofline_decompress", "gpu_launches", "clocks")},
    }
    for k in ("e2e", "cpu_baseline"):
        if k in head:
            line[k] = head[k]
    if len(codecs) > 1:
        line["codecs"] = {c: recs[c] for c in codecs[1:]}
        extra = cascaded_default_options(args, world, rank, local, device)
        if extra is not None and line["codecs"].get("cascaded") is not None:
            line["codecs"]["cascaded"]["default_options"] = extra
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- C4
def run_c4(args):
    """BASELINE.json configs[3]: LZ4 decompress-only of a pre-compressed batch of 4 KiB chunks (4 Mi chunks = 16 GiB
    by default), STRONG scaling: the batch is fixed, one global compressed-size array goes through
    sharding.shard_bounds -> every rank decompresses its contiguous, byte-balanced chunk range -> the per-chunk
    results travel through sharding.gather_sizes (NCCL all_gather; timed on its own).  Setup (untimed): every rank
    generates and compresses the blocks of its own range and of nothing else; the size array is all-gathered."""
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    world, rank, local = dist_setup(args.gpus)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    B = importlib.import_module("hipcomp-core_b200.batched")
    S = importlib.import_module("hipcomp-core_b200.synth")
    SH = importlib.import_module("hipcomp-core_b200.sharding")
    chunk = args.chunk or 4096
    total_mib = args.size_mib or 16384
    block_mib = 256  # generation unit: block b of the global buffer is log_lines(seed + b), on whichever rank needs it
    nblocks = total_mib // block_mib
    cpb = (block_mib << 20) // chunk
    nchunks = nblocks * cpb
    stride = (B.max_compressed_chunk_bytes("lz4", chunk, B.default_opts("lz4")) + 15) // 16 * 16

    def compress_block(b):
        raw = S.log_lines(block_mib << 20, 0xC400 + b, device)
        src = B.DeviceBatch(raw, np.arange(cpb, dtype=np.int64) * chunk, np.full(cpb, chunk, np.int64))
        comp = B.compress("lz4", src, max_chunk=chunk)
        return raw, comp

    # pass 1 (untimed): compressed sizes of an even split of the blocks, all-gathered into the global size array
    blo, bhi = nblocks * rank // world, nblocks * (rank + 1) // world
    local_sizes = []
    for b in range(blo, bhi):
        _, comp = compress_block(b)
        local_sizes.append(comp.sizes.clone())
        del comp
    local_sizes = torch.cat(local_sizes) if local_sizes else torch.zeros(0, dtype=torch.int64, device=device)
    if world > 1:
        gsizes, _, _, _ = SH.gather_sizes(local_sizes)
    else:
        gsizes = local_sizes.cpu().numpy()
    assert len(gsizes) == nchunks
    bounds = SH.shard_bounds(gsizes, world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    n = hi - lo
    # pass 2 (untimed): this rank's byte-balanced range, compressed streams packed back to back
    my_sizes = torch.from_numpy(np.ascontiguousarray(gsizes[lo:hi])).to(device)
    offs = torch.cumsum(my_sizes, 0) - my_sizes
    blob = torch.empty(int(my_sizes.sum().item()) + 64, dtype=torch.uint8, device=device)
    check = torch.zeros(1, dtype=torch.int64, device=device)
    lib = importlib.import_module("hipcomp-core_b200._lib").lib()
    for b in range(lo // cpb, (hi + cpb - 1) // cpb if n else 0):
        raw, comp = compress_block(b)
        c0, c1 = max(lo, b * cpb), min(hi, (b + 1) * cpb)
        k0 = c0 - b * cpb
        rc = lib.b200compGatherChunksAsync(comp.ptrs.data_ptr() + 8 * k0, comp.sizes.data_ptr() + 8 * k0,
                                           offs.data_ptr() + 8 * (c0 - lo), blob.data_ptr(), c1 - c0,
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0
        check += raw[k0 * chunk:(c1 - b * cpb) * chunk].view(torch.int64).sum()
        torch.cuda.synchronize()
        del raw, comp
    comp_batch = B.DeviceBatch(blob, offs.cpu().numpy(), my_sizes)
    out = B.DeviceBatch.empty(n, chunk, device)
    dtemp = torch.empty(max(B.decompress_temp_bytes("lz4", n, chunk), 16), dtype=torch.uint8, device=device)
    actual = torch.zeros(n, dtype=torch.int64, device=device)
    status = torch.zeros(n, dtype=torch.int32, device=device)
    f_dec = lib.hipcompBatchedLZ4DecompressAsync

    def step():
        st = f_dec(comp_batch.ptrs.data_ptr(), comp_batch.sizes.data_ptr(), out.sizes.data_ptr(), actual.data_ptr(), n,
                   dtemp.data_ptr(), dtemp.numel(), out.ptrs.data_ptr(), status.data_ptr(),
                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert st == 0

    clocks = ClockSampler(local)
    clocks.start()
    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    assert int(status.abs().sum().item()) == 0 and bool((actual == chunk).all().item())
    assert int(out.buf[: n * chunk].view(torch.int64).sum().item()) == int(check.item()), "C4: decoded bytes differ"
    barrier(world)
    torch.cuda.synchronize()
    clocks.mark_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    clocks.mark_end()
    barrier(world)
    clk = clocks.stop()
    t = max_over_ranks(e0.elapsed_time(e1) / 1e3, world, device)
    # the size query a decompress-only consumer runs first (hipcompBatchedLZ4GetDecompressSizeAsync), on its own
    qsizes = torch.zeros(n, dtype=torch.int64, device=device)
    f_size = lib.hipcompBatchedLZ4GetDecompressSizeAsync
    q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(4):
        if it == 1:
            q0.record()
        assert f_size(comp_batch.ptrs.data_ptr(), comp_batch.sizes.data_ptr(), qsizes.data_ptr(), n,
                      C.c_void_p(torch.cuda.current_stream().cuda_stream)) == 0
    q1.record()
    torch.cuda.synchronize()
    assert bool((qsizes == chunk).all().item())
    tq = max_over_ranks(q0.elapsed_time(q1) / 3e3, world, device)
    # the size / status gather of the north star, timed on its own (host wall clock, includes the NCCL all_gathers)
    tg = None
    if world > 1:
        SH.gather_sizes(actual[: min(n, 1024)])  # the first collective pays NCCL's lazy initialisation
        barrier(world)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        g_actual, g_off, first, base = SH.gather_sizes(actual)
        torch.cuda.synchronize()
        tg = max_over_ranks(time.perf_counter() - t0, world, device)
        assert len(g_actual) == nchunks and first == lo and int(g_actual.sum()) == nchunks * chunk
    my_bytes = float(int(my_sizes.sum().item()))
    max_share = max_over_ranks(my_bytes, world, device) / max(sum_over_ranks(my_bytes, world, device), 1.0)
    if rank != 0:
        return
    K = args.steps
    peak, peak_src = load_peaks()
    total_raw, total_comp = nchunks * chunk, float(np.sum(gsizes))
    alg = (total_raw + total_comp) / world  # per rank and launch (byte-balanced)
    line = {
        "metric": "lz4_decompress_throughput", "value": total_raw * K / t / GB, "unit": "GB/s", "n_gpus": world, "steps": K,
        "warmup": max(args.warmup, 3), "ms_per_step": t / K * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": f"C4: LZ4 decompress-only, {nchunks} x {chunk} B chunks ({total_mib} MiB) pre-compressed, "
                               f"sharded over {world} GPU(s) by sharding.shard_bounds (byte-balanced contiguous ranges)",
                   "bytes_total": total_raw, "chunk_bytes": chunk, "chunks_total": nchunks,
                   "largest_rank_share_of_compressed_bytes": max_share,
                   "l2": "each rank's step reads and writes far more than the 126 MB L2; no explicit flush", "seed": 0xC400},
        "compression_ratio": total_raw / total_comp,
        "size_gather_ms": None if tg is None else tg * 1e3,
        "size_query_ms": tq * 1e3, "size_query_GBps_of_compressed": total_comp / tq / GB,
        "roofline": {"bound": "hbm", "kernel": "lz4::decompress", "achieved": alg / (t / K) / GB, "peak": peak, "unit": "GB/s",
                     "frac": alg / (t / K) / GB / peak, **load_traffic("lz4", "decompress", total_raw / world),
                     "peak_source": peak_src, "ms_per_launch": t / K * 1e3},
        "gpu_launches": K, "clocks": clk,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- C5
def run_c5(args):
    """BASELINE.json configs[4]: mixed-entropy sweep - zeros / random / text / int columns through all three codecs.
    Each class is --c5-gib GiB for the whole job (strong scaling: every rank takes total / N, in waves of <= 4 GiB);
    one JSON line per (class, codec) with aggregate compress / decompress GB/s, ratio, roofline fractions and, at
    N = 1, the CPU arm on a 256 MiB sample of the same class."""
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    world, rank, local = dist_setup(args.gpus)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    peak, peak_src = load_peaks()
    chunk = args.chunk or 65536
    per_rank = (args.c5_gib << 30) // world
    wave = min(per_rank, 4 << 30)
    nwaves = max(1, per_rank // wave)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    for klass in ("zeros", "random", "text", "ints"):
        for codec in ("lz4", "snappy", "cascaded"):
            a = argparse.Namespace(**vars(args))
            a.data = klass
            wl = dict(WORKLOADS[codec], size_mib=wave >> 20, chunk=chunk)
            t_comp = t_dec = 0.0
            raw_b = comp_b = 0
            cpu_sample = None
            for w in range(nwaves):
                wl_w = dict(wl, seed=wl["seed"] + 77 * w)
                raw = make_data(a, wl_w, rank, device)
                plan = Plan(codec, raw, chunk, (1, 1, 1), 4)
                plan.compress()
                plan.decompress()
                plan.verify()
                barrier(world)
                torch.cuda.synchronize()
                e0, e1, e2 = ev(), ev(), ev()
                e0.record()
                for _ in range(args.steps):
                    plan.compress()
                e1.record()
                for _ in range(args.steps):
                    plan.decompress()
                e2.record()
                torch.cuda.synchronize()
                t_comp += e0.elapsed_time(e1) / 1e3 / args.steps
                t_dec += e1.elapsed_time(e2) / 1e3 / args.steps
                raw_b += raw.numel()
                comp_b += int(plan.comp_sizes.sum().item())
                if world == 1 and w == 0 and not args.no_cpu:
                    cpu_sample = raw[: 256 << 20].cpu().numpy()
                del plan, raw
                torch.cuda.empty_cache()
            tc, td = max_over_ranks(t_comp, world, device), max_over_ranks(t_dec, world, device)
            tot_raw, tot_comp = sum_over_ranks(float(raw_b), world, device), sum_over_ranks(float(comp_b), world, device)
            if rank != 0:
                continue
            line = {"metric": f"{codec}_class_sweep", "workload": "C5", "input_class": klass, "codec": codec, "n_gpus": world,
                    "bytes_total": tot_raw, "chunk_bytes": chunk, "waves_per_gpu": nwaves, "unit": "GB/s",
                    "compress_GBps": tot_raw / tc / GB, "decompress_GBps": tot_raw / td / GB,
                    "compression_ratio": tot_raw / tot_comp,
                    "roofline_compress_frac": (tot_raw + tot_comp) / world / tc / GB / peak,
                    "roofline_decompress_frac": (tot_raw + tot_comp) / world / td / GB / peak, "peak": peak,
                    "peak_source": peak_src, "scaling": "strong", "data": "synthetic"}
            if cpu_sample is not None:
                cpu = CpuCodec(codec, chunk, casc=(4, 1, 1, 1, 4), threads=args.cpu_threads)
                b = cpu.baseline(cpu_sample, with_single=False)
                line["cpu"] = {k: b[k] for k in ("compress_GBps", "decompress_GBps", "compression_ratio", "cores", "kind", "sample")}
            print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- reference arm
def run_reference(args):
    """Times the CPU implementation of the same round trips on the host cores (rank 0 only; the other ranks exit).
    Each step processes the arm's whole config: bytes_per_gpu x n_gpus, as n_gpus passes over rank 0's buffer."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    S = importlib.import_module("hipcomp-core_b200.synth")
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    codecs = [args.codec] if args.codec else ["lz4", "snappy", "cascaded"]
    casc_opts = tuple(int(x) for x in args.casc_opts.split(","))
    ctype, cwidth = CASC_TYPES[args.casc_type]
    W, K = max(args.warmup, 0), args.steps
    recs = {}
    for codec in codecs:
        wl = workload_for(args, codec)
        mib = min(args.cpu_sample_mib or wl["size_mib"], wl["size_mib"])
        gen = {"log_lines": lambda: S.log_lines(mib << 20, wl["seed"], dev),
               "zipf_text": lambda: S.zipf_text(mib << 20, wl["seed"], dev, chunk=wl["chunk"]),
               "sorted_ints": lambda: S.sorted_ints((mib << 20) // 4, wl["seed"], dev).view(torch.uint8)}[wl["gen"]]
        sample = gen().cpu().numpy()
        cpu = CpuCodec(codec, wl["chunk"], casc=(ctype, *casc_opts, cwidth), threads=args.cpu_threads)
        for _ in range(min(W, 1)):
            cpu.roundtrip_seconds(sample)
        base = cpu.baseline(sample, passes=max(args.gpus, 1), steps=K, with_single=(codec == codecs[0]))
        nbytes = wl["size_mib"] << 20
        recs[codec] = {
            "metric": f"{codec}_roundtrip_throughput", "value": base["value"], "unit": "GB/s",
            "ms_per_step": base["seconds_per_step"] * 1e3,
            "dtype": "u8" if codec != "cascaded" else {1: "int8", 2: "int16", 4: "int32", 8: "int64"}[cwidth],
            "config": make_config(wl, max(args.gpus, 1), nbytes, nbytes // wl["chunk"]),
            "same_config": mib == wl["size_mib"],
            "compress_GBps": base["compress_GBps"], "decompress_GBps": base["decompress_GBps"],
            "compression_ratio": base["compression_ratio"], "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
        del sample
    head = recs[codecs[0]]
    line = {"impl": "reference", "metric": head["metric"], "value": head["value"], "unit": "GB/s", "n_gpus": args.gpus,
            "steps": K, "warmup": W, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": head["dtype"], "data": "synthetic",
            **{k: head[k] for k in ("config", "same_config", "compress_GBps", "decompress_GBps", "compression_ratio",
                                    "cpu_baseline", "e2e")},
            "gpu_launches": 0}
    if len(codecs) > 1:
        line["codecs"] = {c: recs[c] for c in codecs[1:]}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "c4":
        run_c4(a)
    elif a.workload == "c5":
        run_c5(a)
    else:
        if a.impl == "refgpu":
            L = importlib.import_module("hipcomp-core_b200._lib")
            L.use(L.load_from(os.path.join(ROOT, "oracle", "_ref", "libhipcomp_ref.so")))
            a.no_e2e = True
        run_b200(a)
