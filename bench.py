#!/usr/bin/env python
"""bench.py — headline benchmark of the batched codec hot path (see DESIGN.md "Measurement").

Workload (BASELINE.json configs[1]): LZ4 batched compress + decompress, 64 KiB chunks of
1 GiB synthetic log text per GPU.  One step = one CompressAsync over the whole batch followed
by one DecompressAsync of its output.  metric = uncompressed GB (1e9 B) per second through that
round trip, whole job (all ranks).  Chunks are independent: ranks shard the batch, there is no
data-path collective ("scaling": "weak": every rank owns its own 1 GiB).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--codec lz4|snappy|cascaded]
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
    ap.add_argument("--codec", default="lz4", choices=["lz4", "snappy", "cascaded"])
    ap.add_argument("--size-mib", type=int, default=None, help="uncompressed MiB per GPU (default: the BASELINE config)")
    ap.add_argument("--chunk", type=int, default=None)
    ap.add_argument("--cpu-sample-mib", type=int, default=None)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--data", default="default", choices=["default", "zeros", "random", "text", "logs", "ints"],
                    help="input class (C5 sweep); default = the codec's BASELINE config generator")
    ap.add_argument("--casc-opts", default="1,1,1", help="cascaded num_RLEs,num_deltas,use_bp (C3 = 1,1,1; library default 2,1,1)")
    ap.add_argument("--e2e-slices", type=int, default=4, help="slices the e2e path pipelines over its streams")
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


# ----------------------------------------------------------------------------- CPU codecs
class CpuCodec:
    """The CPU implementation timed as the reference arm / cpu_baseline.

    lz4     : liblz4.so.1 LZ4_compress_default / LZ4_decompress_safe (the CPU LZ4 block codec the
              north_star names; the reference ships none of its own)
    snappy  : the snappy bundled with pyarrow (stand-in for share/snappy's python-snappy dependency)
    cascaded: oracle/cascaded_oracle.c (CPU restatement; no other CPU implementation exists)
    Chunks are spread over all host threads; ctypes / pyarrow release the GIL inside the codec calls.
    """

    def __init__(self, codec, chunk):
        self.codec, self.chunk = codec, chunk
        self.threads = os.cpu_count() or 1
        if codec == "lz4":
            self.lib = C.CDLL("liblz4.so.1")
            self.lib.LZ4_compress_default.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
            self.lib.LZ4_decompress_safe.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
            self.bound = self.lib.LZ4_compressBound(chunk)
            self.kind, self.desc = "reference", "liblz4.so.1 LZ4_compress_default+LZ4_decompress_safe"
        elif codec == "snappy":
            import pyarrow as pa

            self.pa = pa.Codec("snappy")
            self.kind, self.desc = "reference", "stock snappy (pyarrow bundled) compress+decompress"
        else:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import oracle_lib

            self.o = oracle_lib
            self.o.lib()
            self.kind, self.desc = "port", "oracle/cascaded_oracle.c compress+decompress (RLE=1,delta=1,bp=1)"

    def _run_threads(self, fn, n):
        T = min(self.threads, n)
        errs = []

        def work(k):
            try:
                for i in range(k, n, T):
                    fn(i)
            except Exception as e:  # pragma: no cover
                errs.append(e)

        ts = [threading.Thread(target=work, args=(k,)) for k in range(T)]
        t0 = time.perf_counter()
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        dt = time.perf_counter() - t0
        if errs:
            raise errs[0]
        return dt

    def roundtrip_seconds(self, raw):
        """raw: np.uint8 array, multiple of chunk.  Returns (t_compress, t_decompress, compressed_bytes)."""
        n, ch = len(raw) // self.chunk, self.chunk
        if self.codec == "lz4":
            comp = np.empty((n, self.bound), np.uint8)
            sizes = np.zeros(n, np.int64)
            back = np.empty_like(raw)
            rp, cp, bp = raw.ctypes.data, comp.ctypes.data, back.ctypes.data
            lib, bound = self.lib, self.bound

            def enc(i):
                sizes[i] = lib.LZ4_compress_default(rp + i * ch, cp + i * bound, ch, bound)

            def dec(i):
                if lib.LZ4_decompress_safe(cp + i * bound, bp + i * ch, int(sizes[i]), ch) != ch:
                    raise RuntimeError("liblz4 round trip failed")

            tc = self._run_threads(enc, n)
            td = self._run_threads(dec, n)
            assert back[: ch * 4].tobytes() == raw[: ch * 4].tobytes()
            return tc, td, int(sizes.sum())
        if self.codec == "snappy":
            comp = [None] * n
            mv = memoryview(raw)

            def enc(i):
                comp[i] = self.pa.compress(mv[i * ch : (i + 1) * ch])

            def dec(i):
                self.pa.decompress(comp[i], decompressed_size=ch)

            tc = self._run_threads(enc, n)
            td = self._run_threads(dec, n)
            return tc, td, sum(c.size for c in comp)
        L = self.o.lib()
        stride = ch + 16
        comp = np.zeros((n, stride), np.uint8)
        sizes = np.zeros(n, np.int64)
        back = np.empty_like(raw)
        rp, cp, bp = raw.ctypes.data, comp.ctypes.data, back.ctypes.data

        def enc(i):
            sizes[i] = L.casc_oracle_compress(rp + i * ch, ch, cp + i * stride, 4, 1, 1, 1)

        def dec(i):
            a = C.c_size_t(0)
            if L.casc_oracle_decompress(cp + i * stride, int(sizes[i]), bp + i * ch, ch, 4, C.byref(a)) != 0:
                raise RuntimeError("oracle round trip failed")

        tc = self._run_threads(enc, n)
        td = self._run_threads(dec, n)
        return tc, td, int(sizes.sum())


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
    return S.sorted_ints(nbytes // 4, seed, device).view(torch.uint8)


class Plan:
    """Pre-allocated buffers + argument lists so the timed loop is just the two C-ABI calls."""

    def __init__(self, codec, raw, chunk, casc_opts=(1, 1, 1)):
        B = importlib.import_module("hipcomp-core_b200.batched")
        L = importlib.import_module("hipcomp-core_b200._lib")
        self.B, self.L, self.codec, self.chunk = B, L, codec, chunk
        dev = raw.device
        self.n = raw.numel() // chunk
        self.raw = raw
        self.opts = L.CascadedOpts(4096, L.TYPE_INT, *casc_opts) if codec == "cascaded" else B.default_opts(codec)
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

    def compress(self, src_ptrs=None, stream=None, lo=0, n=None):
        """One CompressAsync over chunks [lo, lo+n) (default: the whole batch)."""
        n = self.n if n is None else n
        s = C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)
        per_chunk_temp = self.temp_bytes // max(self.n, 1)
        st = self.f_comp((src_ptrs if src_ptrs is not None else self.src.ptrs.data_ptr()) + 8 * lo,
                         self.src.sizes.data_ptr() + 8 * lo, self.chunk, n, self.temp.data_ptr() + per_chunk_temp * lo,
                         per_chunk_temp * n, self.comp.ptrs.data_ptr() + 8 * lo, self.comp_sizes.data_ptr() + 8 * lo,
                         self.opts, s)
        assert st == 0, f"CompressAsync -> {st}"

    def decompress(self, comp_ptrs=None, comp_sizes=None, stream=None, lo=0, n=None):
        """One DecompressAsync; comp_ptrs / comp_sizes (if given) already start at chunk lo."""
        n = self.n if n is None else n
        s = C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)
        st = self.f_dec(comp_ptrs if comp_ptrs is not None else self.comp.ptrs.data_ptr() + 8 * lo,
                        comp_sizes.data_ptr() if comp_sizes is not None else self.comp_sizes.data_ptr() + 8 * lo,
                        self.dst.sizes.data_ptr() + 8 * lo, self.actual.data_ptr() + 8 * lo, n, self.dtemp.data_ptr(),
                        self.dtemp_bytes, self.dst.ptrs.data_ptr() + 8 * lo, self.status.data_ptr() + 4 * lo, s)
        assert st == 0, f"DecompressAsync -> {st}"

    def verify(self):
        torch.cuda.synchronize()
        assert int(self.status.abs().sum().item()) == 0, "a chunk failed to decompress"
        assert bool((self.actual == self.chunk).all().item())
        assert torch.equal(self.back, self.raw), "round trip is not bit-exact"


def run_b200(args):
    wl = dict(WORKLOADS[args.codec])
    if args.size_mib:
        wl["size_mib"] = args.size_mib
    if args.chunk:
        wl["chunk"] = args.chunk
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    world, rank, local = dist_setup(args.gpus)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    raw = make_data(args, wl, rank, device)
    nbytes = raw.numel()
    casc_opts = tuple(int(x) for x in args.casc_opts.split(","))
    plan = Plan(args.codec, raw, wl["chunk"], casc_opts)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    clocks = ClockSampler(local)
    clocks.start()
    for _ in range(max(args.warmup, 3)):
        plan.compress()
        plan.decompress()
    plan.verify()
    comp_bytes = int(plan.comp_sizes.sum().item())

    # ---- timed region: K steps, inputs resident in HBM.  1 GiB input + ~1.4 GiB of other buffers per
    # step is far larger than the 126 MB L2, so nothing survives in L2 from one step to the next.
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
    t_comp = sum(a.elapsed_time(b) for a, b, _ in marks) / 1e3
    t_dec = sum(b.elapsed_time(c) for _, b, c in marks) / 1e3
    t_total = marks[0][0].elapsed_time(marks[-1][2]) / 1e3
    t_total_max = max_over_ranks(t_total, world, device)
    t_comp_max = max_over_ranks(t_comp, world, device)
    t_dec_max = max_over_ranks(t_dec, world, device)
    total_bytes = sum_over_ranks(float(nbytes), world, device)
    total_comp = sum_over_ranks(float(comp_bytes), world, device)
    plan.verify()

    # ---- e2e: host buffers in, host buffers out, through the same C-ABI calls
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(plan, raw, args, world, device)

    if rank != 0:
        return
    peak, peak_src = load_peaks()
    K = args.steps
    alg_comp = nbytes + comp_bytes  # compress: raw read + compressed written (per launch, this rank)
    alg_dec = comp_bytes + nbytes
    comp_ms, dec_ms = t_comp / K * 1e3, t_dec / K * 1e3
    dominant = "compress" if t_comp >= t_dec else "decompress"
    ach = (alg_comp / (t_comp / K) if dominant == "compress" else alg_dec / (t_dec / K)) / GB
    line = {
        **({"impl": "refgpu"} if args.impl == "refgpu" else {}),
        "metric": f"{args.codec}_roundtrip_throughput",
        "value": total_bytes * K / t_total_max / GB,
        "unit": "GB/s",
        "n_gpus": world,
        "steps": K,
        "warmup": max(args.warmup, 3),
        "ms_per_step": t_total_max / K * 1e3,
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "u8" if args.codec != "cascaded" else "int32",
        "data": "synthetic",
        "config": {"workload": wl["name"] + ("" if args.data == "default" else f" [input class: {args.data}]")
                   + ("" if args.codec != "cascaded" or casc_opts == (1, 1, 1) else f" [opts {args.casc_opts}]"),
                   "bytes_per_gpu": nbytes, "chunk_bytes": wl["chunk"],
                   "chunks_per_gpu": plan.n, "sharding": f"{world} rank(s), independent chunk ranges, no collective",
                   "l2": "inputs (>= 1 GiB per step) larger than the 126 MB L2; no explicit flush",
                   "seed": wl["seed"]},
        "compress_GBps": total_bytes * K / t_comp_max / GB,
        "decompress_GBps": total_bytes * K / t_dec_max / GB,
        "compression_ratio": total_bytes / total_comp,
        "roofline": {"bound": "hbm", "kernel": f"{args.codec}::{dominant}_kernel", "achieved": ach, "peak": peak,
                     "unit": "GB/s", "frac": ach / peak, **load_traffic(args.codec, dominant, nbytes), "peak_source": peak_src,
                     "ms_per_launch": comp_ms if dominant == "compress" else dec_ms},
        "roofline_decompress": {"bound": "hbm", "achieved": alg_dec / (t_dec / K) / GB, "peak": peak, "unit": "GB/s",
                                "frac": alg_dec / (t_dec / K) / GB / peak, "ms_per_launch": dec_ms},
        "roofline_compress": {"bound": "hbm", "achieved": alg_comp / (t_comp / K) / GB, "peak": peak, "unit": "GB/s",
                              "frac": alg_comp / (t_comp / K) / GB / peak, "ms_per_launch": comp_ms},
        "gpu_launches": 2 * K if args.codec != "cascaded" else 6 * K,  # cascaded: 1 compress + 5 decompress instantiations
        "clocks": clk,
    }
    if e2e:
        line["e2e"] = e2e
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(args, wl, raw)
    print(json.dumps(line), flush=True)


def run_e2e(plan, raw, args, world, device):
    """Host-resident input -> compress -> host-resident compressed blob -> decompress -> host-resident output,
    through the same C-ABI calls, as a caller with host data would drive them: the batch is cut into slices
    that travel down three streams, so the H2D copy of one slice, the kernels of another and the D2H copy of a
    third overlap (PCIe is full duplex).  Every byte still makes the whole trip inside the timed region:
      pinned raw -H2D-> CompressAsync -> gather -D2H-> pinned blob + sizes -H2D-> DecompressAsync -D2H-> pinned raw
    The host needs each slice's compressed size before it can size the blob copies: it waits on that slice's
    event only, while later slices are already queued."""
    n, chunk, nbytes = plan.n, plan.chunk, raw.numel()
    S = max(1, min(args.e2e_slices, n))
    bounds = [n * k // S for k in range(S + 1)]
    h_raw = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    h_raw.copy_(raw)
    h_back = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    h_sizes = torch.empty(n, dtype=torch.int64, pin_memory=True)
    h_status = torch.empty(n, dtype=torch.int32, pin_memory=True)
    h_blob = torch.empty(n * plan.stride, dtype=torch.uint8, pin_memory=True)
    d_in = torch.empty_like(raw)
    d_blob = torch.empty(n * plan.stride, dtype=torch.uint8, device=device)
    d_blob2 = torch.empty_like(d_blob)
    d_sizes2 = torch.empty(n, dtype=torch.int64, device=device)
    in_ptrs = torch.from_numpy(np.arange(n, dtype=np.int64) * chunk + d_in.data_ptr()).to(device)
    streams = [torch.cuda.Stream(device) for _ in range(min(3, S))]
    K = max(2, min(args.steps, 5))
    main = torch.cuda.current_stream()

    def stage1(k):
        lo, hi = bounds[k], bounds[k + 1]
        st = streams[k % len(streams)]
        with torch.cuda.stream(st):
            d_in[lo * chunk:hi * chunk].copy_(h_raw[lo * chunk:hi * chunk], non_blocking=True)
            plan.compress(src_ptrs=in_ptrs.data_ptr(), stream=st, lo=lo, n=hi - lo)
            sz = plan.comp_sizes[lo:hi]
            offs = torch.cumsum(sz, 0) - sz
            rc = plan.f_gather(plan.comp.ptrs.data_ptr() + 8 * lo, sz.data_ptr(), offs.data_ptr(),
                               d_blob.data_ptr() + lo * plan.stride, hi - lo, C.c_void_p(st.cuda_stream))
            assert rc == 0
            h_sizes[lo:hi].copy_(sz, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(st)
        return ev, offs

    def stage2(k, ev):
        lo, hi = bounds[k], bounds[k + 1]
        st = streams[k % len(streams)]
        ev.synchronize()
        total = int(h_sizes[lo:hi].sum().item())
        base = lo * plan.stride
        with torch.cuda.stream(st):
            h_blob[base:base + total].copy_(d_blob[base:base + total], non_blocking=True)
            # decompress direction, from the host copy of the compressed slice
            d_blob2[base:base + total].copy_(h_blob[base:base + total], non_blocking=True)
            d_sizes2[lo:hi].copy_(h_sizes[lo:hi], non_blocking=True)
            sz = d_sizes2[lo:hi]
            ptrs = torch.cumsum(sz, 0) - sz + (d_blob2.data_ptr() + base)
            plan.decompress(comp_ptrs=ptrs.data_ptr(), comp_sizes=sz, stream=st, lo=lo, n=hi - lo)
            h_back[lo * chunk:hi * chunk].copy_(plan.back[lo * chunk:hi * chunk], non_blocking=True)
            h_status[lo:hi].copy_(plan.status[lo:hi], non_blocking=True)
        return total, ptrs

    def step():
        for st in streams:
            st.wait_stream(main)
        depth = len(streams) - 1 if len(streams) > 1 else 0
        pending, keep, total = [], [], 0
        for k in range(S + depth):
            if k < S:
                pending.append(stage1(k))
            j = k - depth
            if j >= 0:
                ev, offs = pending[j]
                t, ptrs = stage2(j, ev)
                total += t
                keep.append((offs, ptrs))
        for st in streams:
            main.wait_stream(st)
        main.synchronize()
        return total

    plan.back.zero_()
    total = step()
    assert int(h_status.abs().sum().item()) == 0 and torch.equal(h_back, h_raw), "e2e round trip failed"
    barrier(world)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        total = step()
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    assert int(h_status.abs().sum().item()) == 0 and torch.equal(h_back, h_raw), "e2e round trip failed"
    t = max_over_ranks(e0.elapsed_time(e1) / 1e3, world, device)
    tot_bytes = sum_over_ranks(float(nbytes), world, device)
    return {"value": tot_bytes * K / t / GB, "unit": "GB/s", "h2d_bytes_per_step": nbytes + total + 8 * n,
            "d2h_bytes_per_step": total + 8 * n + nbytes + 4 * n, "steps": K, "ms_per_step": t / K * 1e3,
            "slices": S, "streams": len(streams),
            "path": "pinned host raw -> H2D -> CompressAsync -> gather -> D2H blob+sizes -> H2D -> DecompressAsync -> "
                    "D2H raw+statuses; slices pipelined over streams, all copies inside the timed region"}


def cpu_baseline(args, wl, raw):
    cpu = CpuCodec(args.codec, wl["chunk"])
    mib = args.cpu_sample_mib or min(wl["size_mib"], max(64, 32 * cpu.threads))
    sample = raw[: mib << 20].cpu().numpy()
    tc, td, cb = cpu.roundtrip_seconds(sample)
    return {"value": len(sample) / (tc + td) / GB, "unit": "GB/s", "cores": cpu.threads, "kind": cpu.kind,
            "sample": f"first {mib} MiB of the same workload, {cpu.desc}, {cpu.threads} threads",
            "compress_GBps": len(sample) / tc / GB, "decompress_GBps": len(sample) / td / GB,
            "compression_ratio": len(sample) / cb}


# ----------------------------------------------------------------------------- reference arm
def run_reference(args):
    """Times the CPU implementation of the same round trip on the host cores (rank 0 only)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    wl = dict(WORKLOADS[args.codec])
    if args.size_mib:
        wl["size_mib"] = args.size_mib
    if args.chunk:
        wl["chunk"] = args.chunk
    cpu = CpuCodec(args.codec, wl["chunk"])
    mib = args.cpu_sample_mib or min(wl["size_mib"], max(64, 32 * cpu.threads))
    S = importlib.import_module("hipcomp-core_b200.synth")
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    gen = {"log_lines": lambda: S.log_lines(mib << 20, wl["seed"], dev),
           "zipf_text": lambda: S.zipf_text(mib << 20, wl["seed"], dev, chunk=wl["chunk"]),
           "sorted_ints": lambda: S.sorted_ints((mib << 20) // 4, wl["seed"], dev).view(torch.uint8)}[wl["gen"]]
    sample = gen().cpu().numpy()
    W, K = max(args.warmup, 1), args.steps
    for _ in range(min(W, 2)):
        cpu.roundtrip_seconds(sample)
    tcs, tds, cb = [], [], 0
    for _ in range(K):
        tc, td, cb = cpu.roundtrip_seconds(sample)
        tcs.append(tc)
        tds.append(td)
    t = sum(tcs) + sum(tds)
    v = len(sample) * K / t / GB
    line = {
        "impl": "reference", "metric": f"{args.codec}_roundtrip_throughput", "value": v, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": t / K * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8" if args.codec != "cascaded" else "int32",
        "data": "synthetic",
        "config": {"workload": wl["name"], "chunk_bytes": wl["chunk"], "seed": wl["seed"],
                   "step": f"bounded sample: first {mib} MiB of the workload per step"},
        "compress_GBps": len(sample) * K / sum(tcs) / GB, "decompress_GBps": len(sample) * K / sum(tds) / GB,
        "compression_ratio": len(sample) / cb,
        "cpu_baseline": {"value": v, "unit": "GB/s", "cores": cpu.threads, "kind": cpu.kind,
                         "sample": f"first {mib} MiB of the same workload, {cpu.desc}, {cpu.threads} threads"},
        "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        if a.impl == "refgpu":
            L = importlib.import_module("hipcomp-core_b200._lib")
            L.use(L.load_from(os.path.join(ROOT, "oracle", "_ref", "libhipcomp_ref.so")))
            a.no_e2e = True
        run_b200(a)
