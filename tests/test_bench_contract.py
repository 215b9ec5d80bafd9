"""bench.py host-side pieces that do not need a GPU: workload table, roofline.traffic lookup, CPU reference arm."""
import importlib.util
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        spec.loader.exec_module(m)
    finally:
        sys.argv = argv
    return m


def test_workloads_name_the_baseline_configs():
    b = _bench()
    assert b.WORKLOADS["lz4"]["name"].startswith("C2") and b.WORKLOADS["lz4"]["size_mib"] == 1024
    assert b.WORKLOADS["snappy"]["name"].startswith("C1") and b.WORKLOADS["snappy"]["size_mib"] == 256
    assert b.WORKLOADS["cascaded"]["name"].startswith("C3") and b.WORKLOADS["cascaded"]["size_mib"] == 4096
    assert all(w["chunk"] == 65536 for w in b.WORKLOADS.values())


def test_traffic_lookup_scales_the_committed_capture():
    b = _bench()
    t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    for codec in ("lz4", "snappy", "cascaded"):
        for direction in ("compress", "decompress"):
            e = t[codec][direction]
            got = b.load_traffic(codec, direction, 2 * e["uncompressed_bytes"])
            assert abs(got["traffic"] - 2 * e["dram_bytes"] / 1e9) < 1e-6 and got["traffic_source"] == e["source"]
            assert os.path.exists(os.path.join(ROOT, e["source"].split(" ")[0]))
    assert b.load_traffic("nope", "compress", 1)["traffic"] is None


def test_reference_arm_runs_without_a_gpu():
    """`bench.py --impl reference` is the CPU arm: it must print one JSON line with impl/cpu_baseline/e2e and no GPU."""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--cpu-sample-mib", "16"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "GB/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0


def test_default_options_subrecord_is_auxiliary():
    """codecs.cascaded.default_options: the (2,1,1) invocation's arguments, a subset of its record, N = 1 only, and an
    exception inside it never reaches the headline line."""
    b = _bench()
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        args = b.parse_args()
    finally:
        sys.argv = argv
    seen = {}

    def fake(sub, codec, world, rank, local, device, with_cpu):
        seen.update(casc_opts=sub.casc_opts, size_mib=sub.size_mib, no_e2e=sub.no_e2e, codec=codec, with_cpu=with_cpu)
        wl = b.workload_for(sub, codec)
        return {"value": 1.0, "config": b.make_config(wl, world, sub.size_mib << 20, 1), "ms_per_step": 1.0,
                "compress_GBps": 2.0, "decompress_GBps": 3.0, "compression_ratio": 7.0, "roofline_compress": {},
                "roofline_decompress": {}, "gpu_launches": 6, "clocks": {}, "e2e": {"never": "copied"}}

    real, b.measure_codec = b.measure_codec, fake
    try:
        rec = b.cascaded_default_options(args, 1, 0, 0, None)
        assert seen == dict(casc_opts="2,1,1", size_mib=1024, no_e2e=True, codec="cascaded", with_cpu=False)
        assert args.casc_opts == "1,1,1" and not args.no_e2e, "the caller's arguments are untouched"
        assert rec["compress_GBps"] == 2.0 and "opts 2,1,1" in rec["config"]["workload"] and "e2e" not in rec
        assert b.cascaded_default_options(args, 2, 0, 0, None) is None

        def boom(*a, **k):
            raise RuntimeError("no device")

        b.measure_codec = boom
        assert "RuntimeError" in b.cascaded_default_options(args, 1, 0, 0, None)["error"]
    finally:
        b.measure_codec = real
